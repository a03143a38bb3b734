import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
e = b200cnn.TrainStepEngine(batch=128, mode="bf16"); e.init_params(0)
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0,256,(128,3,32,32),dtype=np.uint8)).cuda(); lab = torch.as_tensor(rng.integers(0,10,(128,)).astype(np.int32)).cuda()
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
flush = torch.empty(256<<20, dtype=torch.uint8, device='cuda')
p0 = e.params.clone()
for cold in (0,1):
    for _ in range(3):
        e.params.copy_(p0)
        if cold: flush.zero_()
        e.forward(img, lab); torch.cuda.synchronize()
    t = tl.cpu().numpy()[16:25]
    print("cold" if cold else "hot", "tail rows phases (cycles):", (t[1:]-t[:-1]).tolist(), "total", t[8]-t[0])
