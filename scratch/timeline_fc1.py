"""Phase stamps of CTA 0 of the FC1 forward GEMM (the last tc_gemm / conv5x5 kernel of a forward pass)."""
import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B = 128
e = b200cnn.TrainStepEngine(batch=B, mode="bf16"); e.init_params(0)
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
for _ in range(3):
    e.forward(img, lab); torch.cuda.synchronize()
t = tl.cpu().numpy()
print("FC1 fwd CTA0 cycles since prologue done: wait returned %d, first stage landed %d, MMAs issued %d, accum ready %d, epilogue done %d" % tuple(t[1:6] - t[0]), "| first TMEM load back %d, second %d" % (t[6]-t[0], t[7]-t[0]))
