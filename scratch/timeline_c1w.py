import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B, crop = 128, 24
OH = crop // 2
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
w = (torch.randn(5, 5, 3, 64, device='cuda') * 0.05); b = torch.zeros(64, device='cuda')
pool = torch.empty(B, OH, OH, 64, device='cuda'); arg = torch.empty(B, OH, OH, 64, dtype=torch.uint8, device='cuda')
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.b200_conv1_pool_fwd_bf16(img.data_ptr(), w.data_ptr(), b.data_ptr(), pool.data_ptr(), arg.data_ptr(), None, B, 32, 32, crop, crop, st))
dpool = torch.randn(B, OH, OH, 64, device='cuda')
nb = C.c_int64(); lib.b200_conv1_wgrad_bf16_scratch_bytes(B, C.byref(nb)); scr = torch.empty(max(nb.value, 256), dtype=torch.uint8, device='cuda')
dw = torch.empty(5, 5, 3, 64, device='cuda'); db = torch.empty(64, device='cuda')
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for cold in (0, 1):
    for _ in range(3):
        if cold: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.b200_conv1_wgrad_bf16(img.data_ptr(), dpool.data_ptr(), pool.data_ptr(), arg.data_ptr(), dw.data_ptr(), db.data_ptr(), scr.data_ptr(), scr.numel(), B, 32, 32, crop, crop, st))
        e1.record(); torch.cuda.synchronize()
    t = tl.cpu().numpy()
    print("conv1 wgrad", "cold" if cold else "hot", "event us %.1f" % (e0.elapsed_time(e1) * 1000), "cycles since start:", (t[26:32] - t[25]).tolist())
