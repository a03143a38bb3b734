import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B,H,W,Cin,Cout = 128,12,12,64,64
x = torch.randn(B,H,W,Cin, device='cuda'); dy = torch.randn(B,H,W,Cout, device='cuda'); dw = torch.empty(5,5,Cin,Cout, device='cuda')
nb = C.c_int64(); lib.b200_conv5x5_wgrad_bf16_scratch_bytes(B,C.byref(nb)); scr = torch.zeros(nb.value, dtype=torch.uint8, device='cuda')
st = torch.cuda.current_stream().cuda_stream
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
flush = torch.empty(256<<20, dtype=torch.uint8, device='cuda')
fn = lambda: lib.b200_conv5x5_wgrad_bf16(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), scr.data_ptr(), nb.value, B,H,W,Cin,Cout,st)
for cold in (0, 1):
    for _ in range(3):
        if cold: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); rc = fn(); e1.record(); torch.cuda.synchronize()
    t = tl.cpu().numpy(); d = (t[9:14] - t[8])
    print("wgrad", "cold" if cold else "hot", "rc", rc, "event us %.1f" % (e0.elapsed_time(e1)*1000), "cycles since start: zeroed %d img0_filled %d all_filled %d accum_ready %d epilogue_done %d" % tuple(d))
