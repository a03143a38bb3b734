import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B,H,W,Cin,Cout = 128,12,12,64,64
x = torch.randn(B,H,W,Cin, device='cuda'); w = torch.randn(5,5,Cin,Cout, device='cuda')*0.05; b = torch.zeros(Cout, device='cuda')
y = torch.empty(B,H,W,Cout, device='cuda')
nb = C.c_int64(); lib.b200_conv5x5_packed_bytes(Cin,Cout,C.byref(nb)); wp = torch.zeros(nb.value, dtype=torch.uint8, device='cuda')
st = torch.cuda.current_stream().cuda_stream
lib.b200_conv5x5_pack_weights(w.data_ptr(), wp.data_ptr(), Cin, Cout, st)
tl = torch.zeros(8, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
flush = torch.empty(256<<20, dtype=torch.uint8, device='cuda')
for name, fn in (("fwd", lambda: lib.b200_conv5x5_fwd_bf16(x.data_ptr(), wp.data_ptr(), b.data_ptr(), y.data_ptr(), B,H,W,Cin,Cout,1,st)),
                 ("dgrad", lambda: lib.b200_conv5x5_dgrad_bf16(x.data_ptr(), wp.data_ptr(), y.data_ptr(), B,H,W,Cin,Cout,st))):
    for cold in (0, 1):
        for _ in range(3):
            if cold: flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t = tl.cpu().numpy(); d = (t[1:7] - t[0])
        print(name, "cold" if cold else "hot", "event us %.1f" % (e0.elapsed_time(e1)*1000), "cycles since start: zeroed %d filled %d mma_start %d mma_issued %d accum_ready %d epilogue_done %d" % tuple(d))
