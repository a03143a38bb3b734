"""Per-CTA schedule of the conv2 forward / input-gradient kernels at a large batch (needs a B200_TRACE=1 build):
duration of each of the first 512 CTAs, how many share an SM at a time, phase stamps of CTA 0."""
import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
H = W = 12; Cc = 64
st = torch.cuda.current_stream().cuda_stream
x = torch.randn(B, H, W, Cc, device='cuda').relu_()
xb = x.to(torch.bfloat16).contiguous()
w = torch.randn(5, 5, Cc, Cc, device='cuda') * 0.05
b = torch.zeros(Cc, device='cuda')
nb = C.c_int64(); lib.b200_conv5x5_packed_bytes(Cc, Cc, C.byref(nb)); wp = torch.zeros(nb.value, dtype=torch.uint8, device='cuda')
lib.b200_conv5x5_pack_weights(w.data_ptr(), wp.data_ptr(), Cc, Cc, st)
pool = torch.empty(B, 6, 6, Cc, device='cuda'); arg = torch.empty(B, 6, 6, Cc, dtype=torch.uint8, device='cuda')
dpool = torch.randn(B, 6, 6, Cc, device='cuda'); dx = torch.empty(B, H, W, Cc, device='cuda'); db = torch.zeros(Cc, device='cuda')
buf = torch.zeros(64 + 2 * 3 * 512, dtype=torch.int64, device='cuda')
trace = lib.b200_debug_trace(buf.data_ptr()) == 0
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
kernels = {"fwd": (3, lambda: lib.b200_conv5x5_pool_fwd_bf16x(xb.data_ptr(), wp.data_ptr(), b.data_ptr(), None, pool.data_ptr(), arg.data_ptr(), B, H, W, Cc, Cc, st)),
           "fwd_f32src": (3, lambda: lib.b200_conv5x5_pool_fwd_bf16(x.data_ptr(), wp.data_ptr(), b.data_ptr(), None, pool.data_ptr(), arg.data_ptr(), B, H, W, Cc, Cc, st)),
           "dgrad_pool": (9, lambda: lib.b200_conv5x5_dgrad_pool_bf16(dpool.data_ptr(), pool.data_ptr(), arg.data_ptr(), wp.data_ptr(), dx.data_ptr(), db.data_ptr(), B, H, W, Cc, Cc, st))}
for name, (tid, fn) in kernels.items():
    for variant in (0, 1):
        lib.b200_debug_conv_variant(variant)
        ms = []
        for it in range(4):
            buf[0:32:2] = torch.iinfo(torch.int64).max; buf[1:32:2] = 0; buf[32] = tid; buf[33] = -1; buf[64:] = 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record(); rc = fn(); e1.record(); torch.cuda.synchronize()
            assert rc == 0, lib.b200_last_error()
            ms.append(e0.elapsed_time(e1) * 1e3)
        t = tl.cpu().numpy()
        print(f"{name} variant {variant} B={B}: event us {np.round(ms, 1).tolist()}; CTA0 cycles since start: "
              f"t1 {t[1]-t[0]} t2 {t[2]-t[0]} mma_start {t[3]-t[0]} mma_issued {t[4]-t[0]} accum_ready {t[5]-t[0]} epilogue_done {t[6]-t[0]}")
        if trace:
            raw = buf.cpu().numpy()
            c = raw[64:64 + 1536].reshape(-1, 3); c = c[c[:, 0] > 0]
            if len(c):
                t0 = c[:, 0].min()
                s, e = (c[:, 0] - t0) / 1e3, (c[:, 1] - t0) / 1e3
                conc = 0
                for sm in set(c[:, 2].tolist()):
                    m = c[c[:, 2] == sm]
                    ev = sorted([(a, 1) for a in m[:, 0]] + [(a, -1) for a in m[:, 1]])
                    cur = 0
                    for _, d in ev:
                        cur += d; conc = max(conc, cur)
                print(f"   first {len(c)} CTAs: durations us min {np.min(e-s):.1f} median {np.median(e-s):.1f} max {np.max(e-s):.1f}; "
                      f"distinct SMs {len(set(c[:,2].tolist()))}; max CTAs resident on one SM {conc}; last end {e.max():.1f} us")
lib.b200_debug_conv_variant(-1)
