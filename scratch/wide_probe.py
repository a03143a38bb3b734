"""Phase stamps (clock64 of CTA 0) of the wide model's conv2 forward / input-gradient kernels + per-CTA durations."""
import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
e = b200cnn.TrainStepEngine(batch=B, mode="bf16", c1=256, c2=256, f1=768, f2=384); e.init_params(0)
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
buf = torch.zeros(64 + 2 * 3 * 512, dtype=torch.int64, device='cuda')
trace = lib.b200_debug_trace(buf.data_ptr()) == 0
def reset(sel):
    buf[0:32:2] = torch.iinfo(torch.int64).max; buf[1:32:2] = 0; buf[32] = sel; buf[33] = -1; buf[64:] = 0
def ctas(name):
    raw = buf.cpu().numpy()
    c = raw[64:64 + 1536].reshape(-1, 3); c = c[c[:, 0] > 0]
    if len(c):
        t0 = c[:, 0].min(); s, en = (c[:, 0] - t0) / 1e3, (c[:, 1] - t0) / 1e3
        print(f"   {name}: first {len(c)} CTAs: durations us min {np.min(en-s):.1f} median {np.median(en-s):.1f} max {np.max(en-s):.1f}; last end {en.max():.1f} us")
for it in range(3):
    reset(3); torch.cuda.synchronize()
    e.forward(img, lab); torch.cuda.synchronize()
    t = tl.cpu().numpy()
    print("conv2 fwd   CTA0 cycles: strip/t1 %d t2 %d mma_start %d mma_issued %d accum_ready %d epilogue_done %d" % tuple(t[1:7] - t[0]))
    if trace: ctas("conv2_fwd")
    reset(9); torch.cuda.synchronize()
    e.backward(); torch.cuda.synchronize()
    t = tl.cpu().numpy()
    print("conv2 dgrad CTA0 cycles: strip/t1 %d t2 %d mma_start %d mma_issued %d accum_ready %d epilogue_done %d" % tuple(t[1:7] - t[0]))
    if trace: ctas("conv2_dgrad")
