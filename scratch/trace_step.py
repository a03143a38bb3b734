"""Kernel schedule of one captured bf16 train step (needs a B200_TRACE=1 build): start / end of every kernel in us
relative to the first kernel's start, from %globaltimer stamps."""
import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
lib = _lib.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
names = ["pack", "w2img", "conv1_fwd", "conv2_fwd", "fc1_fwd", "tail_rows", "tail_grads", "fc1_dw", "fc1_dx",
         "conv2_dgrad", "conv2_wgrad", "wgrad_reduce", "conv1_wgrad", "sgd", "p2p_fc", "p2p_conv"]
e = b200cnn.TrainStepEngine(batch=B, mode="bf16"); e.init_params(0)
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
buf = torch.zeros(64 + 2 * 3 * 512, dtype=torch.int64, device='cuda')
sel2 = int(sys.argv[3]) if len(sys.argv) > 3 else -1
sel = int(sys.argv[2]) if len(sys.argv) > 2 else -1      # kernel whose CTAs log individually
_lib.check(lib.b200_debug_trace(buf.data_ptr()))
def reset():
    buf[0:32:2] = torch.iinfo(torch.int64).max; buf[1:32:2] = 0; buf[32] = sel; buf[33] = sel2; buf[64:] = 0      # stamps are < 2^63: signed view orders the same
tl = torch.zeros(32, dtype=torch.int64, device='cuda')
lib.b200_debug_conv_timeline(C.cast(tl.data_ptr(), C.POINTER(C.c_int64)))
e.train_step(img, lab, 0.1); torch.cuda.synchronize()
g = e.capture(lambda: e.train_step(img, lab, 0.1))
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for cold in (0, 1):
    for _ in range(5):
        if cold: flush.zero_()
        reset(); torch.cuda.synchronize(); g.replay(); torch.cuda.synchronize()
    raw = buf.cpu().numpy()
    t = raw[:32].reshape(-1, 2).astype(np.float64)
    t = t[t[:, 1] > 0] if False else t
    t0 = t[t[:, 1] > 0, 0].min()
    print("cold" if cold else "hot", "step: %.1f us from first start to last end" % ((t[:, 1].max() - t0) / 1e3))
    tt = tl.cpu().numpy()
    print("  [%s] conv2 dgrad CTA0: zeroed %d filled %d mma_start %d mma_issued %d accum_ready %d epilogue_done %d" % (("cold" if cold else "hot",) + tuple(tt[1:7] - tt[0])))
    print("  [%s] conv2 wgrad CTA0: zeroed %d img0 %d all_filled %d accum_ready %d epilogue_done %d" % (("cold" if cold else "hot",) + tuple(tt[9:14] - tt[8])))
    print("  [%s] conv1 wgrad CTA0: staged+xstrip %d gathered %d accum_ready %d epilogue_done %d" % (("cold" if cold else "hot",) + tuple(tt[26:30] - tt[25])))
    print("  [%s] conv1 fwd CTA0: strip+filter ready %d accum_ready %d conv tile in smem %d" % ("cold" if cold else "hot", tt[15]-tt[14], tt[30]-tt[14], tt[31]-tt[14]))
    print("  [%s] tail rows phases:" % ("cold" if cold else "hot"), (tt[17:25] - tt[16:24]).tolist())
    for i in np.argsort(t[:, 0]):
        if t[i, 1] <= 0: continue
        print("  %-13s start %7.1f  end %7.1f  (%5.1f us)" % (names[i], (t[i, 0] - t0) / 1e3, (t[i, 1] - t0) / 1e3, (t[i, 1] - t[i, 0]) / 1e3))
# phase stamps (clock64) of CTA 0 of the LAST conv5x5 kernel that ran (the dgrad of the step) and of the tail rows kernel
g.replay(); torch.cuda.synchronize()
t = tl.cpu().numpy()
print("conv2 dgrad CTA0 cycles since start: zeroed %d filled %d mma_start %d mma_issued %d accum_ready %d epilogue_done %d" % tuple(t[1:7] - t[0]))
print("conv2 wgrad CTA0 cycles since start: zeroed %d img0 %d all_filled %d accum_ready %d epilogue_done %d" % tuple(t[9:14] - t[8]))
print("conv1 wgrad CTA0 cycles since start: staged+xstrip %d gathered %d accum_ready %d epilogue_done %d" % tuple(t[26:30] - t[25]))
print("conv1 fwd CTA0 cycles since start: strip+filter ready %d accum_ready %d conv tile in smem %d" % (t[15]-t[14], t[30]-t[14], t[31]-t[14]))
print("tail rows phases:", (t[17:25] - t[16:24]).tolist())

if sel >= 0:
    c = raw[64:64 + 1536].reshape(-1, 3); c = c[c[:, 0] > 0]
    st = (c[:, 0] - c[:, 0].min()) / 1e3; en = (c[:, 1] - c[:, 0].min()) / 1e3
    print("%s: %d CTAs; start spread %.1f us; durations min %.1f median %.1f max %.1f us; last end %.1f" % (names[sel], len(c), st.max(), (en - st).min(), np.median(en - st), (en - st).max(), en.max()))
    order = np.argsort(st)
    print("  starts (us):", np.round(st[order][::8], 1).tolist())
    print("  durations (us) in start order:", np.round((en - st)[order][::8], 1).tolist())
    print("  distinct SMs:", len(set(c[:, 2].tolist())))

if sel >= 0 and sel2 >= 0:
    c2 = raw[64 + 1536:].reshape(-1, 3); c2 = c2[c2[:, 0] > 0]
    t0 = min(c[:, 0].min(), c2[:, 0].min())
    both = 0
    for a in c:
        for b in c2:
            if a[2] == b[2] and a[0] < b[1] and b[0] < a[1]: both += 1
    print("%s x %s: CTA pairs resident on the same SM at the same time: %d" % (names[sel], names[sel2], both))
    sms2 = {}
    for b in c2: sms2.setdefault(int(b[2]), []).append(((b[0]-t0)/1e3, (b[1]-t0)/1e3))
    for a in c[np.argsort(c[:, 0])][:6]:
        print("  %s CTA on SM %d: %.1f-%.1f; %s there: %s" % (names[sel], a[2], (a[0]-t0)/1e3, (a[1]-t0)/1e3, names[sel2], [(round(x,1), round(y,1)) for x, y in sms2.get(int(a[2]), [])]))
    for a in c[np.argsort(c[:, 0])][-4:]:
        print("  %s CTA on SM %d: %.1f-%.1f; %s there: %s" % (names[sel], a[2], (a[0]-t0)/1e3, (a[1]-t0)/1e3, names[sel2], [(round(x,1), round(y,1)) for x, y in sms2.get(int(a[2]), [])]))
