"""Kernel schedule of one captured multi-GPU bf16 train step (B200_TRACE=1 build), rank 0's view; run under torchrun."""
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, '.')
import b200cnn
from b200cnn import _lib
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = _lib.load()
B = 128
names = ["pack", "w2img", "conv1_fwd", "conv2_fwd", "fc1_fwd", "tail_rows", "tail_grads", "fc1_dw", "fc1_dx",
         "conv2_dgrad", "conv2_wgrad", "wgrad_reduce", "conv1_wgrad", "sgd", "p2p_fc", "p2p_conv"]
total = b200cnn.TrainStepEngine(batch=8, mode="fp32").total
ex = b200cnn.P2PExchange(total, torch.device("cuda", local), overlap=(os.environ.get("OVERLAP", "1") == "1"))
e = b200cnn.TrainStepEngine(batch=B, mode="bf16", device=local, params=ex.params, grads=ex.grads); e.init_params(0)
params0 = e.params.clone()
rng = np.random.default_rng(rank)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
buf = torch.zeros(64 + 2 * 3 * 512, dtype=torch.int64, device='cuda')
_lib.check(lib.b200_debug_trace(buf.data_ptr()))
def reset():
    buf[0:32:2] = torch.iinfo(torch.int64).max; buf[1:32:2] = 0; buf[32] = -1; buf[33] = -1; buf[64:] = 0
def step():
    e.params.copy_(params0)
    ex.train_step_(e, img, lab, 0.1)
g = e.capture(step)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for cold in (0, 1):
    for _ in range(6):
        if cold: flush.zero_()
        reset(); torch.cuda.synchronize(); dist.barrier(); g.replay(); torch.cuda.synchronize()
    raw = buf.cpu().numpy()
    t = raw[:32].reshape(-1, 2).astype(np.float64)
    ok = t[:, 1] > 0
    t0 = t[ok, 0].min()
    if rank == 0:
        print("world", world, "overlap", ex.overlap, "cold" if cold else "hot", "step: %.1f us from first start to last end" % ((t[ok, 1].max() - t0) / 1e3))
        for i in np.argsort(t[:, 0]):
            if ok[i]:
                print("  %-13s start %7.1f  end %7.1f  (%5.1f us)" % (names[i], (t[i, 0] - t0) / 1e3, (t[i, 1] - t0) / 1e3, (t[i, 1] - t[i, 0]) / 1e3))
torch.cuda.synchronize(); dist.barrier()
del g
os._exit(0)
