"""Multi-GPU equivalence check, run under torchrun on >= 2 GPUs (not collected by pytest: one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29761 \
        tests/multi_gpu_check.py

For the same per-rank batches it checks that (1) the NCCL path (allreduce + sgd kernel), (2) the fused
peer-memory kernel (b200_p2p_sgd_apply) and (3) the oracle's allreduce_step give the same parameters, that the
replicas stay bit-identical, and times both exchange paths inside CUDA graphs."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200cnn  # noqa: E402
from oracle import ref_numpy as rn  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B, mode = 128, os.environ.get("B200CNN_MODE", "bf16")
    total = b200cnn.TrainStepEngine(batch=8, mode="fp32").total
    use_mc = os.environ.get("B200CNN_MULTICAST", "1") == "1"
    ex = b200cnn.P2PExchange(total, dev, use_multicast=use_mc)
    eng = b200cnn.TrainStepEngine(batch=B, mode=mode, device=local, params=ex.params, grads=ex.grads)
    params0 = rn.init_params(0)
    rng = np.random.default_rng(77 + rank)
    img = np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
    lab = rng.integers(0, 10, (B,)).astype(np.int32)
    imgd, labd = torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda()
    lr = 0.1

    def grads_step():
        eng.forward(imgd, labd)
        eng.backward()

    # (1) NCCL allreduce + sgd kernel
    eng.load_params(params0)
    grads_step()
    g_local = eng.grads.clone()
    ex.allreduce_(eng.grads)
    eng.sgd_apply(lr, ex.grad_scale)
    torch.cuda.synchronize()
    p_nccl = eng.params.clone()
    # (2) fused peer-memory kernel on the SAME local gradients (the bias-gradient atomics of the bf16 backward
    #     make two backward passes differ in the last bit, so the gradients are reused, not recomputed)
    eng.load_params(params0)
    eng.grads.copy_(g_local)
    torch.cuda.synchronize(); dist.barrier()
    ex.apply_(eng, lr)
    torch.cuda.synchronize()
    dist.barrier()
    p_p2p = eng.params.clone()
    diff = float((p_p2p - p_nccl).abs().max())
    scale = float(p_nccl.abs().max())
    same_replicas = ex.replicas_identical(eng.params)
    # (3) oracle semantics: SUM of the workers' gradients applied once with lr
    gl = [torch.zeros_like(g_local) for _ in range(world)]
    dist.all_gather(gl, g_local)
    want = torch.as_tensor(np.concatenate([np.pad(params0[k].reshape(-1), (0, (-params0[k].size) % 64))
                                           for k in rn.PARAM_NAMES])).cuda() - lr * sum(gl)
    err_oracle = float((p_p2p - want).abs().max()) / scale
    # a second step on top (epoch counter, flags reuse)
    grads_step()
    ex.apply_(eng, lr)
    torch.cuda.synchronize()
    same2 = ex.replicas_identical(eng.params)

    # (4) the same step with the exchange overlapped with the backward pass (b200_train_step_p2p): from params0 on the
    #     same batch it must land on the parameters of (2) -- up to the last-bit jitter of the backward's atomics -- and
    #     leave all replicas bit-identical, twice in a row (epoch counters of both flag channels)
    eng.load_params(params0)
    torch.cuda.synchronize(); dist.barrier()
    ex.overlap = True       # exercise the overlapped variant whatever the world size
    ex.train_step_(eng, imgd, labd, lr)
    torch.cuda.synchronize(); dist.barrier()
    diff_ovl = float((eng.params - p_p2p).abs().max())
    same3 = ex.replicas_identical(eng.params)
    ex.train_step_(eng, imgd, labd, lr)
    torch.cuda.synchronize(); dist.barrier()
    same4 = ex.replicas_identical(eng.params)

    # (5) the same step with the exchange as one kernel at the end and conv1's partial sums folded into it
    #     (b200_train_step_p2p_single): bit-identical to (2), replicas identical
    eng.load_params(params0)
    torch.cuda.synchronize(); dist.barrier()
    ex.overlap = False
    ex.train_step_(eng, imgd, labd, lr)
    torch.cuda.synchronize(); dist.barrier()
    diff_single = float((eng.params - p_p2p).abs().max())
    same6 = ex.replicas_identical(eng.params)
    ex.overlap = True

    # timing inside graphs
    def timed(fn, n=200):
        g = eng.capture(fn)
        for _ in range(10):
            g.replay()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            g.replay()
        e1.record(); torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1) / n * 1000], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    t_nccl = timed(lambda: (ex.allreduce_(eng.grads), eng.sgd_apply(lr, ex.grad_scale)))
    t_p2p = timed(lambda: ex.apply_(eng, lr))
    eng.load_params(params0)

    def full_p2p():
        eng.forward(imgd, labd); eng.backward(); ex.apply_(eng, lr)

    def full_nccl():
        eng.forward(imgd, labd); eng.backward(); ex.allreduce_(eng.grads); eng.sgd_apply(lr, ex.grad_scale)
    t_step_p2p = timed(full_p2p, 100)
    eng.load_params(params0)
    t_step_nccl = timed(full_nccl, 100)
    eng.load_params(params0)
    t_step_ovl = timed(lambda: ex.train_step_(eng, imgd, labd, lr), 100)
    same5 = ex.replicas_identical(eng.params)
    eng.load_params(params0)
    ex.overlap = False
    t_step_single = timed(lambda: ex.train_step_(eng, imgd, labd, lr), 100)
    same5 = same5 and ex.replicas_identical(eng.params)
    ex.overlap = True
    if rank == 0:
        print(f"world={world} multicast={ex.multicast} |p2p - nccl|max = {diff:.3e} (scale {scale:.3f}), "
              f"vs oracle sum-update rel {err_oracle:.2e}, replicas identical: {same_replicas} / {same2}")
        print(f"exchange+update only: nccl allreduce + sgd kernel {t_nccl:.1f} us, fused p2p kernel {t_p2p:.1f} us")
        print(f"whole step (batch {B}/GPU): nccl path {t_step_nccl:.1f} us, fused p2p path {t_step_p2p:.1f} us, "
              f"overlapped p2p path {t_step_ovl:.1f} us -> {B * world / t_step_ovl * 1e6:.0f} img/s, "
              f"single-kernel path with folded reduction {t_step_single:.1f} us")
        print(f"overlapped exchange: |overlapped - fused|max = {diff_ovl:.3e}, single-kernel step: {diff_single:.3e}, "
              f"replicas identical: {same3} / {same4} / {same5} / {same6}")
    ok = diff <= 1e-6 * scale and err_oracle <= 1e-6 and same_replicas and same2 and \
        diff_ovl <= 1e-5 * scale and same3 and same4 and same5 and diff_single <= 1e-5 * scale and same6
    torch.cuda.synchronize(); dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
