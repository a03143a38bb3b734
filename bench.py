#!/usr/bin/env python
"""bench.py -- train images/sec of the cifar10cnn.py CNN step on N B200s (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
  python bench.py --impl reference ...                      (CPU port of the reference path)

One "step" = one `mon_sess.run([train_op, loss])` (cifar10cnn.py:230) per worker: crop/cast,
forward, backward, gradient exchange, SGD apply on one batch of 128 synthetic CIFAR records.
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "train images/sec (CIFAR-10 CNN, bs128/GPU)"
# 2 x train MACs (24x24 crop), layer-boundary activation traffic model, parameter bytes
MODELS = {
    "base": dict(dims=dict(c1=64, c2=64, f1=384, f2=192), flop=105_295_104, act=352_542, param_bytes=4_273_192,
                 name="conv5x5x64-pool-conv5x5x64-pool-fc384-fc192-fc10"),
    # BASELINE.json configs[4]: conv channels x4, FC x2
    "wide": dict(dims=dict(c1=256, c2=256, f1=768, f2=384), flop=1_504_074_240, act=1_392_798, param_bytes=36_143_656,
                 name="width-scaled conv5x5x256-pool-conv5x5x256-pool-fc768-fc384-fc10"),
}
FLOP_PER_IMAGE = MODELS["base"]["flop"]
ACT_ELEMS_PER_IMAGE = MODELS["base"]["act"]
PARAM_BYTES = MODELS["base"]["param_bytes"]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops_burst": d["bf16_tflops"],
                "tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def synthetic_batch(rng, batch):
    """CIFAR-record-shaped bytes (cifar10cnn.py:57-62): uint8 [B,3,32,32] uniform 0..255, int32 labels 0..9."""
    import numpy as np
    return (rng.integers(0, 256, (batch, 3, 32, 32), dtype=np.uint8),
            rng.integers(0, 10, (batch,)).astype(np.int32))


def cpu_baseline(batch, budget_s, threads=None, dims=None):
    """The reference's CPU path restated in torch-CPU fp32 (oracle/ref_torch.py, kind 'port';
    TensorFlow 1.x itself is not installable here) timed on this host's cores."""
    import numpy as np
    import torch
    from oracle import ref_numpy as rn, ref_torch as rt
    threads = threads or os.cpu_count()
    tr = rt.CpuTrainer(rn.init_params(0, **(dims or {})), lr=0.1, threads=threads)
    rng = np.random.default_rng(1234)
    pool = [synthetic_batch(rng, batch) for _ in range(4)]
    tr.step(*pool[0])
    tr.reset()
    n, t0 = 0, time.perf_counter()
    while True:
        tr.step(*pool[n % 4])
        tr.reset()                                             # raw-byte inputs diverge: fresh weights
        n += 1
        el = time.perf_counter() - t0
        if el >= budget_s or n >= 2000:
            break
    return {"value": batch * n / el, "unit": "images/s", "cores": threads, "kind": "port",
            "sample": f"{n} steps of batch {batch} in {el:.1f} s (torch-CPU fp32 restatement of cifar10cnn.py:94-176, "
                      f"{threads} threads, {os.cpu_count()} logical cores)"}


def _cpu_ps_worker(idx, shared, p0, batch, steps, warmup, threads, barrier, out_q, seed):
    """One CPU worker of the config-1-shaped baseline (README:11-13): pull the variables from the
    shared-memory "ps" copy, forward + backward on its own batch, push/apply without a lock
    (cifar10cnn.py:162-163 executed on the ps, :195-196)."""
    import numpy as np
    import torch
    torch.set_num_threads(threads)
    from oracle import ref_torch as rt
    from oracle.ref_numpy import PARAM_NAMES
    rng = np.random.default_rng(seed)
    pool = [synthetic_batch(rng, batch) for _ in range(4)]

    def step(i):
        p = {k: shared[k].clone().requires_grad_(True) for k in PARAM_NAMES}          # pull
        x = rt.crop_cast(pool[i % 4][0], dtype=torch.float32)
        loss = rt.cifar_loss(rt.create_cnn(p, x), pool[i % 4][1])
        grads = torch.autograd.grad(loss, [p[k] for k in PARAM_NAMES])
        with torch.no_grad():
            for k, g in zip(PARAM_NAMES, grads):
                shared[k].sub_(0.1 * g)                                                # push + apply, no lock
            if idx == 0:        # raw-byte inputs diverge within a few updates: keep the arithmetic ordinary
                for k in PARAM_NAMES:
                    shared[k].copy_(p0[k])
        return float(loss.detach())

    for i in range(warmup):
        step(i)
    barrier.wait()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    out_q.put((idx, t0, time.perf_counter()))


def cpu_ps_workers(batch, n_workers, steps, warmup=2):
    """1 coordinator ("ps": this process, owner of the shared-memory variables) + n_workers worker processes,
    host threads split evenly, asynchronous lock-free updates.  -> (images/s, description)"""
    import torch
    import torch.multiprocessing as mp
    from oracle import ref_numpy as rn
    cores = os.cpu_count() or 1
    threads = max(1, cores // n_workers)
    ctx = mp.get_context("spawn")
    p0 = {k: torch.tensor(v) for k, v in rn.init_params(0).items()}
    shared = {k: v.clone().share_memory_() for k, v in p0.items()}
    barrier, q = ctx.Barrier(n_workers), ctx.Queue()
    procs = [ctx.Process(target=_cpu_ps_worker,
                         args=(i, shared, p0, batch, steps, warmup, threads, barrier, q, 1234 + i))
             for i in range(n_workers)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    el = max(r[2] for r in res) - min(r[1] for r in res)
    return batch * steps * n_workers / el, (
        f"{steps} steps of batch {batch} on each of {n_workers} worker processes ({threads} threads each, "
        f"{cores} logical cores) + 1 coordinator holding the variables in shared memory, asynchronous lock-free "
        f"updates (the README's 1 ps + N workers recipe; torch-CPU fp32 restatement, TensorFlow 1.x unavailable)")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    import torch
    from oracle import ref_numpy as rn, ref_torch as rt
    threads = os.cpu_count()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    warm = min(args.warmup, 5)
    if world > 1:
        # N GPUs are compared with N CPU workers (1 ps + N workers on this host), not with one process
        steps = max(8, min(args.steps, 400) // world)
        t0 = time.perf_counter()
        v, sample = cpu_ps_workers(args.batch, world, steps, warmup=min(warm, 3))
        ms = 1e3 * args.batch * world / v
    else:
        tr = rt.CpuTrainer(rn.init_params(0, **MODELS[args.model]["dims"]), lr=0.1, threads=threads)
        rng = np.random.default_rng(1234)
        pool = [synthetic_batch(rng, args.batch) for _ in range(4)]
        steps = min(args.steps, 400)
        if args.batch > 128:
            steps = max(4, steps * 128 // args.batch)
        if args.model == "wide":
            steps = max(4, steps // 12)
        for i in range(warm):
            tr.step(*pool[i % 4])
            tr.reset()
        t0 = time.perf_counter()
        for i in range(steps):
            tr.step(*pool[i % 4])
            tr.reset()
        el = time.perf_counter() - t0
        v = args.batch * steps / el
        ms = 1e3 * el / steps
        sample = (f"{steps} steps of batch {args.batch}, single process, {threads} threads; the reference's own "
                  f"TensorFlow-1.x runtime is unavailable on this box, this is the torch-CPU restatement")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "batch_per_gpu": args.batch, "host_threads": threads,
                   "cpu_workers": world},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_name(args):
    return (f"cifar10cnn.py CNN train step ({MODELS[args.model]['name']}, softmax-xent, "
            f"SGD lr 0.1), 32x32x3 uint8 records centre-cropped to 24x24, batch {args.batch}/GPU")


# algorithmic work per launch site: ("tensor", MACs per image) or ("hbm", bytes per image, fixed bytes)
def site_work(mode, dims=None, param_bytes=PARAM_BYTES):
    act = 4 if mode == "fp32" else 2
    d = dims or MODELS["base"]["dims"]
    C1, C2, F1, F2 = d["c1"], d["c2"], d["f1"], d["f2"]
    c1, p1, c2, p2 = 576 * C1, 144 * C1, 144 * C2, 36 * C2
    m_c1, m_c2, m_f1, m_f2, m_f3 = 576 * 75 * C1, 144 * 25 * C1 * C2, p2 * F1, F1 * F2, F2 * 10      # MACs per image
    return {
        "conv1_fwd": ("tensor", m_c1), "conv1_wgrad": ("tensor", m_c1),
        "conv2_fwd": ("tensor", m_c2), "conv2_dgrad": ("tensor", m_c2),
        "conv2_wgrad": ("tensor", m_c2),
        "full1_fwd": ("tensor", m_f1), "full1_bwd": ("tensor", 2 * m_f1),
        "full1_dw": ("tensor", m_f1), "full1_dx": ("tensor", m_f1),
        "conv1_pool1_fwd": ("tensor", m_c1), "conv2_pool2_fwd": ("tensor", m_c2), "fc_tail": ("tensor", 3 * (m_f2 + m_f3)),
        "weights_prep": ("hbm", 0, 25 * C1 * C2 * 6 + F1 * F2 * 6),
        "full2_fwd": ("tensor", m_f2), "full2_bwd": ("tensor", 2 * m_f2),
        "full3_fwd": ("tensor", m_f3), "full3_bwd": ("tensor", 2 * m_f3),
        "crop_cast": ("hbm", 3072 + 1728 * act, 0),
        "pool1_fwd": ("hbm", c1 * act + p1 * act + p1, 0), "pool2_fwd": ("hbm", c2 * act + p2 * act + p2, 0),
        "pool1_bwd": ("hbm", p1 * act + p1 + c1 * act * 2, 0), "pool2_bwd": ("hbm", p2 * act + p2 + c2 * act * 2, 0),
        "head": ("hbm", 10 * 4 * 2 + 4 + 4, 0), "conv2_dgrad_prep": ("hbm", 0, 2 * 25 * C1 * C2 * 4),
        "sgd_apply": ("hbm", 0, 3 * param_bytes),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--model", default="base", choices=sorted(MODELS),
                    help="base = the script's own widths; wide = BASELINE.json's width-scaled model (conv x4, FC x2)")
    ap.add_argument("--mode", default=os.environ.get("B200CNN_MODE", "auto"), choices=["auto", "fp32", "bf16"])
    ap.add_argument("--reduce", default="sum", choices=["sum", "mean"])
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N>1: fused peer-memory allreduce+SGD kernel (default) or NCCL allreduce + SGD kernel")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sites-json", default=None, help="also write the per-site timing table here")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import b200cnn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if args.gpus != world and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE {world}; reporting n_gpus={world}", file=sys.stderr)
    mode = b200cnn.DEFAULT_MODE if args.mode == "auto" else args.mode
    B, K, W = args.batch, args.steps, args.warmup
    lr = 0.1
    peaks = load_peaks()

    ex, exchange = None, "none"
    if world > 1 and args.exchange == "p2p":
        try:
            total = b200cnn.TrainStepEngine(batch=8, mode="fp32", device=local, **MODELS[args.model]["dims"]).total
            ex = b200cnn.P2PExchange(total, torch.device("cuda", local), reduce=args.reduce)
            exchange = "p2p-fused" + ("-nvls" if ex.multicast else "") + ("-overlapped" if ex.overlap else "")
        except Exception as e:          # no symmetric memory on this box: fall back to NCCL (still all CUDA)
            if rank == 0:
                print(f"P2P exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
            ex = None
    if ex is None:
        ex = b200cnn.GradientExchange(reduce=args.reduce)
        exchange = "nccl" if world > 1 else "none"
    fused = isinstance(ex, b200cnn.P2PExchange)
    model = MODELS[args.model]
    eng = b200cnn.TrainStepEngine(batch=B, mode=mode, device=local, params=ex.params if fused else None,
                                  grads=ex.grads if fused else None, **model["dims"])
    eng.init_params(0)
    params0 = eng.params.clone()
    ex.broadcast_params_(eng.params)
    if fused and ex.overlap and args.model == "base":
        ex.bind(eng)                   # inbox of the one-shot conv-gradient exchange (collective, outside any capture)
        if ex._inbox is not None:
            exchange += "-push"

    rng = np.random.default_rng(1234 + rank)
    POOL, RING = 16, 8
    pool = []
    for _ in range(POOL):
        i, l = synthetic_batch(rng, B)
        pool.append((torch.as_tensor(i).cuda(), torch.as_tensor(l).cuda()))
    host_ring = []
    for _ in range(RING):
        i, l = synthetic_batch(rng, B)
        host_ring.append((torch.as_tensor(i).pin_memory(), torch.as_tensor(l).pin_memory(),
                          torch.zeros(1).pin_memory()))
    s_img = torch.empty_like(pool[0][0])
    s_lab = torch.empty_like(pool[0][1])

    def device_step():
        if world == 1:
            eng.train_step(s_img, s_lab, lr)
        else:
            if fused and args.model == "base":
                ex.train_step_(eng, s_img, s_lab, lr)      # exchange split in two and overlapped with the backward
            else:
                eng.forward_backward(s_img, s_lab)
                exchange_and_update()

    def exchange_and_update():
        if fused:
            ex.apply_(eng, lr)                       # one kernel: allreduce over peer memory + SGD
        else:
            ex.allreduce_(eng.grads)
            eng.sgd_apply(lr, ex.grad_scale)

    # end-to-end path: the public host-buffer call with prefetch (b200_train_step_host_prefetch) -- graph j trains on the
    # batch staged in slot j % 2 while host batch j + 1 is copied (pinned -> device) into the other slot on a copy
    # stream, and returns its loss to pinned host memory.  Every step's inputs cross PCIe inside the timed region.
    dev_in = [(torch.empty_like(pool[0][0]), torch.empty_like(pool[0][1])) for _ in range(2)]      # N > 1: torch-side slots
    copy_stream = torch.cuda.Stream()

    def stage_first():
        img, lab, _ = host_ring[0]
        if world == 1:
            eng.stage_host_batch(0, img, lab)
        else:
            dev_in[0][0].copy_(img, non_blocking=True)
            dev_in[0][1].copy_(lab, non_blocking=True)

    def host_step(j):
        nimg, nlab, _ = host_ring[(j + 1) % RING]
        out = host_ring[j][2]

        def fn():
            if world == 1:
                eng.train_step_host_prefetch(j % 2, nimg, nlab, lr, loss_out=out)
                return
            cur = torch.cuda.current_stream()
            copy_stream.wait_stream(cur)
            with torch.cuda.stream(copy_stream):
                dev_in[(j + 1) % 2][0].copy_(nimg, non_blocking=True)
                dev_in[(j + 1) % 2][1].copy_(nlab, non_blocking=True)
            img, lab = dev_in[j % 2]
            if fused and args.model == "base":
                ex.train_step_(eng, img, lab, lr)
            else:
                eng.forward_backward(img, lab)
                exchange_and_update()
            cur.wait_stream(copy_stream)
            out.copy_(eng.loss(), non_blocking=True)
        return fn

    # Synthetic raw-byte inputs with lr 0.1 move the weights by >10x their scale in ONE update: the
    # second step from the same weights already runs on a dead or overflowing network.  Every timed step therefore
    # starts from the initial weights: region 1 restores them between its per-step event pairs (untimed); the
    # back-to-back regions 2 and 3 replay graphs whose first node is a device-side copy of the saved weights, and
    # that copy's cost is timed separately and reported (`restore_us_per_step`, included in value_hot_l2 and e2e).
    def reset():
        eng.params.copy_(params0)

    def restoring(fn):
        def run():
            reset()
            fn()
        return run

    s_img.copy_(pool[0][0]); s_lab.copy_(pool[0][1])
    graph = eng.capture(device_step)
    graph_r = eng.capture(restoring(device_step))
    graph_restore = eng.capture(reset)
    stage_first()
    host_graphs = [eng.capture(restoring(host_step(j))) for j in range(RING)]
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    flush_rd = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")      # 256 MiB, only ever read
    flush_mode = os.environ.get("B200_BENCH_FLUSH", "write")

    def flush_l2():
        """untimed, between timed steps: write a 256 MiB buffer (twice the 126 MB L2) so that nothing of the previous
        step survives in L2.  B200_BENCH_FLUSH=write+read (experiment) follows it with a read pass over a second 256 MiB
        buffer, which leaves the L2 full of CLEAN lines: the step then does not pay the write-back of the memset's dirty
        lines on every allocation."""
        flush_buf.zero_()
        if flush_mode == "write+read":
            flush_rd.sum()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up --------------------------------------------------------------------------
    for i in range(W):
        s_img.copy_(pool[i % POOL][0]); s_lab.copy_(pool[i % POOL][1])
        reset()
        graph.replay()
        graph_r.replay()
        host_graphs[i % RING].replay()
    reset()
    barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    # ---- timed region 1: device-resident inputs, L2 flushed between steps -------------------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    for i in range(K):
        reset()                                             # untimed: every step starts from the initial weights
        flush_l2()                                          # untimed: evicts weights/activations from the 126 MB L2
        a, b = ev[i]
        a.record()
        s_img.copy_(pool[i % POOL][0]); s_lab.copy_(pool[i % POOL][1])
        graph.replay()
        b.record()
    barrier()
    ms_cold = sum(a.elapsed_time(b) for a, b in ev)

    # ---- timed region 2: same steps back to back, L2 warm (extra information) ---------------------
    reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        s_img.copy_(pool[i % POOL][0]); s_lab.copy_(pool[i % POOL][1])
        graph_r.replay()
    e1.record()
    barrier()
    ms_hot = e0.elapsed_time(e1)
    loss_hot = float(eng.loss().item())
    e0.record()
    for i in range(K):
        graph_restore.replay()
    e1.record()
    barrier()
    ms_restore = e0.elapsed_time(e1)

    # ---- timed region 3: end to end through the public host-buffer call ---------------------------
    reset()
    barrier()
    e0.record()
    stage_first()                                           # the first batch's copy is inside the timed region too
    loss_seen = 0.0
    done = [torch.cuda.Event() for _ in range(RING)]
    for i in range(K):
        host_graphs[i % RING].replay()                      # H2D records+labels, step, D2H loss
        done[i % RING].record()
        if i:                                               # the loss of every step is read on the host (:230) -- one step
            done[(i - 1) % RING].synchronize()              # behind, so that the next step is already queued while the
            loss_seen = float(host_ring[(i - 1) % RING][2][0])      # host waits (no launch gap on the GPU)
    done[(K - 1) % RING].synchronize()
    loss_seen = float(host_ring[(K - 1) % RING][2][0])
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    if world > 1:      # after a real exchange step: every replica holds bit-identical parameters
        assert ex.replicas_identical(eng.params), "replicas diverged"

    # ---- live per-site timing for the roofline (eager, CUDA events between launch sites) --------
    reset()
    n_prof = 10
    acc = {}
    for i in range(n_prof + 2):
        reset()
        flush_l2()
        sites = eng.profile_step(pool[i % POOL][0], pool[i % POOL][1], lr)
        if i >= 2:
            for n, ms in sites:
                acc[n] = acc.get(n, 0.0) + ms / n_prof
    work = site_work(mode, model["dims"], model["param_bytes"])
    table = []
    for n, ms in acc.items():
        w = work.get(n)
        row = {"site": n, "ms": ms}
        if w and w[0] == "tensor":
            row.update(bound="tensor", tflops=2 * w[1] * B / (ms * 1e-3) / 1e12)
        elif w:
            row.update(bound="hbm", gbs=(w[1] * B + w[2]) / (ms * 1e-3) / 1e9)
        table.append(row)
    table.sort(key=lambda r: -r["ms"])
    # the roofline object is about ONE kernel: sites that time two launches together (the FC tail's rows + gradient
    # kernels, conv2's wgrad + partial reduction, conv1's wgrad + partial reduction -- folded into the update kernel in
    # the captured step, a launch of its own here --, the weight-image kernels, the layer-by-layer schedule's FC
    # backward products) stay in `sites` but are not candidates
    composite = {"fc_tail", "conv2_wgrad", "conv1_wgrad", "weights_prep", "full1_bwd", "full2_bwd", "full3_bwd", "full1_fwd"}
    dom = next((r for r in table if r["site"] not in composite), table[0])
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):          # dram__bytes_read+write per launch from the committed ncu --set full capture
        traffic = json.load(open(tpath)).get(dom["site"], {}).get("dram_bytes_per_launch")
    if dom.get("bound") == "tensor":
        roof = {"bound": "tensor", "kernel": dom["site"], "achieved": dom["tflops"], "peak": peaks["tflops_burst"],
                "unit": "TFLOP/s", "frac": dom["tflops"] / peaks["tflops_burst"], "traffic": traffic,
                "peak_source": peaks["source"] + " (burst: kernel timed alone)", "kernel_ms": dom["ms"]}
    else:
        roof = {"bound": "hbm", "kernel": dom["site"], "achieved": dom.get("gbs"), "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": (dom.get("gbs") or 0) / peaks["hbm_gbs"], "traffic": traffic,
                "peak_source": peaks["source"], "kernel_ms": dom["ms"]}

    # max over ranks
    t = torch.tensor([ms_cold, ms_hot, ms_e2e, ms_restore], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_cold, ms_hot, ms_e2e, ms_restore = (float(x) for x in t.cpu())
    import math
    assert math.isfinite(loss_seen) and math.isfinite(loss_hot), ("non-finite loss in a timed region", loss_seen, loss_hot)

    if rank == 0:
        imgs = B * world * K
        step_flops = model["flop"] * B
        act_b = 4 if mode == "fp32" else 2
        step_bytes = model["act"] * act_b * B + 3 * model["param_bytes"]
        ms_step = ms_cold / K
        out = {
            "metric": METRIC, "value": imgs / (ms_cold * 1e-3), "unit": "images/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if mode == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args), "model_widths": model["dims"], "batch_per_gpu": B, "global_batch": B * world,
                       "mode": mode, "parallelism": f"dp{world}", "reduce": args.reduce, "exchange": exchange,
                       "l2": "flushed between timed steps (256 MiB memset, untimed%s); per-step CUDA events summed"
                             % (" + 256 MiB read pass" if flush_mode == "write+read" else ""),
                       "cuda_graph": True},
            "value_hot_l2": imgs / (ms_hot * 1e-3),
            "restore_us_per_step": 1e3 * ms_restore / K,
            "clocks": clocks,
            "e2e": {"value": imgs / (ms_e2e * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": B * 3072 + B * 4, "d2h_bytes_per_step": 4,
                    "prefetch": True,
                    "note": "public host-buffer call with prefetch (b200_train_step_host_prefetch): every step copies the NEXT "
                            "batch's pinned host records H2D on a copy stream while it trains on the batch staged by the "
                            "previous call, then copies its loss D2H; K steps move K + 1 batches across PCIe inside the "
                            "timed region; the host reads each step's loss one "
                            "step behind (the next step is queued before it waits, so the GPU sees no launch gap); L2 not "
                            "flushed (inputs arrive from the host each step); each step starts with a device-side "
                            "restore of the initial weights (restore_us_per_step, inside the timed region)",
                    "last_loss": loss_seen},
            # kernels of this repo per step: the single-GPU step ends in sgd_kernel; with workers that launch is replaced
            # by the fused peer-memory update (two launches when the exchange is overlapped) or follows NCCL's kernels
            "gpu_launches": (eng.launches_per_step if world == 1 else
                             eng.launches_per_step - 1 + (2 if (fused and ex.overlap) else 1)) * K,
            "roofline": roof,
            "step_roofline": {"tensor_frac_of_sustained": step_flops / (ms_step * 1e-3) / 1e12 / peaks["tflops_sustained"],
                              "hbm_frac_layer_boundary_model": step_bytes / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"],
                              "flop_per_step": step_flops, "model_bytes_per_step": step_bytes},
            "sites": [{k: (round(v, 5) if isinstance(v, float) else v) for k, v in r.items()} for r in table[:8]],
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(B, args.cpu_seconds, dims=model["dims"])
            if B == 128 and args.model == "base":        # BASELINE.json configs[0] shape: 1 ps + 2 workers on localhost (README:11-13)
                try:
                    v2, sample2 = cpu_ps_workers(B, 2, max(8, int(args.cpu_seconds * out["cpu_baseline"]["value"] / B / 4)))
                    out["cpu_baseline_1ps_2workers"] = {"value": v2, "unit": "images/s", "cores": os.cpu_count(),
                                                        "kind": "port", "sample": sample2}
                except Exception as e:      # reported baseline only: never fail the bench line for it
                    out["cpu_baseline_1ps_2workers"] = {"unavailable": f"{type(e).__name__}: {e}"}
        if args.sites_json:
            os.makedirs(os.path.dirname(os.path.abspath(args.sites_json)), exist_ok=True)
            json.dump({"mode": mode, "batch": B, "sites": table, "ms_per_step_graph_cold": ms_step,
                       "ms_per_step_graph_hot": ms_hot / K}, open(args.sites_json, "w"), indent=1)
        print(json.dumps(out))
    if world > 1:
        # CUDA graphs that captured NCCL work are still alive here; tearing the communicator down under
        # them can deadlock, so release them first and leave without running further destructors.
        torch.cuda.synchronize()
        dist.barrier()
        del graph, graph_r, graph_restore, host_graphs
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
