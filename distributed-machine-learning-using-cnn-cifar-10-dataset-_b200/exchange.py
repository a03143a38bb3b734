"""The parameter-server gradient push/apply of cifar10cnn.py (:162-163 executed on /job:ps,
placement :195-196) recast as one collective per step.

The reference is asynchronous: each worker pulls the variables, pushes its own mean-gradient
and the ps applies it without locking.  Here every worker contributes its gradient to one
allreduce and applies the result locally with the fused SGD kernel, so replicas stay
bit-identical and no parameter pull is needed.  reduce="sum" (default) applies the SUM of the N
worker gradients with the script's lr -- exactly N reference updates computed from one
parameter snapshot -- and advances global_step by N; reduce="mean" averages instead.

torch.distributed is plumbing: NCCL over NVLink on the GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradientExchange:
    def __init__(self, reduce="sum", group=None):
        if reduce not in ("sum", "mean"):
            raise ValueError("reduce must be 'sum' or 'mean'")
        self.reduce = reduce
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0

    @property
    def grad_scale(self):
        """factor the SGD kernel folds into lr (b200_sgd_apply's grad_scale)."""
        return 1.0 if self.reduce == "sum" else 1.0 / self.world

    @property
    def global_step_increment(self):
        """one allreduce round = `world` worker updates of the reference's shared global_step (:163)."""
        return self.world

    def allreduce_(self, flat_grads: torch.Tensor):
        """in-place sum of the flat gradient buffer over all workers (bucket = the whole 4.27 MB:
        the message is latency-bound, one launch is cheapest)."""
        if self.world > 1:
            dist.all_reduce(flat_grads, op=dist.ReduceOp.SUM, group=self.group)
        return flat_grads

    def broadcast_params_(self, flat_params: torch.Tensor, src=0):
        """chief (task 0, cifar10cnn.py:222) publishes the initial / restored variables once."""
        if self.world > 1:
            dist.broadcast(flat_params, src=src, group=self.group)
        return flat_params

    def replicas_identical(self, flat_params: torch.Tensor) -> bool:
        """checksum allreduce: True when every replica holds bit-identical parameters."""
        if self.world == 1:
            return True
        v = flat_params.view(torch.int32).to(torch.int64)
        s = torch.stack([v.sum(), (v * (torch.arange(v.numel(), device=v.device) % 8191 + 1)).sum()])
        lo, hi = s.clone(), s.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
        return bool((lo == hi).all())


class P2PExchange(GradientExchange):
    """The same exchange as ONE kernel per GPU over NVLink peer memory (csrc/p2p_update.cu): the all-reduce
    of the flat gradient buffer is fused with the SGD update and the updated shard is written straight into
    every worker's parameter buffer -- no NCCL call, no separate update launch.  With an NVSwitch multicast
    mapping the switch does the additions (multimem.ld_reduce / multimem.st).

    torch.distributed._symmetric_memory is the plumbing: it allocates the peer-mapped gradient / parameter
    buffers and exchanges their addresses.  Usage: create it first, build the engine on `self.params` /
    `self.grads` (TrainStepEngine(params=..., grads=...)), then call apply_(engine, lr) once per step on
    every worker."""

    def __init__(self, total_floats, device, reduce="sum", group=None, use_multicast=True, overlap=None):
        super().__init__(reduce=reduce, group=group)
        if self.world < 2:
            raise ValueError("P2PExchange needs an initialised process group with at least 2 workers")
        if self.world > 8 or total_floats % (4 * self.world):
            raise ValueError("P2PExchange supports 2..8 workers and sizes divisible by 4*world")
        import ctypes as C
        import os
        import torch.distributed._symmetric_memory as symm_mem
        grp = group if group is not None else dist.group.WORLD
        self.grads = symm_mem.empty(total_floats, dtype=torch.float32, device=device)
        self.params = symm_mem.empty(total_floats, dtype=torch.float32, device=device)
        self.grads.zero_()
        self.params.zero_()
        self._hg = symm_mem.rendezvous(self.grads, grp)
        self._hp = symm_mem.rendezvous(self.params, grp)
        n = self.world
        arr = C.c_void_p * n
        self._peer_grads = arr(*[int(x) for x in self._hg.buffer_ptrs])
        self._peer_params = arr(*[int(x) for x in self._hp.buffer_ptrs])
        # flags: a private corner of the gradient handle's signal pads (torch's own barrier uses the start)
        off = 2048
        if self._hg.signal_pad_size < off + 128:
            raise RuntimeError("signal pad too small")
        self._peer_pads = arr(*[int(x) + off for x in self._hg.signal_pad_ptrs])
        mc_ok = use_multicast and self._hg.has_multicast_support and self._hp.has_multicast_support
        self.multicast = bool(mc_ok and self._hg.multicast_ptr and self._hp.multicast_ptr)
        self._mc_grads = int(self._hg.multicast_ptr) if self.multicast else None
        self._mc_params = int(self._hp.multicast_ptr) if self.multicast else None
        # inbox of the one-shot conv-gradient exchange (second flag channel): [2 parities][world][conv range] floats
        self._inbox = None                                # allocated by bind(), once the engine's conv range is known
        self._grp = grp
        self._device = device
        self._state = torch.zeros(8, dtype=torch.int32, device=device)      # (epoch, arrivals, reducers done, -) per flag channel
        # Overlapping the exchange with the backward pass costs a second pair of flag barriers.  Measured (whole step,
        # batch 128/GPU, L2 flushed): 2 GPUs 140 us; 4 GPUs 142 us overlapped vs 148 us as one kernel at the end; 8 GPUs
        # 144 vs 150 -- so it is the default at every world size (round 1 had it off at eight workers, where the
        # reduction launch in front of the second channel ate the gain; that reduction now runs inside the kernel).
        import os
        env = os.environ.get("B200CNN_OVERLAP")          # experiments: force the overlapped (1) / single-kernel (0) exchange
        self.overlap = (True if env is None else env == "1") if overlap is None else bool(overlap)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)

    INBOX_MAX_FLOATS = 1 << 18       # 1 MB of conv gradients (the script's widths: 107,392 floats)

    def bind(self, engine):
        """Give `engine` (built on self.params / self.grads) the inbox of the one-shot conv-gradient exchange: 2 x world x
        (offset of full_weight_1) floats of symmetric memory.  Collective: every worker calls it once."""
        import ctypes as C
        import os
        import torch.distributed._symmetric_memory as symm_mem
        from ._lib import check
        if getattr(engine, "_inbox_bound", False) or os.environ.get("B200CNN_INBOX", "1") != "1":
            return
        n_conv = int(engine.offsets[4])                  # full_weight_1 starts where the conv tensors end
        if n_conv > self.INBOX_MAX_FLOATS:               # every rank receives world x the range: only pays while it is small
            engine._inbox_bound = True
            return
        floats = 2 * self.world * n_conv
        if self._inbox is not None and self._inbox_floats != floats:
            raise ValueError("P2PExchange.bind: engines of one exchange must have the same conv-gradient range")
        if self._inbox is None:                          # (a second engine on this exchange shares the inbox: one step at a time)
            self._inbox = symm_mem.empty(floats, dtype=torch.float32, device=self._device)
            self._inbox.zero_()
            hi = symm_mem.rendezvous(self._inbox, self._grp)
            arr = (C.c_void_p * self.world)(*[int(x) for x in hi.buffer_ptrs])
            mc = int(hi.multicast_ptr) if (self.multicast and hi.has_multicast_support and hi.multicast_ptr) else None
            self._inbox_handle, self._inbox_peers, self._inbox_mc, self._inbox_floats = hi, arr, mc, floats
        check(engine.lib.b200_p2p_set_inbox(engine.h, self._inbox.data_ptr(), self._inbox_mc, self._inbox_peers, self.world,
                                            floats))
        engine._inbox_bound = True
        torch.cuda.synchronize(self._device)
        dist.barrier(group=self.group)

    def apply_(self, engine, lr):
        """allreduce + update in one launch on the current stream (replaces allreduce_ + engine.sgd_apply)."""
        from ._lib import check
        check(engine.lib.b200_p2p_sgd_apply(engine.h, self._peer_grads, self._peer_params, self._peer_pads,
                                            self._mc_grads, self._mc_params, self._state.data_ptr(), self.rank,
                                            self.world, lr, self.grad_scale, torch.cuda.current_stream().cuda_stream))

    def train_step_(self, engine, images, labels, lr):
        """forward + backward + exchange + update with the exchange overlapped with the conv layers' backward
        kernels (b200_train_step_p2p); same result as engine.forward_backward(...) followed by apply_(engine, lr)."""
        from ._lib import check
        engine._check_inputs(images, labels)
        if self.overlap:
            self.bind(engine)
        fn = engine.lib.b200_train_step_p2p if self.overlap else engine.lib.b200_train_step_p2p_single
        check(fn(engine.h, images.data_ptr(), labels.data_ptr(), self._peer_grads,
                                             self._peer_params, self._peer_pads, self._mc_grads, self._mc_params,
                                             self._state.data_ptr(), self.rank, self.world, lr, self.grad_scale,
                                             torch.cuda.current_stream().cuda_stream))
