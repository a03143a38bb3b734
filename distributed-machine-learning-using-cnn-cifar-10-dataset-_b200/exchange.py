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
