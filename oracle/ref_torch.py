"""Second, independent CPU restatement (TEST INFRASTRUCTURE ONLY) of cifar10cnn.py:94-176,
written with torch-CPU ops and autograd instead of hand-derived gradients.

Two jobs:
  1. cross-check oracle/ref_numpy.py (agreement required to <= 1e-10 relative in fp64) --
     the two share no arithmetic code, so a semantic slip in one shows up as a mismatch;
  2. the timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`), in fp32 with
     all host threads, because the reference's own TensorFlow-1.x CPU path cannot be
     installed here.  kind = "port".

PARITY UNPINNED by the reference (no tests or golden vectors exist there); see the header
of ref_numpy.py.  Never imported by the product path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from .ref_numpy import PARAM_NAMES


def to_torch(params, dtype=torch.float64, requires_grad=True):
    return {k: torch.tensor(np.asarray(v), dtype=dtype).requires_grad_(requires_grad)
            for k, v in params.items()}


def crop_cast(images_u8_chw, crop_h=24, crop_w=24, dtype=torch.float64):
    """cifar10cnn.py:57-68: CHW uint8 -> HWC float, centre crop.  Returned NCHW for torch."""
    x = torch.as_tensor(np.asarray(images_u8_chw))
    oh = (x.shape[2] - crop_h) // 2
    ow = (x.shape[3] - crop_w) // 2
    return x[:, :, oh:oh + crop_h, ow:ow + crop_w].to(dtype).contiguous()


def _pool(x):
    """max_pool ksize 3, stride 2, SAME (cifar10cnn.py:113/:123): pad 0 before, 1 after on
    even sizes, padded cells are -inf so they never win."""
    H, W = x.shape[2], x.shape[3]
    OH, OW = -(-H // 2), -(-W // 2)
    pth = max((OH - 1) * 2 + 3 - H, 0)
    ptw = max((OW - 1) * 2 + 3 - W, 0)
    x = F.pad(x, (ptw // 2, ptw - ptw // 2, pth // 2, pth - pth // 2), value=float("-inf"))
    return F.max_pool2d(x, kernel_size=3, stride=2)


def create_cnn(p, x_nchw):
    """cifar10cnn.py:94-147 with NCHW activations; filters HWIO -> OIHW; flatten in (h,w,c)
    order (:126) via permute."""
    w1 = p["conv1_kernel"].permute(3, 2, 0, 1)
    w2 = p["conv2_kernel"].permute(3, 2, 0, 1)
    h = F.relu(F.conv2d(x_nchw, w1, p["conv1_bias"], padding=2))
    h = _pool(h)
    h = F.relu(F.conv2d(h, w2, p["conv2_bias"], padding=2))
    h = _pool(h)
    h = h.permute(0, 2, 3, 1).reshape(x_nchw.shape[0], -1)
    h = F.relu(h @ p["full_weight_1"] + p["full_bias_1"])
    h = F.relu(h @ p["full_weight_2"] + p["full_bias_2"])
    return F.relu(h @ p["full_weight_3"] + p["full_bias_3"])          # :145 ReLU on logits


def cifar_loss(logits, labels):
    """cifar10cnn.py:150-157."""
    t = torch.as_tensor(np.asarray(labels)).reshape(-1).long()
    return F.cross_entropy(logits, t, reduction="mean")


def loss_and_grads(params, images_u8_chw, labels, dtype=torch.float64, crop=(24, 24)):
    p = to_torch(params, dtype)
    x = crop_cast(images_u8_chw, crop[0], crop[1], dtype)
    logits = create_cnn(p, x)
    loss = cifar_loss(logits, labels)
    loss.backward()
    return (float(loss.detach()), logits.detach().numpy(),
            {k: p[k].grad.detach().numpy() for k in PARAM_NAMES})


class CpuTrainer:
    """fp32 training step on the host cores -- the timed CPU baseline.  One step =
    crop/cast + forward + backward + SGD apply (cifar10cnn.py:230 for one worker)."""

    def __init__(self, params, lr=0.1, threads=None):
        if threads:
            torch.set_num_threads(threads)
        self.p = to_torch(params, torch.float32)
        self.p0 = {k: v.detach().clone() for k, v in self.p.items()}
        self.lr = lr

    def reset(self):
        """restore the initial weights (raw-byte inputs with lr 0.1 diverge within a few steps, as
        they do in the reference -- SURVEY.md Q11 -- and inf/nan arithmetic would distort timing)."""
        with torch.no_grad():
            for k, v in self.p.items():
                v.copy_(self.p0[k])

    def step(self, images_u8_chw, labels):
        x = crop_cast(images_u8_chw, dtype=torch.float32)
        loss = cifar_loss(create_cnn(self.p, x), labels)
        grads = torch.autograd.grad(loss, [self.p[k] for k in PARAM_NAMES])
        with torch.no_grad():
            for k, g in zip(PARAM_NAMES, grads):
                self.p[k].sub_(self.lr * g)
        return float(loss.detach())
