"""Generates tests/golden/*.npz.  Run from the repo root: python tests/golden/make_golden.py

The reference (cifar10cnn.py) cannot be imported here (TensorFlow 1.x is not installable on
Python 3.12), so the vectors come from oracle/ref_torch.py -- the torch-autograd fp64
restatement -- and are then used to pin oracle/ref_numpy.py (hand-derived gradients) and the
CUDA path.  Gradients are stored as digests (sum, abs-sum, 256 sampled entries) to keep the
fixtures small; logits, loss, predictions and inputs are stored in full.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import ref_numpy as rn, ref_torch as rt  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sample_idx(n, k=256):
    return np.unique(np.linspace(0, n - 1, k).astype(np.int64))


def digest(arr):
    f = np.asarray(arr, dtype=np.float64).reshape(-1)
    return np.array([f.sum(), np.abs(f).sum()]), f[sample_idx(f.size)]


def make(name, B, kind, seed):
    rng = np.random.default_rng(seed)
    params = rn.init_params(0)
    labels = rng.integers(0, 10, (B, 1)).astype(np.int32)
    if kind == "u8":      # CIFAR-record-shaped raw bytes (cifar10cnn.py:57-62)
        images = rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)
    else:                 # low-contrast bytes: keeps activations O(1), gradients well conditioned
        images = np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
    loss, logits, grads = rt.loss_and_grads(params, images, labels)
    out = {"images": images, "labels": labels, "logits": logits, "loss": np.float64(loss),
           "pred": logits.argmax(1).astype(np.int32)}
    for k in rn.PARAM_NAMES:
        out["psum_" + k], out["psmp_" + k] = digest(params[k])
        out["gsum_" + k], out["gsmp_" + k] = digest(grads[k])
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "loss", loss, "zero-logit frac", float((logits == 0).mean()))


if __name__ == "__main__":
    make("step_u8_b8.npz", 8, "u8", 101)
    make("step_soft_b8.npz", 8, "soft", 202)
