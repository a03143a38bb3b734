"""Checkpoint / resume with the contract of cifar10cnn.py:222 (`MonitoredTrainingSession(checkpoint_dir=
FLAGS.log_dir, is_chief=...)`): the chief restores the newest checkpoint in --log_dir at start and saves one
every `save_secs` (TF default 600 s) and at the end.  Files are plain .npz keyed by the variable names a TF
checkpoint of the reference would carry (`model_definition/<scope>/<name>`, :204 + :105-144, plus `global_step`
:29), in TF layouts (HWIO, [in,out]), so they can be converted either way with a few lines of numpy."""
from __future__ import annotations

import glob
import os
import re
import time

import numpy as np

from .engine import PARAM_NAMES, PARAM_SCOPES

PREFIX = "model.ckpt-"


def variable_names():
    return [f"model_definition/{s}/{n}" for s, n in zip(PARAM_SCOPES, PARAM_NAMES)]


def save(checkpoint_dir, params, global_step):
    """params: {short name: array}.  Writes model.ckpt-<step>.npz atomically and returns the path."""
    os.makedirs(checkpoint_dir, exist_ok=True)
    blob = {tf: np.asarray(params[n], dtype=np.float32) for tf, n in zip(variable_names(), PARAM_NAMES)}
    blob["global_step"] = np.int64(global_step)
    blob["Variable"] = np.int32(0)            # generation_num (:216): never incremented by the reference
    path = os.path.join(checkpoint_dir, f"{PREFIX}{int(global_step)}.npz")
    tmp = path + ".tmp.npz"
    np.savez(tmp, **blob)
    os.replace(tmp, path)
    return path


def latest(checkpoint_dir):
    best, best_step = None, -1
    for p in glob.glob(os.path.join(checkpoint_dir, PREFIX + "*.npz")):
        m = re.search(re.escape(PREFIX) + r"(\d+)\.npz$", p)
        if m and int(m.group(1)) > best_step:
            best, best_step = p, int(m.group(1))
    return best


def load(path):
    """-> ({short name: array}, global_step)"""
    z = np.load(path)
    params = {n: z[tf] for tf, n in zip(variable_names(), PARAM_NAMES)}
    return params, int(z["global_step"])


class Saver:
    """chief-only periodic saver (CheckpointSaverHook stand-in)."""

    def __init__(self, checkpoint_dir, save_secs=600.0):
        self.dir, self.save_secs = checkpoint_dir, save_secs
        self._last = time.time()

    def maybe_save(self, engine, global_step, force=False):
        if not self.dir or (not force and time.time() - self._last < self.save_secs):
            return None
        self._last = time.time()
        return save(self.dir, engine.get_params(), global_step)
