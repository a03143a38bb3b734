"""A minimal stand-in for the TF-1.x graph/session surface that cifar10cnn.py uses, so that the
drop-in script keeps the reference's structure: input_pipeline -> create_cnn -> cifar_loss /
batch_accuracy -> train_step -> `mon_sess.run([train_op, loss])`.

Nodes are plain Python objects; all arithmetic happens in the CUDA library when the session runs."""
from __future__ import annotations

import numpy as np
import torch

from . import checkpoint
from .data import BatchSource
from .engine import TrainStepEngine
from .exchange import GradientExchange

LEARNING_RATE, LR_DECAY, NUM_GENS_TO_WAIT = 0.1, 0.9, 250.0       # cifar10cnn.py:21-23


class ImagesNode:
    def __init__(self, source):
        self.source = source


class LabelsNode:
    def __init__(self, source):
        self.source = source


class LogitsNode:
    def __init__(self, images, engine):
        self.images, self.engine = images, engine


class LossNode:
    def __init__(self, logits, labels):
        self.logits, self.labels = logits, labels


class AccuracyNode:
    def __init__(self, logits, labels):
        self.logits, self.labels = logits, labels


class Variable:
    """tf.Variable(0, trainable=False) of :216 -- never incremented by the reference."""

    def __init__(self, value=0):
        self.value = value


class TrainOp:
    def __init__(self, loss, generation_num):
        self.loss, self.generation_num = loss, generation_num

    def learning_rate(self):
        """exponential_decay(..., staircase=True) of :161, evaluated in fp32 like TF."""
        p = np.floor(np.float32(self.generation_num.value) / np.float32(NUM_GENS_TO_WAIT))
        return float(np.float32(LEARNING_RATE) * np.power(np.float32(LR_DECAY), p, dtype=np.float32))


class ModelScope:
    """tf.variable_scope('model_definition') with reuse (:204-210): towers built in one scope share
    the ten variables; each tower has its own activations."""

    def __init__(self, mode=None, seed=0):
        self.mode, self.seed = mode, seed
        self.engines = []

    def tower(self, batch_size):
        from . import DEFAULT_MODE
        first = self.engines[0] if self.engines else None
        eng = TrainStepEngine(batch=batch_size, mode=self.mode or DEFAULT_MODE,
                              params=None if first is None else first.params)
        if first is None:
            eng.init_params(self.seed)
        self.engines.append(eng)
        return eng


def input_pipeline(batch_size, train_logical=True, seed=0, data_dir=None, synthetic=None):
    """cifar10cnn.py:72-91."""
    kw = {} if data_dir is None else {"data_dir": data_dir}
    src = BatchSource(batch_size, train_logical, seed=seed, synthetic=synthetic, **kw)
    return ImagesNode(src), LabelsNode(src)


class MonitoredTrainingSession:
    """run() executes the fused step; should_stop() is StopAtStepHook(last_step) (:219).  global_step
    counts WORKER updates like the reference's shared counter: one allreduce round applies
    `world` of them."""

    def __init__(self, is_chief=True, last_step=20000, reduce="sum", checkpoint_dir=None, save_secs=600.0):
        self.is_chief = is_chief
        self.last_step = last_step
        self.exchange = GradientExchange(reduce=reduce)
        self.global_step = 0
        self._synced = False
        self._engine = None
        self._saver = checkpoint.Saver(checkpoint_dir, save_secs) if (is_chief and checkpoint_dir) else None
        self._restore_from = checkpoint.latest(checkpoint_dir) if (is_chief and checkpoint_dir) else None

    def __enter__(self):
        return self

    def __exit__(self, exc_type, *exc):
        if exc_type is None and self._saver and self._engine is not None:
            self._saver.maybe_save(self._engine, self.global_step, force=True)       # final checkpoint, like TF
        return False

    def _first_sync(self, eng):
        """chief restores the newest checkpoint (or keeps its fresh init) and publishes variables and
        global_step to the other workers (:222)."""
        if self._restore_from:
            params, step = checkpoint.load(self._restore_from)
            eng.load_params(params)
            self.global_step = step
        self.exchange.broadcast_params_(eng.params)
        gs = torch.tensor([self.global_step], dtype=torch.int64, device=eng.params.device)
        if self.exchange.world > 1:
            import torch.distributed as dist
            dist.broadcast(gs, src=0)
        self.global_step = int(gs.item())
        self._synced = True

    def should_stop(self):
        return self.global_step >= self.last_step

    def _forward(self, logits_node, labels_node):
        if logits_node.images.source is not labels_node.source:
            raise ValueError("images and labels must come from the same input_pipeline")
        img, lab = labels_node.source.next()
        logits_node.engine.forward(img, lab)
        return logits_node.engine

    def run(self, fetches):
        single = not isinstance(fetches, (list, tuple))
        items = [fetches] if single else list(fetches)
        out = [None] * len(items)
        train = [f for f in items if isinstance(f, TrainOp)]
        done = {}
        if train:
            op = train[0]
            eng = op.loss.logits.engine
            self._engine = eng
            if not self._synced:
                self._first_sync(eng)
                if self.should_stop():
                    return None if single else [None] * len(items)
            self._forward(op.loss.logits, op.loss.labels)
            eng.backward()
            self.exchange.allreduce_(eng.grads)
            eng.sgd_apply(op.learning_rate(), self.exchange.grad_scale)
            self.global_step += self.exchange.global_step_increment
            done[id(op.loss.logits)] = eng
            if self._saver:
                self._saver.maybe_save(eng, self.global_step)
        for i, f in enumerate(items):
            if isinstance(f, TrainOp):
                continue
            if isinstance(f, (LossNode, AccuracyNode)):
                eng = done.get(id(f.logits))
                if eng is None:                                  # a separate run() draws a new batch
                    eng = self._forward(f.logits, f.labels)
                    done[id(f.logits)] = eng
                if isinstance(f, LossNode):
                    out[i] = float(eng.loss().item())
                else:
                    out[i] = float(eng.n_correct().item()) / eng.batch
            else:
                raise TypeError(f"cannot fetch {type(f).__name__}")
        return out[0] if single else out
