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
    the ten variables; each tower has its own activations.  With more than one worker (an initialised NCCL
    process group) the variables and gradients of the first tower live in peer-mapped symmetric memory, so
    that the ps push/apply can run as the fused peer-memory kernel (exchange.P2PExchange); where symmetric
    memory is unavailable the session falls back to an NCCL allreduce + SGD kernel."""

    def __init__(self, mode=None, seed=0, reduce="sum"):
        self.mode, self.seed, self.reduce = mode, seed, reduce
        self.engines = []
        self.exchange = None

    def _make_exchange(self):
        import torch.distributed as dist
        from .exchange import P2PExchange
        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            return GradientExchange(reduce=self.reduce)
        if dist.get_backend() == "nccl":
            try:
                total = TrainStepEngine(batch=8, mode="fp32").total
                return P2PExchange(total, torch.device("cuda", torch.cuda.current_device()), reduce=self.reduce)
            except Exception as e:       # no symmetric memory / NVLink peer access here: NCCL is still all-CUDA
                import sys
                print(f"b200cnn: peer-memory exchange unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
        return GradientExchange(reduce=self.reduce)

    def tower(self, batch_size):
        from . import DEFAULT_MODE
        first = self.engines[0] if self.engines else None
        kw = {}
        if first is None:
            self.exchange = self._make_exchange()
            if hasattr(self.exchange, "params"):
                kw = {"params": self.exchange.params, "grads": self.exchange.grads}
        else:
            kw = {"params": first.params}
        eng = TrainStepEngine(batch=batch_size, mode=self.mode or DEFAULT_MODE, **kw)
        eng.exchange = self.exchange
        if first is None:
            eng.init_params(self.seed)
        self.engines.append(eng)
        return eng


def input_pipeline(batch_size, train_logical=True, seed=0, data_dir=None, synthetic=None):
    """cifar10cnn.py:72-91."""
    kw = {} if data_dir is None else {"data_dir": data_dir}
    src = BatchSource(batch_size, train_logical, seed=seed, synthetic=synthetic, **kw)
    return ImagesNode(src), LabelsNode(src)


class LossValue:
    """What `mon_sess.run([train_op, loss])` (:230) returns for the loss: a number that is fetched from its pinned
    host slot the first time somebody looks at it.  The step itself never waits for the GPU, so the next step is
    queued while this one runs; float(), comparisons, arithmetic and printing all resolve it."""
    __slots__ = ("_buf", "_event", "_value")

    def __init__(self, buf, event):
        self._buf, self._event, self._value = buf, event, None

    def resolve(self):
        if self._value is None:
            self._event.synchronize()
            self._value = float(self._buf[0])
            self._buf = self._event = None
        return self._value

    def __float__(self):
        return self.resolve()

    def __eq__(self, other):
        return self.resolve() == float(other)

    def __lt__(self, other):
        return self.resolve() < float(other)

    def __le__(self, other):
        return self.resolve() <= float(other)

    def __gt__(self, other):
        return self.resolve() > float(other)

    def __ge__(self, other):
        return self.resolve() >= float(other)

    def __hash__(self):
        return hash(self.resolve())

    def __add__(self, other):
        return self.resolve() + other

    __radd__ = __add__

    def __sub__(self, other):
        return self.resolve() - other

    def __rsub__(self, other):
        return other - self.resolve()

    def __mul__(self, other):
        return self.resolve() * other

    __rmul__ = __mul__

    def __truediv__(self, other):
        return self.resolve() / other

    def __repr__(self):
        return repr(self.resolve())

    def __format__(self, spec):
        return format(self.resolve(), spec)


class _GraphStep:
    """The body of `mon_sess.run([train_op, loss])` as it is benchmarked: one CUDA graph per input buffer holding
    forward + backward + exchange + update (+ the loss copy to a pinned slot), the next batch drawn from the shuffle
    queue on a side stream while the current step runs, no host synchronisation."""

    def __init__(self, eng, exchange, source, lr, staleness=0):
        self.eng, self.ex, self.source, self.lr = eng, exchange, source, lr
        dev = eng.params.device
        # asynchronous-ps fidelity (cifar10cnn.py:162-163 without SyncReplicasOptimizer): a gradient reaches the variables
        # `staleness` exchange rounds after the snapshot it was computed on.  A FIFO of that many flat gradient buffers
        # sits between the backward pass and the exchange; 0 = the synchronous default.
        self.stale_k = int(staleness)
        if self.stale_k:
            self._fifo = torch.zeros(self.stale_k, eng.grads.numel(), dtype=torch.float32, device=dev)
            self._fresh = torch.zeros_like(eng.grads)
        B = eng.batch
        c = eng.cfg
        self.bufs = [(torch.empty(B, c.c_in, c.img_h, c.img_w, dtype=torch.uint8, device=dev),
                      torch.empty(B, dtype=torch.int32, device=dev)) for _ in range(2)]
        for img, lab in self.bufs:
            img.zero_(); lab.zero_()
        self.loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.side = torch.cuda.Stream(device=dev)
        self.pending = [None, None]
        self.used = [False, False]
        self.i = 0
        if exchange.world > 1 and not self.stale_k and getattr(exchange, "overlap", False) and hasattr(exchange, "bind"):
            exchange.bind(eng)              # inbox of the one-shot conv-gradient exchange (collective; outside the capture)
        # capturing runs the step once eagerly (library side streams are created then) and once more under
        # capture-time validation: keep the variables as they were
        keep = eng.params.clone()
        self.graphs = [eng.capture(lambda k=k: self._body(k)) for k in range(2)]
        eng.params.copy_(keep)
        if self.stale_k:
            self._fifo.zero_()                                     # (the capture runs pushed garbage through the line)
        torch.cuda.synchronize(dev)
        self._prefetch(0)

    def _body(self, k):
        img, lab = self.bufs[k]
        eng, ex = self.eng, self.ex
        if self.stale_k:
            eng.forward_backward(img, lab)
            self._fresh.copy_(eng.grads)
            eng.grads.copy_(self._fifo[0])                         # the gradient computed stale_k rounds ago
            for j in range(self.stale_k - 1):
                self._fifo[j].copy_(self._fifo[j + 1])
            self._fifo[self.stale_k - 1].copy_(self._fresh)
            if ex.world > 1 and hasattr(ex, "apply_"):
                ex.apply_(eng, self.lr)
            else:
                ex.allreduce_(eng.grads)
                eng.sgd_apply(self.lr, ex.grad_scale)
        elif ex.world == 1:
            eng.train_step(img, lab, self.lr)
        elif hasattr(ex, "train_step_"):
            ex.train_step_(eng, img, lab, self.lr)                 # fused peer-memory exchange + update
        else:
            eng.forward_backward(img, lab)
            ex.allreduce_(eng.grads)
            eng.sgd_apply(self.lr, ex.grad_scale)
        self.loss_host[k].copy_(eng.loss(), non_blocking=True)

    def _prefetch(self, k):
        """draw the next batch into buffer k on the side stream (after the step that last read it)"""
        if self.used[k]:
            self.side.wait_event(self.done[k])
        img, lab = self.bufs[k]
        self.source.next_into(img, lab, stream=self.side.cuda_stream)
        self.ready[k].record(self.side)

    def step(self):
        k = self.i & 1
        main = torch.cuda.current_stream()
        main.wait_event(self.ready[k])
        if self.pending[k] is not None:
            self.pending[k].resolve()                               # two steps old: frees the pinned slot
        self.graphs[k].replay()
        self.done[k].record(main)
        self.used[k] = True
        lv = LossValue(self.loss_host[k], self.done[k])
        self.pending[k] = lv
        self._prefetch(1 - k)
        self.i += 1
        return lv


class MonitoredTrainingSession:
    """run() executes the fused step; should_stop() is StopAtStepHook(last_step) (:219).  global_step
    counts WORKER updates like the reference's shared counter: one exchange round applies
    `world` of them."""

    def __init__(self, is_chief=True, last_step=20000, reduce="sum", checkpoint_dir=None, save_secs=600.0, staleness=0):
        if staleness < 0 or staleness > 64:
            raise ValueError("staleness must be in 0..64 exchange rounds")
        self.staleness = int(staleness)
        self.is_chief = is_chief
        self.last_step = last_step
        self.reduce = reduce
        self.exchange = None
        self.global_step = 0
        self._synced = False
        self._engine = None
        self._steps = {}
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
        self.exchange = getattr(eng, "exchange", None) or GradientExchange(reduce=self.reduce)
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

    def _train(self, op):
        eng = op.loss.logits.engine
        if op.loss.logits.images.source is not op.loss.labels.source:
            raise ValueError("images and labels must come from the same input_pipeline")
        lr = op.learning_rate()
        key = (id(eng), lr)
        st = self._steps.get(key)
        if st is None:
            if self.exchange.world > 1:
                import torch.distributed as dist
                dist.barrier()
            st = self._steps[key] = _GraphStep(eng, self.exchange, op.loss.labels.source, lr, self.staleness)
        return st.step()

    def run(self, fetches):
        single = not isinstance(fetches, (list, tuple))
        items = [fetches] if single else list(fetches)
        out = [None] * len(items)
        train = [f for f in items if isinstance(f, TrainOp)]
        done = {}
        step_loss = None
        if train:
            op = train[0]
            eng = op.loss.logits.engine
            self._engine = eng
            if not self._synced:
                self._first_sync(eng)
                if self.should_stop():
                    return None if single else [None] * len(items)
            step_loss = self._train(op)
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
                    step_is_train = False
                else:
                    step_is_train = train and f.logits is train[0].loss.logits
                if isinstance(f, LossNode):
                    out[i] = step_loss if step_is_train else float(eng.loss().item())
                else:
                    out[i] = float(eng.n_correct().item()) / eng.batch
            else:
                raise TypeError(f"cannot fetch {type(f).__name__}")
        return out[0] if single else out
