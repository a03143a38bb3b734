"""TrainStepEngine: the host-side owner of one GPU's replica of the cifar10cnn.py model.

PyTorch is the allocator / stream / graph shell only: it allocates the flat parameter,
gradient and workspace buffers, hands raw pointers to libb200cnn.so (include/b200cnn.h) and
captures the launches into a CUDA graph.  All arithmetic happens in the library's kernels.
There is no CPU path: constructing an engine without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import Config, check

# creation order of the variables in create_cnn (cifar10cnn.py:106-144); the scopes are the ones a
# TF checkpoint of the reference would carry (model_definition/<scope>/<name>, :204 + :105-142).
PARAM_NAMES = ("conv1_kernel", "conv1_bias", "conv2_kernel", "conv2_bias", "full_weight_1",
               "full_bias_1", "full_weight_2", "full_bias_2", "full_weight_3", "full_bias_3")
PARAM_SCOPES = ("conv1", "conv1", "conv2", "conv2", "full1", "full1", "full2", "full2", "full3", "full3")

_DT = {4: torch.float32, 1: torch.uint8, 2: torch.bfloat16}
_INT_TENSORS = {"pred", "n_correct", "n_bad", "stage_labels", "stage_labels1"}


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("b200cnn needs a CUDA (sm_100a) device: there is no CPU fallback on the product path")


class TrainStepEngine:
    def __init__(self, batch=128, img_hw=(32, 32), crop_hw=(24, 24), c_in=3, c1=64, c2=64, f1=384,
                 f2=192, n_cls=10, mode="fp32", device=None, params=None, grads=None):
        _require_cuda()
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        torch.cuda.set_device(self.device)
        self.mode = {"fp32": _lib.MODE_FP32, "bf16": _lib.MODE_BF16}[mode]
        self.cfg = Config(batch, img_hw[0], img_hw[1], crop_hw[0], crop_hw[1], c_in, c1, c2, f1, f2, n_cls,
                          self.mode)
        self.batch = batch
        self.n_cls = n_cls
        h = C.c_void_p()
        check(self.lib.b200_create(C.byref(self.cfg), C.byref(h)))
        self.h = h
        offs = (C.c_int64 * 10)()
        sizes = (C.c_int64 * 10)()
        total = C.c_int64()
        check(self.lib.b200_param_layout(h, offs, sizes, C.byref(total)))
        self.offsets, self.sizes, self.total = list(offs), list(sizes), total.value
        h2 = -(-(-(-crop_hw[0] // 2)) // 2)
        w2 = -(-(-(-crop_hw[1] // 2)) // 2)
        self.shapes = dict(zip(PARAM_NAMES, [
            (5, 5, c_in, c1), (c1,), (5, 5, c1, c2), (c2,), (h2 * w2 * c2, f1), (f1,), (f1, f2), (f2,),
            (f2, n_cls), (n_cls,)]))
        if params is not None:      # a second tower sharing the variables (scope.reuse_variables(), :209)
            if params.numel() != self.total or params.device != self.device or params.dtype != torch.float32:
                raise ValueError("shared parameter buffer does not match this engine")
            self.params = params
        else:
            self.params = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        if grads is not None:       # e.g. a peer-mapped symmetric buffer (exchange.P2PExchange)
            if grads.numel() != self.total or grads.device != self.device or grads.dtype != torch.float32:
                raise ValueError("gradient buffer does not match this engine")
            self.grads = grads
        else:
            self.grads = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        nb = C.c_int64()
        check(self.lib.b200_workspace_bytes(h, C.byref(nb)))
        self.workspace = torch.zeros(nb.value + 256, dtype=torch.uint8, device=self.device)
        pad = (-self.workspace.data_ptr()) % 256
        self._ws_base = self.workspace.data_ptr() + pad
        check(self.lib.b200_bind_params(h, self.params.data_ptr()))
        check(self.lib.b200_bind_grads(h, self.grads.data_ptr()))
        check(self.lib.b200_bind_workspace(h, self._ws_base, nb.value))
        n = C.c_int32()
        check(self.lib.b200_launches_per_step(h, C.byref(n)))
        self.launches_per_step = n.value
        self._loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.lib.b200_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # ---- parameters -----------------------------------------------------------------------
    def _view(self, flat, name):
        i = PARAM_NAMES.index(name)
        return flat[self.offsets[i]:self.offsets[i] + self.sizes[i]].view(self.shapes[name])

    def param(self, name):
        return self._view(self.params, name)

    def grad(self, name):
        return self._view(self.grads, name)

    def load_params(self, params):
        """params: {name: array}, TF layouts (HWIO, [in,out]).  Identical arrays are loaded into
        the oracle in the parity tests (RNG streams are not comparable, cifar10cnn.py:97-101)."""
        for k in PARAM_NAMES:
            v = torch.as_tensor(np.ascontiguousarray(np.asarray(params[k], dtype=np.float32)))
            if tuple(v.shape) != tuple(self.shapes[k]):
                raise ValueError(f"{k}: shape {tuple(v.shape)} != {self.shapes[k]}")
            self.param(k).copy_(v.to(self.device))

    def init_params(self, seed=0):
        """trunc-normal(0.05) weights / 0.1 biases, cifar10cnn.py:97-101, generated on the host."""
        rng = np.random.default_rng(seed)
        out = {}
        for k in PARAM_NAMES:
            shp = self.shapes[k]
            if len(shp) == 1:
                out[k] = np.full(shp, 0.1, np.float32)
            else:
                v = rng.standard_normal(shp)
                bad = np.abs(v) > 2.0
                while bad.any():
                    v[bad] = rng.standard_normal(int(bad.sum()))
                    bad = np.abs(v) > 2.0
                out[k] = (v * 0.05).astype(np.float32)
        self.load_params(out)
        return out

    def get_params(self):
        return {k: self.param(k).detach().cpu().numpy().copy() for k in PARAM_NAMES}

    def get_grads(self):
        return {k: self.grad(k).detach().cpu().numpy().copy() for k in PARAM_NAMES}

    # ---- the step -----------------------------------------------------------------------------
    @staticmethod
    def _stream():
        return torch.cuda.current_stream().cuda_stream

    def _check_inputs(self, images, labels):
        if images.dtype != torch.uint8 or not images.is_cuda or not images.is_contiguous():
            raise ValueError("images must be a contiguous uint8 CUDA tensor [B,3,H,W]")
        if labels.dtype != torch.int32 or not labels.is_cuda or not labels.is_contiguous():
            raise ValueError("labels must be a contiguous int32 CUDA tensor [B] or [B,1]")
        c = self.cfg
        if images.numel() != c.batch * c.c_in * c.img_h * c.img_w or labels.numel() != c.batch:
            raise ValueError("batch shape does not match the engine")

    def forward(self, images, labels):
        self._check_inputs(images, labels)
        check(self.lib.b200_forward(self.h, images.data_ptr(), labels.data_ptr(), self._stream()))

    def forward_backward(self, images, labels):
        """forward + backward in one call (lets independent gradient kernels overlap); same results."""
        self._check_inputs(images, labels)
        check(self.lib.b200_forward_backward(self.h, images.data_ptr(), labels.data_ptr(), self._stream()))

    def backward(self):
        check(self.lib.b200_backward(self.h, self._stream()))

    def sgd_apply(self, lr, grad_scale=1.0):
        check(self.lib.b200_sgd_apply(self.h, lr, grad_scale, self._stream()))

    def train_step(self, images, labels, lr=0.1):
        self._check_inputs(images, labels)
        check(self.lib.b200_train_step(self.h, images.data_ptr(), labels.data_ptr(), lr, self._stream()))

    def train_step_host(self, images_pinned, labels_pinned, lr=0.1, loss_out=None, sync=True):
        """The user-facing call with HOST buffers (pinned uint8 [B,3,H,W], int32 [B]): H2D copy,
        step, D2H of the loss.  Returns the python float loss when sync=True."""
        if images_pinned.is_cuda or labels_pinned.is_cuda or not images_pinned.is_pinned():
            raise ValueError("train_step_host takes pinned host tensors")
        out = self._loss_host if loss_out is None else loss_out
        check(self.lib.b200_train_step_host(self.h, images_pinned.data_ptr(), labels_pinned.data_ptr(), lr,
                                            out.data_ptr(), self._stream()))
        if sync:
            torch.cuda.current_stream().synchronize()
            return float(out[0])
        return None

    def stage_host_batch(self, slot, images_pinned, labels_pinned):
        """copy a pinned host batch into staging slot 0 / 1 (primes train_step_host_prefetch)"""
        if images_pinned.is_cuda or labels_pinned.is_cuda or not images_pinned.is_pinned():
            raise ValueError("stage_host_batch takes pinned host tensors")
        check(self.lib.b200_stage_host_batch(self.h, slot, images_pinned.data_ptr(), labels_pinned.data_ptr(),
                                             self._stream()))

    def train_step_host_prefetch(self, slot, next_images_pinned, next_labels_pinned, lr=0.1, loss_out=None, sync=False):
        """Train on the batch already staged in `slot` while the NEXT pinned host batch is copied into the other
        slot (an input pipeline with prefetch); the loss goes to `loss_out` (pinned)."""
        out = self._loss_host if loss_out is None else loss_out
        ni = None if next_images_pinned is None else next_images_pinned.data_ptr()
        nl = None if next_labels_pinned is None else next_labels_pinned.data_ptr()
        check(self.lib.b200_train_step_host_prefetch(self.h, slot, ni, nl, lr, out.data_ptr(), self._stream()))
        if sync:
            torch.cuda.current_stream().synchronize()
            return float(out[0])
        return None

    def capture(self, fn):
        """Capture the launches `fn()` makes on the current stream into a CUDA graph (after one
        eager warm-up call) and return the graph; graph.replay() re-issues them."""
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g

    def profile_step(self, images, labels, lr=0.1):
        """One eager step with per-site CUDA-event timing -> [(site name, milliseconds)]."""
        self._check_inputs(images, labels)
        ms = (C.c_float * 64)()
        n = C.c_int32()
        check(self.lib.b200_profile_step(self.h, images.data_ptr(), labels.data_ptr(), lr, self._stream(), ms, 64,
                                         C.byref(n)))
        return [(self.lib.b200_site_name(self.h, i).decode(), float(ms[i])) for i in range(n.value)]

    # ---- results / internals ----------------------------------------------------------------------
    def tensor(self, name):
        p = C.c_void_p()
        nb = C.c_int64()
        es = C.c_int32()
        check(self.lib.b200_tensor_ptr(self.h, name.encode(), C.byref(p), C.byref(nb), C.byref(es)))
        off = p.value - self.workspace.data_ptr()
        raw = self.workspace[off:off + nb.value]
        if name in _INT_TENSORS:
            return raw.view(torch.int32)
        return raw.view(_DT[es.value])

    def logits(self):
        return self.tensor("logits").view(self.batch, self.n_cls)

    def loss(self):
        return self.tensor("loss")

    def n_correct(self):
        return self.tensor("n_correct")

    def pred(self):
        return self.tensor("pred")
