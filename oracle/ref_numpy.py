"""CPU oracle (TEST INFRASTRUCTURE ONLY) for the cifar10cnn.py training step.

This file is a plain numpy restatement of the arithmetic the reference script asks
TensorFlow 1.x to do.  It is imported only by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py -- never by the product path, which is
CUDA-only and raises when its extension is missing.

PARITY UNPINNED: the reference (/root/reference/cifar10cnn.py, 274 lines + a README)
has no tests, seeds, golden vectors or fixtures, and TensorFlow 1.x (un-pinned, not
vendored; `tf.contrib` at cifar10cnn.py:29 proves 1.x) cannot be installed on
Python 3.12, so nothing published pins these numbers.  What pins them instead:
  * oracle/ref_torch.py, an independent autograd restatement (agreement <= 1e-10 in fp64),
  * central finite differences (tests/test_oracle.py),
  * hand-computable known-answer tests for the TF op semantics that libraries disagree on
    (asymmetric SAME pooling pad, first-max tie rule, ReLU-grad at 0, first-index argmax).

Every function cites the reference line it restates.  Layouts are TensorFlow's:
activations NHWC, conv filters HWIO, FC weights [in, out].  Gradients are derived by
hand (no autograd) so that the cross-check against ref_torch.py is independent.
"""
from __future__ import annotations

import numpy as np
from numpy.lib.stride_tricks import as_strided

# ---- constants, cifar10cnn.py:9-27 -------------------------------------------------------
BATCH_SIZE = 128          # :13
GENERATIONS = 20000       # :14
IMAGE_HEIGHT = 32         # :15
IMAGE_WIDTH = 32          # :16
CROP_HEIGHT = 24          # :17
CROP_WIDTH = 24           # :18
NUM_CHANNELS = 3          # :19
NUM_TARGETS = 10          # :20
LEARNING_RATE = 0.1       # :21
LR_DECAY = 0.9            # :22
NUM_GENS_TO_WAIT = 250.0  # :23
RECORD_LENGTH = IMAGE_HEIGHT * IMAGE_WIDTH * NUM_CHANNELS + 1  # :24-25

# Variable creation order inside create_cnn (cifar10cnn.py:105-144); the product keeps it.
PARAM_NAMES = (
    "conv1_kernel", "conv1_bias", "conv2_kernel", "conv2_bias",
    "full_weight_1", "full_bias_1", "full_weight_2", "full_bias_2",
    "full_weight_3", "full_bias_3",
)


def param_shapes(crop_h=CROP_HEIGHT, crop_w=CROP_WIDTH, c1=64, c2=64, f1=384, f2=192,
                 n_cls=NUM_TARGETS):
    """Shapes of the ten variables (cifar10cnn.py:106-144).  `reshaped_dim` (:127) is
    ceil(ceil(H/2)/2) * ceil(ceil(W/2)/2) * c2 because both pools are SAME / stride 2."""
    h2 = -(-(-(-crop_h // 2)) // 2)
    w2 = -(-(-(-crop_w // 2)) // 2)
    flat = h2 * w2 * c2
    return {
        "conv1_kernel": (5, 5, NUM_CHANNELS, c1), "conv1_bias": (c1,),
        "conv2_kernel": (5, 5, c1, c2), "conv2_bias": (c2,),
        "full_weight_1": (flat, f1), "full_bias_1": (f1,),
        "full_weight_2": (f1, f2), "full_bias_2": (f2,),
        "full_weight_3": (f2, n_cls), "full_bias_3": (n_cls,),
    }


def init_params(seed=0, dtype=np.float32, **shape_kw):
    """cifar10cnn.py:97-101: weights truncated-normal(stddev 0.05) -- samples beyond two
    standard deviations are re-drawn -- and biases constant 0.1.  TF's RNG stream is not
    reproducible outside TF, so parity tests load these arrays into both sides."""
    rng = np.random.default_rng(seed)
    out = {}
    for name, shp in param_shapes(**shape_kw).items():
        if len(shp) == 1:
            out[name] = np.full(shp, 0.1, dtype=dtype)
        else:
            v = rng.standard_normal(shp)
            bad = np.abs(v) > 2.0
            while bad.any():
                v[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(v) > 2.0
            out[name] = (v * 0.05).astype(dtype)
    return out


# ---- input path, cifar10cnn.py:57-68 --------------------------------------------------------
def crop_cast(images_u8_chw, crop_h=CROP_HEIGHT, crop_w=CROP_WIDTH, dtype=np.float64):
    """[B,3,32,32] uint8 CHW record payload -> [B,crop_h,crop_w,3] float HWC.
    :61-62 reshape to CHW, :65 transpose to HWC, :66 cast, :68 resize_image_with_crop_or_pad,
    which is a deterministic CENTRE crop: offset (32-24)//2 = 4 on both axes."""
    x = np.asarray(images_u8_chw)
    assert x.dtype == np.uint8 and x.ndim == 4 and x.shape[1] == NUM_CHANNELS
    oh = (x.shape[2] - crop_h) // 2
    ow = (x.shape[3] - crop_w) // 2
    return np.ascontiguousarray(
        x[:, :, oh:oh + crop_h, ow:ow + crop_w].transpose(0, 2, 3, 1)).astype(dtype)


# ---- conv2d SAME 5x5 stride 1, cifar10cnn.py:107 and :118 -------------------------------------
def _im2col5(x):
    """[B,H,W,C] -> [B*H*W, 25*C] patches with 2 zeros of padding on each side
    (SAME with a 5x5 window and stride 1: total pad 4, 2 before / 2 after)."""
    B, H, W, C = x.shape
    xp = np.zeros((B, H + 4, W + 4, C), dtype=x.dtype)
    xp[:, 2:2 + H, 2:2 + W, :] = x
    s = xp.strides
    win = as_strided(xp, shape=(B, H, W, 5, 5, C), strides=(s[0], s[1], s[2], s[1], s[2], s[3]),
                     writeable=False)
    return win.reshape(B * H * W, 25 * C)


def conv2d_same(x, w):
    """y[n,i,j,k] = sum_{di,dj,c} x[n,i+di-2,j+dj-2,c] * w[di,dj,c,k]  (cross-correlation)."""
    B, H, W, C = x.shape
    K = w.shape[3]
    return (_im2col5(x) @ w.reshape(25 * C, K)).reshape(B, H, W, K)


def conv2d_same_grad_filter(x, dy):
    """Conv2DBackpropFilter: dw[di,dj,c,k] = sum_{n,i,j} x[n,i+di-2,j+dj-2,c] * dy[n,i,j,k]."""
    B, H, W, C = x.shape
    K = dy.shape[3]
    return (_im2col5(x).T @ dy.reshape(B * H * W, K)).reshape(5, 5, C, K)


def conv2d_same_grad_input(dy, w):
    """Conv2DBackpropInput: dx[n,p,q,c] = sum_{di,dj,k} dy[n,p-di+2,q-dj+2,k] * w[di,dj,c,k],
    i.e. a SAME correlation of dy with the filter rotated 180 degrees and in/out swapped."""
    wr = np.ascontiguousarray(w[::-1, ::-1].transpose(0, 1, 3, 2))
    return conv2d_same(dy, wr)


# ---- max_pool 3x3 stride 2 SAME, cifar10cnn.py:113 and :123 -----------------------------------
def maxpool_3x3s2_same(x):
    """out = ceil(in/2); pad_total = max((out-1)*2+3-in, 0), pad_before = pad_total//2 (0 for
    even sizes), the rest after.  Padded cells never win (-inf).  Returns (y, arg) where
    arg in 0..8 is kh*3+kw of the FIRST maximal cell in (h, w) scan order -- the rule of
    TF's CPU MaxPoolGrad (strict '<' update)."""
    B, H, W, C = x.shape
    OH, OW = -(-H // 2), -(-W // 2)
    pth = max((OH - 1) * 2 + 3 - H, 0)
    ptw = max((OW - 1) * 2 + 3 - W, 0)
    pbh, pbw = pth // 2, ptw // 2
    xp = np.full((B, H + pth, W + ptw, C), -np.inf, dtype=x.dtype)
    xp[:, pbh:pbh + H, pbw:pbw + W, :] = x
    s = xp.strides
    win = as_strided(xp, shape=(B, OH, OW, 3, 3, C),
                     strides=(s[0], 2 * s[1], 2 * s[2], s[1], s[2], s[3]), writeable=False)
    flat = win.reshape(B, OH, OW, 9, C)
    arg = flat.argmax(axis=3)                      # numpy argmax = first maximal element
    y = np.take_along_axis(flat, arg[:, :, :, None, :], axis=3)[:, :, :, 0, :]
    return y, arg.astype(np.uint8)


def maxpool_3x3s2_same_grad(dy, arg, in_hw):
    """MaxPoolGrad: every window sends its gradient to its recorded argmax cell; windows
    overlap (stride 2 < 3) so cells chosen by several windows accumulate."""
    B, OH, OW, C = dy.shape
    H, W = in_hw
    pth = max((OH - 1) * 2 + 3 - H, 0)
    ptw = max((OW - 1) * 2 + 3 - W, 0)
    pbh, pbw = pth // 2, ptw // 2
    dxp = np.zeros((B, H + pth, W + ptw, C), dtype=dy.dtype)
    for k in range(9):
        kh, kw = divmod(k, 3)
        dxp[:, kh:kh + 2 * OH:2, kw:kw + 2 * OW:2, :] += dy * (arg == k)
    return dxp[:, pbh:pbh + H, pbw:pbw + W, :]


# ---- loss / accuracy / learning rate ----------------------------------------------------------
def softmax_xent(logits, labels):
    """sparse_softmax_cross_entropy_with_logits (cifar10cnn.py:154) + reduce_mean (:156).
    Returns (mean loss, per-row loss, backprop = softmax - onehot)."""
    labels = np.asarray(labels).reshape(-1).astype(np.int64)   # :152 squeeze + cast
    B = logits.shape[0]
    assert ((labels >= 0) & (labels < logits.shape[1])).all(), "label out of range"
    m = logits.max(axis=1, keepdims=True)
    e = np.exp(logits - m)
    s = e.sum(axis=1, keepdims=True)
    rows = np.log(s[:, 0]) - (logits[np.arange(B), labels] - m[:, 0])
    back = e / s
    back[np.arange(B), labels] -= 1.0
    return rows.mean(), rows, back


def batch_accuracy(logits, labels):
    """cifar10cnn.py:166-176.  argmax = first maximal index (:173); returns
    (accuracy, predictions int32, number correct)."""
    labels = np.asarray(labels).reshape(-1).astype(np.int32)
    pred = logits.argmax(axis=1).astype(np.int32)
    ok = pred == labels
    return ok.astype(np.float32).mean(), pred, int(ok.sum())


def learning_rate(generation_num=0, base=LEARNING_RATE, decay=LR_DECAY, steps=NUM_GENS_TO_WAIT):
    """exponential_decay(..., staircase=True), cifar10cnn.py:161, evaluated in fp32 like TF.
    generation_num is never incremented in the reference (:216), so this is 0.1 in practice."""
    p = np.floor(np.float32(generation_num) / np.float32(steps))
    return np.float32(base) * np.power(np.float32(decay), p, dtype=np.float32)


# ---- the model, cifar10cnn.py:94-147 ----------------------------------------------------------
def forward(params, x, labels=None):
    """create_cnn on x = [B,H,W,3] float (already cropped).  Keeps every layer-boundary
    tensor so backward() and the per-kernel parity tests can use them."""
    dt = x.dtype
    p = {k: np.asarray(v, dtype=dt) for k, v in params.items()}
    a = {"x": x}
    a["conv1"] = np.maximum(conv2d_same(x, p["conv1_kernel"]) + p["conv1_bias"], 0)      # :107-110
    a["pool1"], a["arg1"] = maxpool_3x3s2_same(a["conv1"])                                 # :113
    a["conv2"] = np.maximum(conv2d_same(a["pool1"], p["conv2_kernel"]) + p["conv2_bias"], 0)  # :118-121
    a["pool2"], a["arg2"] = maxpool_3x3s2_same(a["conv2"])                                 # :123
    a["flat"] = a["pool2"].reshape(x.shape[0], -1)                                         # :126 (h,w,c) order
    a["full1"] = np.maximum(a["flat"] @ p["full_weight_1"] + p["full_bias_1"], 0)          # :133
    a["full2"] = np.maximum(a["full1"] @ p["full_weight_2"] + p["full_bias_2"], 0)         # :139
    a["logits"] = np.maximum(a["full2"] @ p["full_weight_3"] + p["full_bias_3"], 0)        # :145 (ReLU on logits)
    if labels is not None:
        a["loss"], a["loss_rows"], a["xent_back"] = softmax_xent(a["logits"], labels)       # :150-157
        a["accuracy"], a["pred"], a["n_correct"] = batch_accuracy(a["logits"], labels)     # :166-176
    return a


def backward(params, a):
    """The gradient graph `minimize` builds (cifar10cnn.py:162-163), hand-derived.
    ReluGrad uses the OUTPUT: dx = dy * [y > 0] (zero gradient at exactly 0)."""
    dt = a["x"].dtype
    p = {k: np.asarray(v, dtype=dt) for k, v in params.items()}
    B = a["x"].shape[0]
    g = {}
    d = a["xent_back"] / B                                   # reduce_mean upstream 1/B
    d = d * (a["logits"] > 0)                                # ReLU on logits (:145)
    a_d = {"dlogits": d}
    g["full_weight_3"] = a["full2"].T @ d
    g["full_bias_3"] = d.sum(axis=0)
    d = (d @ p["full_weight_3"].T) * (a["full2"] > 0)
    a_d["dfull2"] = d
    g["full_weight_2"] = a["full1"].T @ d
    g["full_bias_2"] = d.sum(axis=0)
    d = (d @ p["full_weight_2"].T) * (a["full1"] > 0)
    a_d["dfull1"] = d
    g["full_weight_1"] = a["flat"].T @ d
    g["full_bias_1"] = d.sum(axis=0)
    d = (d @ p["full_weight_1"].T).reshape(a["pool2"].shape)
    a_d["dpool2"] = d
    d = maxpool_3x3s2_same_grad(d, a["arg2"], a["conv2"].shape[1:3]) * (a["conv2"] > 0)
    a_d["dconv2"] = d
    g["conv2_kernel"] = conv2d_same_grad_filter(a["pool1"], d)
    g["conv2_bias"] = d.sum(axis=(0, 1, 2))
    d = conv2d_same_grad_input(d, p["conv2_kernel"])
    a_d["dpool1"] = d
    d = maxpool_3x3s2_same_grad(d, a["arg1"], a["conv1"].shape[1:3]) * (a["conv1"] > 0)
    a_d["dconv1"] = d
    g["conv1_kernel"] = conv2d_same_grad_filter(a["x"], d)    # no input gradient: x is data
    g["conv1_bias"] = d.sum(axis=(0, 1, 2))
    a.update(a_d)
    return g


def sgd_apply(params, grads, lr):
    """ApplyGradientDescent x10 (cifar10cnn.py:162-163): var <- var - lr * grad, in the
    variable's own dtype (fp32 in the reference)."""
    return {k: (params[k] - np.asarray(lr, dtype=params[k].dtype) * grads[k].astype(params[k].dtype))
            for k in params}


def train_step(params, images_u8_chw, labels, lr=None, dtype=np.float64, crop=(CROP_HEIGHT, CROP_WIDTH)):
    """One `mon_sess.run([train_op, loss])` (cifar10cnn.py:230) for one worker.
    Returns (new_params, loss, grads, activations)."""
    x = crop_cast(images_u8_chw, crop[0], crop[1], dtype=dtype)
    a = forward(params, x, labels)
    g = backward(params, a)
    lr = learning_rate(0) if lr is None else lr
    return sgd_apply(params, g, lr), a["loss"], g, a


def allreduce_step(params, per_rank_grads, lr, mode="sum"):
    """Synchronous stand-in for N asynchronous worker pushes computed from one parameter
    snapshot (cifar10cnn.py:162-163 applied on the ps, :195-196): the N mean-gradients are
    summed (default) or averaged and applied once.  global_step advances by N."""
    n = len(per_rank_grads)
    tot = {k: sum(g[k] for g in per_rank_grads) for k in params}
    if mode == "mean":
        tot = {k: v / n for k, v in tot.items()}
    return sgd_apply(params, tot, lr)
