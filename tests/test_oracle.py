"""Pins oracle/ref_numpy.py: golden vectors (made by the independent torch-autograd
restatement), finite differences, and hand-computable known-answer tests for the TF op
semantics listed in SURVEY.md App. A.  CPU only."""
import os

import numpy as np
import pytest

from oracle import ref_numpy as rn
from oracle import ref_torch as rt

GOLDEN = ["step_u8_b8.npz", "step_soft_b8.npz"]


def _sample_idx(n, k=256):
    return np.unique(np.linspace(0, n - 1, k).astype(np.int64))


@pytest.mark.parametrize("name", GOLDEN)
def test_numpy_oracle_matches_golden(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    params = rn.init_params(0)
    for k in rn.PARAM_NAMES:      # the weight generator itself is pinned by the fixture
        f = params[k].astype(np.float64).reshape(-1)
        np.testing.assert_allclose([f.sum(), np.abs(f).sum()], z["psum_" + k], rtol=1e-12)
        np.testing.assert_array_equal(f[_sample_idx(f.size)], z["psmp_" + k])
    _, loss, g, a = rn.train_step(params, z["images"], z["labels"])
    assert abs(loss - float(z["loss"])) <= 1e-10 * abs(float(z["loss"]))
    np.testing.assert_allclose(a["logits"], z["logits"], rtol=1e-10, atol=1e-10)
    np.testing.assert_array_equal(a["pred"], z["pred"])            # bit-exact argmax
    for k in rn.PARAM_NAMES:
        f = g[k].reshape(-1)
        scale = np.abs(z["gsmp_" + k]).max() + 1e-300
        np.testing.assert_allclose(f[_sample_idx(f.size)], z["gsmp_" + k], rtol=0, atol=1e-10 * scale)
        np.testing.assert_allclose([f.sum(), np.abs(f).sum()], z["gsum_" + k],
                                   rtol=1e-9, atol=1e-9 * z["gsum_" + k][1])


def test_numpy_vs_torch_full_batch():
    rng = np.random.default_rng(7)
    params = rn.init_params(0)
    img = rng.integers(0, 256, (32, 3, 32, 32), dtype=np.uint8)
    lab = rng.integers(0, 10, (32, 1)).astype(np.int32)
    _, loss, g, a = rn.train_step(params, img, lab)
    l2, logits2, g2 = rt.loss_and_grads(params, img, lab)
    assert abs(loss - l2) <= 1e-10 * abs(l2)
    np.testing.assert_allclose(a["logits"], logits2, rtol=1e-10, atol=1e-10)
    for k in rn.PARAM_NAMES:
        assert np.abs(g[k] - g2[k]).max() <= 1e-10 * np.abs(g2[k]).max()


def test_finite_differences():
    """Central differences on the fp64 oracle, a few entries of every tensor."""
    rng = np.random.default_rng(3)
    params = {k: v.astype(np.float64) for k, v in rn.init_params(0).items()}
    x = rng.normal(0, 1, (3, 24, 24, 3))
    lab = rng.integers(0, 10, (3,))
    a = rn.forward(params, x, lab)
    g = rn.backward(params, a)
    for k in rn.PARAM_NAMES:
        flat = params[k].reshape(-1)
        # probe entries with a non-negligible gradient so the relative check has meaning
        order = np.argsort(-np.abs(g[k].reshape(-1)))[:3]
        for i in order:
            old = flat[i]
            h = 1e-6
            flat[i] = old + h
            lp = rn.forward(params, x, lab)["loss"]
            flat[i] = old - h
            lm = rn.forward(params, x, lab)["loss"]
            flat[i] = old
            fd = (lp - lm) / (2 * h)
            assert abs(fd - g[k].reshape(-1)[i]) <= 1e-5 * max(abs(fd), 1e-3), (k, i, fd, g[k].reshape(-1)[i])


def test_crop_is_centre_and_hwc():
    img = np.arange(3 * 32 * 32, dtype=np.uint32).reshape(1, 3, 32, 32)
    img8 = (img % 251).astype(np.uint8)
    out = rn.crop_cast(img8)
    assert out.shape == (1, 24, 24, 3)
    for c in range(3):
        np.testing.assert_array_equal(out[0, :, :, c], img8[0, c, 4:28, 4:28])


def test_pool_same_padding_is_after_only():
    """4x4 input -> 2x2; the second window in each axis sees rows/cols 2,3 and one padded
    cell (never wins).  TF pads 0 before / 1 after for even sizes."""
    x = np.arange(16, dtype=np.float64).reshape(1, 4, 4, 1)
    y, arg = rn.maxpool_3x3s2_same(x)
    np.testing.assert_array_equal(y[0, :, :, 0], [[10, 11], [14, 15]])
    np.testing.assert_array_equal(arg[0, :, :, 0], [[8, 7], [5, 4]])   # kh*3+kw
    xn = -np.ones((1, 4, 4, 1))                      # all-negative: padding must not yield 0 / -inf
    y, _ = rn.maxpool_3x3s2_same(xn)
    assert (y == -1).all()


def test_pool_grad_first_max_wins_and_overlaps_accumulate():
    x = np.zeros((1, 4, 4, 1))
    x[0, 2, 2, 0] = 5.0                               # cell shared by all four windows
    y, arg = rn.maxpool_3x3s2_same(x)
    assert (y == 5).all()
    dy = np.array([1.0, 10.0, 100.0, 1000.0]).reshape(1, 2, 2, 1)
    dx = rn.maxpool_3x3s2_same_grad(dy, arg, (4, 4))
    assert dx[0, 2, 2, 0] == 1111.0 and dx.sum() == 1111.0
    # ties: constant input -> the FIRST cell in (h, w) scan order of each window takes it all
    y, arg = rn.maxpool_3x3s2_same(np.ones((1, 4, 4, 1)))
    assert (arg == 0).all()
    dx = rn.maxpool_3x3s2_same_grad(dy, arg, (4, 4))
    assert dx[0, 0, 0, 0] == 1 and dx[0, 0, 2, 0] == 10 and dx[0, 2, 0, 0] == 100 and dx[0, 2, 2, 0] == 1000


def test_conv_delta_filter_and_padding():
    x = np.random.default_rng(0).normal(size=(2, 6, 6, 3))
    w = np.zeros((5, 5, 3, 4))
    w[2, 2, 1, 0] = 1.0                               # centre tap -> copy of channel 1
    w[0, 0, 0, 1] = 1.0                               # top-left tap -> shift by (+2,+2), zeros enter
    y = rn.conv2d_same(x, w)
    np.testing.assert_array_equal(y[..., 0], x[..., 1])
    np.testing.assert_array_equal(y[:, 2:, 2:, 1], x[:, :4, :4, 0])
    assert (y[:, :2, :, 1] == 0).all() and (y[:, :, :2, 1] == 0).all()


def test_xent_argmax_ties_and_all_zero_row():
    logits = np.zeros((2, 10))
    logits[1, 3] = logits[1, 7] = 2.0                 # tie -> first index (cifar10cnn.py:173)
    loss, rows, back = rn.softmax_xent(logits, [4, 7])
    assert abs(rows[0] - np.log(10)) < 1e-15
    np.testing.assert_allclose(back[0], np.full(10, 0.1) - np.eye(10)[4])
    acc, pred, n = rn.batch_accuracy(logits, [[0], [7]])
    np.testing.assert_array_equal(pred, [0, 3])
    assert n == 1 and acc == 0.5
    with pytest.raises(AssertionError):
        rn.softmax_xent(logits, [10, 0])


def test_relu_grad_is_zero_at_zero():
    params = {k: np.zeros(s) for k, s in rn.param_shapes().items()}   # all outputs exactly 0
    x = np.ones((2, 24, 24, 3))
    a = rn.forward(params, x, [1, 2])
    g = rn.backward(params, a)
    assert all((v == 0).all() for v in g.values())
    assert abs(a["loss"] - np.log(10)) < 1e-15


def test_learning_rate_staircase():
    assert rn.learning_rate(0) == np.float32(0.1)
    assert rn.learning_rate(249) == np.float32(0.1)
    np.testing.assert_allclose(rn.learning_rate(250), 0.09, rtol=1e-6)
    np.testing.assert_allclose(rn.learning_rate(750), 0.1 * 0.9 ** 3, rtol=1e-6)


def test_param_count_and_init():
    p = rn.init_params(0)
    assert sum(v.size for v in p.values()) == 1068298
    assert np.abs(p["full_weight_1"]).max() <= 0.1 + 1e-7
    assert (p["conv2_bias"] == np.float32(0.1)).all()
    p32 = rn.init_params(0, crop_h=32, crop_w=32)
    assert p32["full_weight_1"].shape == (4096, 384)


def test_allreduce_sum_equals_sequential_same_snapshot():
    rng = np.random.default_rng(5)
    params = {k: v.astype(np.float64) for k, v in rn.init_params(0).items()}
    gs = []
    for r in range(2):
        img = np.clip(np.rint(rng.normal(8, 2.5, (4, 3, 32, 32))), 0, 255).astype(np.uint8)
        lab = rng.integers(0, 10, (4,))
        gs.append(rn.train_step(params, img, lab)[2])
    both = rn.allreduce_step(params, gs, 0.1, "sum")
    seq = rn.sgd_apply(rn.sgd_apply(params, gs[0], 0.1), gs[1], 0.1)
    for k in params:
        np.testing.assert_allclose(both[k], seq[k], rtol=1e-13, atol=1e-15)
