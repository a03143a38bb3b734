"""GPU parity tests: every kernel is called through the C ABI (include/b200cnn.h) and compared
with the CPU oracle (oracle/ref_numpy.py, fp64) on the same seeded inputs.

Tolerances (fp32 validation mode): relative error normalised by the tensor's max-abs
<= 1e-5 (accumulation-order noise of fp32 sums up to 73,728 terms long); integer / index results
(crop bytes, pool argmax, predictions, n_correct) bit-exact.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import b200cnn
from b200cnn import _lib
from oracle import ref_numpy as rn

pytestmark = pytest.mark.gpu

RTOL_F32 = 1e-5


def dev(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def relerr(got, want):
    want = np.asarray(want, dtype=np.float64)
    got = np.asarray(got, dtype=np.float64)
    return np.abs(got - want).max() / (np.abs(want).max() + 1e-30)


def stream():
    return torch.cuda.current_stream().cuda_stream


@pytest.fixture(scope="module")
def lib():
    return _lib.load()


def ck(rc):
    _lib.check(rc)


# ---------------------------------------------------------------------------------------------
def test_crop_cast_bit_exact(lib):
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (5, 3, 32, 32), dtype=np.uint8)
    out = torch.empty(5, 24, 24, 3, device="cuda")
    ck(lib.b200_crop_cast_f32(dev(img).data_ptr(), out.data_ptr(), 5, 32, 32, 24, 24, 3, stream()))
    np.testing.assert_array_equal(out.cpu().numpy(), rn.crop_cast(img, dtype=np.float32))


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(3, 24, 24, 3, 64), (3, 12, 12, 64, 64), (2, 7, 9, 5, 12)])
def test_conv_fwd_dgrad_wgrad(lib, B, H, W, Cin, Cout):
    rng = np.random.default_rng(1)
    x = rng.normal(size=(B, H, W, Cin)).astype(np.float32)
    w = (rng.normal(size=(5, 5, Cin, Cout)) * 0.05).astype(np.float32)
    b = rng.normal(size=(Cout,)).astype(np.float32)
    dy = rng.normal(size=(B, H, W, Cout)).astype(np.float32)
    xd, wd_, bd, dyd = dev(x), dev(w), dev(b), dev(dy)
    y = torch.empty(B, H, W, Cout, device="cuda")
    ck(lib.b200_conv5x5_fwd_f32(xd.data_ptr(), wd_.data_ptr(), bd.data_ptr(), y.data_ptr(), B, H, W, Cin, Cout, 1,
                                stream()))
    want = np.maximum(rn.conv2d_same(x.astype(np.float64), w.astype(np.float64)) + b, 0)
    assert relerr(y.cpu().numpy(), want) <= RTOL_F32
    # no bias / no relu variant
    ck(lib.b200_conv5x5_fwd_f32(xd.data_ptr(), wd_.data_ptr(), None, y.data_ptr(), B, H, W, Cin, Cout, 0, stream()))
    assert relerr(y.cpu().numpy(), rn.conv2d_same(x.astype(np.float64), w.astype(np.float64))) <= RTOL_F32
    # dgrad
    dx = torch.empty(B, H, W, Cin, device="cuda")
    wscr = torch.empty(25 * Cin * Cout, device="cuda")
    ck(lib.b200_conv5x5_dgrad_f32(dyd.data_ptr(), wd_.data_ptr(), wscr.data_ptr(), dx.data_ptr(), B, H, W, Cin, Cout,
                                  stream()))
    assert relerr(dx.cpu().numpy(), rn.conv2d_same_grad_input(dy.astype(np.float64), w.astype(np.float64))) <= RTOL_F32
    # wgrad + bias grad
    nb = C.c_int64()
    ck(lib.b200_conv5x5_wgrad_scratch_bytes(B, H, W, Cin, Cout, C.byref(nb)))
    scr = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    dw = torch.empty(5, 5, Cin, Cout, device="cuda")
    db = torch.empty(Cout, device="cuda")
    ck(lib.b200_conv5x5_wgrad_f32(xd.data_ptr(), dyd.data_ptr(), dw.data_ptr(), db.data_ptr(), scr.data_ptr(), nb.value,
                                  B, H, W, Cin, Cout, stream()))
    assert relerr(dw.cpu().numpy(), rn.conv2d_same_grad_filter(x.astype(np.float64), dy.astype(np.float64))) <= RTOL_F32
    assert relerr(db.cpu().numpy(), dy.astype(np.float64).sum((0, 1, 2))) <= RTOL_F32
    # too-small scratch is rejected, not overrun
    assert lib.b200_conv5x5_wgrad_f32(xd.data_ptr(), dyd.data_ptr(), dw.data_ptr(), db.data_ptr(), scr.data_ptr(), 16,
                                      B, H, W, Cin, Cout, stream()) == _lib.E_INVALID


@pytest.mark.parametrize("B,H,W,Cc", [(3, 24, 24, 64), (2, 12, 12, 64), (2, 5, 7, 3), (1, 4, 4, 1)])
def test_maxpool_fwd_bwd_exact(lib, B, H, W, Cc):
    rng = np.random.default_rng(2)
    x = np.maximum(rng.normal(size=(B, H, W, Cc)), 0).astype(np.float32)      # post-ReLU: many exact-0 ties
    x[0, : min(H, 3), : min(W, 3), 0] = 7.0                                      # planted positive ties
    OH, OW = -(-H // 2), -(-W // 2)
    y = torch.empty(B, OH, OW, Cc, device="cuda")
    arg = torch.empty(B, OH, OW, Cc, dtype=torch.uint8, device="cuda")
    xd = dev(x)
    ck(lib.b200_maxpool_fwd_f32(xd.data_ptr(), y.data_ptr(), arg.data_ptr(), B, H, W, Cc, stream()))
    wy, warg = rn.maxpool_3x3s2_same(x)
    np.testing.assert_array_equal(y.cpu().numpy(), wy)
    np.testing.assert_array_equal(arg.cpu().numpy(), warg)
    dy = rng.normal(size=(B, OH, OW, Cc)).astype(np.float32)
    dx = torch.empty(B, H, W, Cc, device="cuda")
    ck(lib.b200_maxpool_bwd_f32(dev(dy).data_ptr(), arg.data_ptr(), xd.data_ptr(), dx.data_ptr(), B, H, W, Cc, 1,
                                stream()))
    want = rn.maxpool_3x3s2_same_grad(dy.astype(np.float64), warg, (H, W)) * (x > 0)
    np.testing.assert_allclose(dx.cpu().numpy(), want, rtol=0, atol=2e-6)
    ck(lib.b200_maxpool_bwd_f32(dev(dy).data_ptr(), arg.data_ptr(), None, dx.data_ptr(), B, H, W, Cc, 0, stream()))
    want = rn.maxpool_3x3s2_same_grad(dy.astype(np.float64), warg, (H, W))
    np.testing.assert_allclose(dx.cpu().numpy(), want, rtol=0, atol=2e-6)
    assert abs(float(dx.sum()) - dy.astype(np.float64).sum()) < 1e-3     # every window's gradient lands once


@pytest.mark.parametrize("B,K,N", [(128, 2304, 384), (128, 384, 192), (128, 192, 10), (5, 33, 7)])
def test_fc_fwd_bwd(lib, B, K, N):
    rng = np.random.default_rng(3)
    x = np.maximum(rng.normal(size=(B, K)), 0).astype(np.float32)
    w = (rng.normal(size=(K, N)) * 0.05).astype(np.float32)
    b = rng.normal(size=(N,)).astype(np.float32)
    dy = rng.normal(size=(B, N)).astype(np.float32)
    xd, wd_, bd, dyd = dev(x), dev(w), dev(b), dev(dy)
    y = torch.empty(B, N, device="cuda")
    ck(lib.b200_fc_fwd_f32(xd.data_ptr(), wd_.data_ptr(), bd.data_ptr(), y.data_ptr(), B, K, N, 1, stream()))
    x64, w64, dy64 = x.astype(np.float64), w.astype(np.float64), dy.astype(np.float64)
    assert relerr(y.cpu().numpy(), np.maximum(x64 @ w64 + b, 0)) <= RTOL_F32
    nb = C.c_int64()
    ck(lib.b200_colsum_scratch_bytes(B, N, C.byref(nb)))
    scr = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    dx, dw, db = torch.empty(B, K, device="cuda"), torch.empty(K, N, device="cuda"), torch.empty(N, device="cuda")
    for mask in (1, 0):
        ck(lib.b200_fc_bwd_f32(xd.data_ptr(), wd_.data_ptr(), dyd.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                               db.data_ptr(), scr.data_ptr(), nb.value, B, K, N, mask, stream()))
        wdx = dy64 @ w64.T
        if mask:
            wdx = wdx * (x > 0)
        assert relerr(dx.cpu().numpy(), wdx) <= RTOL_F32
        assert relerr(dw.cpu().numpy(), x64.T @ dy64) <= RTOL_F32
        assert relerr(db.cpu().numpy(), dy64.sum(0)) <= RTOL_F32


def test_head_loss_argmax_dlogits(lib):
    rng = np.random.default_rng(4)
    B = 300
    logits = np.maximum(rng.normal(size=(B, 10)) * 3, 0).astype(np.float32)   # ReLU'd logits: exact zeros, ties
    logits[0] = 0.0                                                            # all-zero row -> argmax 0, loss log 10
    logits[1, 3] = logits[1, 7] = 9.0                                          # tie -> first index
    labels = rng.integers(0, 10, (B,)).astype(np.int32)
    labels[1] = 7
    ld, lab = dev(logits), dev(labels)
    loss = torch.zeros(1, device="cuda")
    dl = torch.empty(B, 10, device="cuda")
    pred = torch.empty(B, dtype=torch.int32, device="cuda")
    nc = torch.zeros(1, dtype=torch.int32, device="cuda")
    nbad = torch.zeros(1, dtype=torch.int32, device="cuda")
    ck(lib.b200_head_f32(ld.data_ptr(), lab.data_ptr(), loss.data_ptr(), dl.data_ptr(), pred.data_ptr(),
                         nc.data_ptr(), nbad.data_ptr(), B, 10, stream()))
    wl, _, back = rn.softmax_xent(logits.astype(np.float64), labels)
    _, wpred, wn = rn.batch_accuracy(logits, labels)
    assert abs(float(loss) - wl) <= 2e-6 * abs(wl)
    np.testing.assert_array_equal(pred.cpu().numpy(), wpred)                   # bit-exact argmax, first index on ties
    assert int(nc) == wn and int(nbad) == 0
    assert pred[0] == 0 and pred[1] == 3
    want = back / B * (logits > 0)
    np.testing.assert_allclose(dl.cpu().numpy(), want, rtol=0, atol=1e-6 * np.abs(want).max())
    # out-of-range labels are counted and contribute nothing (TF raises; we flag)
    labels[5], labels[6] = 10, -1
    ck(lib.b200_head_f32(ld.data_ptr(), dev(labels).data_ptr(), loss.data_ptr(), dl.data_ptr(), pred.data_ptr(),
                         nc.data_ptr(), nbad.data_ptr(), B, 10, stream()))
    assert int(nbad) == 2 and float(dl[5].abs().sum()) == 0.0
    assert lib.b200_head_f32(ld.data_ptr(), lab.data_ptr(), loss.data_ptr(), dl.data_ptr(), pred.data_ptr(),
                             nc.data_ptr(), nbad.data_ptr(), B, 17, stream()) == _lib.E_INVALID


def test_sgd_apply(lib):
    rng = np.random.default_rng(5)
    for n in (1068352, 1027, 3):
        p = rng.normal(size=n).astype(np.float32)
        g = rng.normal(size=n).astype(np.float32)
        pd = dev(p)
        ck(lib.b200_sgd_apply_f32(pd.data_ptr(), dev(g).data_ptr(), n, 0.1, 1.0, stream()))
        want = p.astype(np.float64) - np.float64(np.float32(0.1)) * g
        np.testing.assert_allclose(pd.cpu().numpy(), want, rtol=0, atol=1.5e-7 * np.abs(want).max())


# ---------------------------------------------------------------------------------------------
def _engine(batch, mode="fp32", **kw):
    e = b200cnn.TrainStepEngine(batch=batch, mode=mode, **kw)
    return e


def _oracle_backward_on_gpu_forward(e, params, labels, a_ref, fused_conv1=False):
    """Backward parity is judged given the GPU's OWN forward values: max-pool argmax and ReLU masks
    are discontinuous, and a near-tie that fp32 resolves differently from fp64 (expected a few
    times per batch: ~1e5 pooling windows, values O(100), fp32 spacing 1e-5) legitimately moves a
    gradient to a neighbouring cell.  So: (1) forward tensors match the oracle to tolerance,
    (2) the GPU's argmax is exactly the first-max of the GPU's own conv output (checked here,
    bit-exact), (3) gradients match the oracle's backward evaluated on the GPU's forward tensors."""
    a = {"x": a_ref["x"]}
    names = ("pool1", "pool2", "full1", "full2", "logits") if fused_conv1 else \
        ("conv1", "pool1", "conv2", "pool2", "full1", "full2", "logits")
    for name in names:
        a[name] = e.tensor(name).cpu().numpy().astype(np.float64).reshape(a_ref[name].shape)
    if fused_conv1:
        # conv1 never reaches HBM in the fused kernel (its own test checks the pool rule on a debug
        # copy); backward only needs its sign at the selected cells, which is the sign of pool1
        # (the same holds for conv2, fused with pool2)
        for k in ("1", "2"):
            a["arg" + k] = e.tensor("arg" + k).cpu().numpy().reshape(a_ref["arg" + k].shape)
            a["conv" + k] = rn.maxpool_3x3s2_same_grad(a["pool" + k], a["arg" + k], a_ref["conv" + k].shape[1:3])
    for name, src in () if fused_conv1 else (("arg1", "conv1"), ("arg2", "conv2")):
        a[name] = e.tensor(name).cpu().numpy().reshape(a_ref[name].shape)
        y, arg = rn.maxpool_3x3s2_same(a[src])
        np.testing.assert_array_equal(a[name], arg)                   # first-max rule on its own values
        np.testing.assert_array_equal(a["pool" + name[-1]], y)
    a["flat"] = a["pool2"].reshape(a["pool2"].shape[0], -1)
    a["loss"], a["loss_rows"], a["xent_back"] = rn.softmax_xent(a["logits"], labels)
    g = rn.backward({k: np.asarray(v, np.float64) for k, v in params.items()}, a)
    return g, a


def _grad_check(e, g_oracle, tol):
    got = e.get_grads()
    for k in rn.PARAM_NAMES:
        err = relerr(got[k], g_oracle[k])
        assert err <= tol, (k, err)


@pytest.mark.parametrize("name", ["step_u8_b8.npz", "step_soft_b8.npz"])
def test_full_step_matches_golden_fixture(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name))
    e = _engine(8)
    params = rn.init_params(0)
    e.load_params(params)
    img, lab = dev(z["images"]), dev(z["labels"].reshape(-1))
    e.forward(img, lab)
    e.backward()
    torch.cuda.synchronize()
    assert relerr(e.logits().cpu().numpy(), z["logits"]) <= RTOL_F32
    assert abs(float(e.loss()) - float(z["loss"])) <= RTOL_F32 * abs(float(z["loss"]))
    gap = np.sort(z["logits"], axis=1)
    if ((gap[:, -1] - gap[:, -2]) > 1e-3 * np.abs(z["logits"]).max()).all():
        np.testing.assert_array_equal(e.pred().cpu().numpy(), z["pred"])
    got = e.get_grads()
    for k in rn.PARAM_NAMES:
        f = got[k].astype(np.float64).reshape(-1)
        idx = np.unique(np.linspace(0, f.size - 1, 256).astype(np.int64))
        scale = max(np.abs(z["gsmp_" + k]).max(), z["gsum_" + k][1] / f.size)
        # gradients below pool2 depend on discrete argmax choices; one fp32-vs-fp64 near-tie moves
        # O(1%) of a batch-8 gradient (measured with the oracle itself run in fp32), so the golden
        # digests pin them loosely here and test_full_step_b128_vs_oracle pins them tightly.
        loose = k.startswith("conv")
        np.testing.assert_allclose(f[idx], z["gsmp_" + k], rtol=0, atol=(5e-2 if loose else RTOL_F32 * 10) * scale)
        assert abs(np.abs(f).sum() - z["gsum_" + k][1]) <= (5e-2 if loose else 1e-4) * z["gsum_" + k][1]


@pytest.mark.parametrize("kind", ["u8", "soft"])
def test_full_step_b128_vs_oracle(kind):
    """BASELINE.json config 2: single B200, batch 128, fp32 validation mode."""
    rng = np.random.default_rng(11)
    B = 128
    if kind == "u8":
        img = rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)
    else:
        img = np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
    lab = rng.integers(0, 10, (B, 1)).astype(np.int32)
    params = rn.init_params(0)
    newp, loss, g, a = rn.train_step(params, img, lab, lr=np.float32(0.1))
    e = _engine(B)
    e.load_params(params)
    imgd, labd = dev(img), dev(lab.reshape(-1))
    e.forward(imgd, labd)
    e.backward()
    torch.cuda.synchronize()
    for name in ("conv1", "pool1", "conv2", "pool2", "full1", "full2", "logits"):
        assert relerr(e.tensor(name).cpu().numpy().reshape(a[name].shape), a[name]) <= RTOL_F32, name
    np.testing.assert_array_equal(e.tensor("x").cpu().numpy().reshape(a["x"].shape), a["x"])
    assert abs(float(e.loss()) - loss) <= RTOL_F32 * abs(loss)
    # bit-exact predictions wherever the top-2 gap exceeds the fp32 error bound
    srt = np.sort(a["logits"], axis=1)
    safe = (srt[:, -1] - srt[:, -2]) > 1e-4 * np.abs(a["logits"]).max()
    safe |= (srt[:, -1] == 0)                                      # all-zero row: argmax 0 on both sides
    np.testing.assert_array_equal(e.pred().cpu().numpy()[safe], a["pred"][safe])
    if safe.all():
        assert int(e.n_correct()) == a["n_correct"]
    g_on_gpu_fwd, a2 = _oracle_backward_on_gpu_forward(e, params, lab, a)
    flips = int((a2["arg1"] != a["arg1"]).sum() + (a2["arg2"] != a["arg2"]).sum())
    assert flips <= 64, flips                                        # near-ties only, a handful per batch
    for name in ("dlogits", "dfull2", "dfull1", "dpool2", "dconv2", "dpool1", "dconv1"):
        assert relerr(e.tensor(name).cpu().numpy().reshape(a2[name].shape), a2[name]) <= 2 * RTOL_F32, name
    _grad_check(e, g_on_gpu_fwd, 2 * RTOL_F32)
    for k in ("full_weight_1", "full_bias_1", "full_weight_2", "full_bias_2", "full_weight_3", "full_bias_3"):
        assert relerr(e.get_grads()[k], g[k]) <= 2 * RTOL_F32, k      # tie-insensitive: pure fp64 oracle
    e.sgd_apply(0.1)
    torch.cuda.synchronize()
    got = e.get_params()
    newp2 = rn.sgd_apply(params, g_on_gpu_fwd, np.float32(0.1))
    for k in rn.PARAM_NAMES:
        assert relerr(got[k], newp2[k]) <= 2 * RTOL_F32, k
    # padding floats between tensors stay zero (they ride along in the allreduce)
    used = torch.zeros(e.total, dtype=torch.bool)
    for o, s in zip(e.offsets, e.sizes):
        used[o:o + s] = True
    assert float(e.params.cpu()[~used].abs().sum()) == 0.0 and float(e.grads.cpu()[~used].abs().sum()) == 0.0


def test_train_step_equals_three_calls_and_graph_replay():
    rng = np.random.default_rng(12)
    B = 32
    img = dev(np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8))
    lab = dev(rng.integers(0, 10, (B,)).astype(np.int32))
    params = rn.init_params(0)
    e1, e2, e3 = _engine(B), _engine(B), _engine(B)
    for e in (e1, e2, e3):
        e.load_params(params)
    e1.forward(img, lab); e1.backward(); e1.sgd_apply(0.1)
    e2.train_step(img, lab, 0.1)
    torch.cuda.synchronize()
    assert torch.equal(e1.params, e2.params) and torch.equal(e1.loss(), e2.loss())
    # CUDA-graph replay issues the same kernels: results identical to eager, step after step
    g = e3.capture(lambda: e3.train_step(img, lab, 0.1))
    e3.load_params(params)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(e1.params, e3.params)
    e1.train_step(img, lab, 0.1)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(e1.params, e3.params) and torch.equal(e1.loss(), e3.loss())


def test_train_step_host_matches_device_path():
    rng = np.random.default_rng(13)
    B = 16
    img = torch.as_tensor(np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)).pin_memory()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).pin_memory()
    params = rn.init_params(0)
    e1, e2 = _engine(B), _engine(B)
    e1.load_params(params); e2.load_params(params)
    e1.train_step(img.cuda(), lab.cuda(), 0.1)
    loss = e2.train_step_host(img, lab, 0.1)
    torch.cuda.synchronize()
    assert torch.equal(e1.params, e2.params)
    assert loss == float(e1.loss())


def test_train_step_host_prefetch_matches_device_path():
    """b200_stage_host_batch + b200_train_step_host_prefetch: the step runs on the batch staged by the previous call
    while the next pinned host batch travels on the copy stream; results equal the device-input path, eagerly and
    from captured graphs replayed in a ring."""
    rng = np.random.default_rng(16)
    B, n = 16, 5
    batches = [(torch.as_tensor(np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)).pin_memory(),
                torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).pin_memory()) for _ in range(n)]
    params = rn.init_params(0)
    e1, e2, e3 = _engine(B), _engine(B), _engine(B)
    for e in (e1, e2, e3):
        e.load_params(params)
    want = []
    for img, lab in batches:
        e1.train_step(img.cuda(), lab.cuda(), 0.01)
        want.append(float(e1.loss()))
    e2.stage_host_batch(0, *batches[0])
    got = []
    for j in range(n):
        nxt = batches[j + 1] if j + 1 < n else (None, None)
        got.append(e2.train_step_host_prefetch(j % 2, nxt[0], nxt[1], 0.01, sync=True))
    assert got == want
    assert torch.equal(e1.params, e2.params)
    # captured: graph j trains on slot j % 2 and prefetches batch (j + 1) % 4
    outs = [torch.zeros(1).pin_memory() for _ in range(4)]
    e3.stage_host_batch(0, *batches[0])
    graphs = [e3.capture(lambda j=j: e3.train_step_host_prefetch(j % 2, *batches[(j + 1) % 4], 0.01, loss_out=outs[j]))
              for j in range(4)]
    e3.load_params(params)
    e3.stage_host_batch(0, *batches[0])
    seen = []
    for j in range(4):
        graphs[j].replay()
        torch.cuda.synchronize()
        seen.append(float(outs[j][0]))
    assert seen == want[:4]


def test_two_steps_follow_the_oracle_trajectory():
    """Two consecutive SGD steps on well-conditioned inputs stay within tolerance of the oracle
    (raw 0..255 inputs with lr 0.1 diverge after one step in the reference too -- SURVEY Q11)."""
    rng = np.random.default_rng(14)
    B = 64
    params = rn.init_params(0)
    e = _engine(B)
    e.load_params(params)
    p64 = {k: v.astype(np.float64) for k, v in params.items()}
    for _ in range(2):
        img = np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
        lab = rng.integers(0, 10, (B,)).astype(np.int32)
        p64, loss, _, _ = rn.train_step(p64, img, lab, lr=0.01)
        e.train_step(dev(img), dev(lab), 0.01)
        torch.cuda.synchronize()
        assert abs(float(e.loss()) - loss) <= 1e-4 * abs(loss)
    got = e.get_params()
    for k in rn.PARAM_NAMES:       # loose: a pool near-tie resolved differently moves ~1% of a conv gradient
        assert relerr(got[k], p64[k]) <= 2e-3, k


def test_uncropped_32x32_shape():
    """create_cnn is size-agnostic (cifar10cnn.py:126-127): 32x32 input -> flatten 4096."""
    rng = np.random.default_rng(15)
    B = 4
    img = np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
    lab = rng.integers(0, 10, (B,)).astype(np.int32)
    params = rn.init_params(0, crop_h=32, crop_w=32)
    e = _engine(B, crop_hw=(32, 32))
    e.load_params(params)
    e.forward(dev(img), dev(lab)); e.backward()
    torch.cuda.synchronize()
    _, loss, g, a = rn.train_step(params, img, lab, crop=(32, 32))
    assert relerr(e.logits().cpu().numpy(), a["logits"]) <= RTOL_F32
    g2, _ = _oracle_backward_on_gpu_forward(e, params, lab, a)
    _grad_check(e, g2, 2 * RTOL_F32)


def test_bad_inputs_raise():
    e = _engine(8)
    with pytest.raises(ValueError):
        e.forward(torch.zeros(8, 3, 32, 32, device="cuda"), torch.zeros(8, dtype=torch.int32, device="cuda"))
    with pytest.raises(ValueError):
        e.forward(torch.zeros(4, 3, 32, 32, dtype=torch.uint8, device="cuda"),
                  torch.zeros(4, dtype=torch.int32, device="cuda"))
    e2 = _engine(8)
    with pytest.raises(_lib.B200Error):
        e2.backward()                                  # backward before forward
    e2.forward(torch.zeros(8, 3, 32, 32, dtype=torch.uint8, device="cuda"), torch.zeros(8, dtype=torch.int32, device="cuda"))
    e2.backward()
    with pytest.raises(_lib.B200Error):
        e2.backward()                                  # one backward per forward


@pytest.mark.parametrize("B", [128, 6, 2048])
def test_fc_tail_fused(lib, B):
    """full2 / logits / loss / accuracy and every gradient above full1 in two launches."""
    F1, F2, NC = 384, 192, 10
    rng = np.random.default_rng(61)
    full1 = np.maximum(rng.normal(size=(B, F1)), 0).astype(np.float32)
    w2 = (rng.normal(size=(F1, F2)) * 0.05).astype(np.float32)
    b2 = (rng.normal(size=(F2,)) * 0.1).astype(np.float32)
    w3 = (rng.normal(size=(F2, NC)) * 0.1).astype(np.float32)
    b3 = (rng.normal(size=(NC,)) * 0.1).astype(np.float32)
    labels = rng.integers(0, NC, (B,)).astype(np.int32)
    t = lambda *s, dt=torch.float32: torch.full(s, -7, dtype=dt, device="cuda")
    full2, logits, dlogits, dfull2, dfull1 = t(B, F2), t(B, NC), t(B, NC), t(B, F2), t(B, F1)
    pred, loss, nc, nbad = t(B, dt=torch.int32), t(1), t(1, dt=torch.int32), t(1, dt=torch.int32)
    dw2, db2, dw3, db3, db1 = t(F1, F2), t(F2), t(F2, NC), t(NC), t(F1)
    nb = C.c_int64()
    ck(lib.b200_fc_tail_scratch_bytes(B, F1, F2, NC, C.byref(nb)))
    scr = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    f1d, w2d, b2d, w3d, b3d, labd = dev(full1), dev(w2), dev(b2), dev(w3), dev(b3), dev(labels)   # keep alive
    ck(lib.b200_fc_tail_f32(f1d.data_ptr(), w2d.data_ptr(), b2d.data_ptr(), w3d.data_ptr(),
                            b3d.data_ptr(), labd.data_ptr(), full2.data_ptr(), logits.data_ptr(),
                            dlogits.data_ptr(), dfull2.data_ptr(), dfull1.data_ptr(), pred.data_ptr(), loss.data_ptr(),
                            nc.data_ptr(), nbad.data_ptr(), dw2.data_ptr(), db2.data_ptr(), dw3.data_ptr(),
                            db3.data_ptr(), db1.data_ptr(), scr.data_ptr(), nb.value, B, F1, F2, NC, stream()))
    torch.cuda.synchronize()
    x = full1.astype(np.float64)
    # the two products with W2 (full2 and dfull1) run on bf16 tensor-core MMAs: their operands (W2, full1, dfull2)
    # are rounded to bf16, accumulation and everything else is fp32 -- so against the oracle on the rounded
    # operands the usual fp32 tolerance holds
    bf = lambda v: torch.as_tensor(np.asarray(v, dtype=np.float32)).to(torch.bfloat16).to(torch.float64).numpy()
    w2 = bf(w2)
    f2 = np.maximum(bf(x) @ w2 + b2, 0)
    lg = np.maximum(f2 @ w3 + b3, 0)
    wl, _, back = rn.softmax_xent(lg, labels)
    _, wpred, wn = rn.batch_accuracy(lg, labels)
    d = back / B * (lg > 0)
    d2 = (d @ w3.T.astype(np.float64)) * (f2 > 0)
    # dfull1's oracle starts from the GPU's own dfull2 (checked against d2 below): a last-bit difference in an fp32
    # value that sits on a bf16 rounding boundary would otherwise flip the operand the MMA sees
    d1 = (bf(dfull2.cpu().numpy()) @ w2.T.astype(np.float64)) * (x > 0)
    for got, want, name in ((full2, f2, "full2"), (logits, lg, "logits"), (dlogits, d, "dlogits"), (dfull2, d2, "dfull2"),
                            (dfull1, d1, "dfull1"), (dw2, x.T @ d2, "dw2"), (db2, d2.sum(0), "db2"),
                            (dw3, f2.T @ d, "dw3"), (db3, d.sum(0), "db3"), (db1, d1.sum(0), "db1")):
        assert relerr(got.cpu().numpy(), want) <= RTOL_F32, name
    assert abs(float(loss) - wl) <= RTOL_F32 * abs(wl)
    srt = np.sort(lg, axis=1)
    safe = ((srt[:, -1] - srt[:, -2]) > 1e-4) | (srt[:, -1] == 0)
    np.testing.assert_array_equal(pred.cpu().numpy()[safe], wpred[safe])
    if safe.all():
        assert int(nc) == wn
    assert int(nbad) == 0
