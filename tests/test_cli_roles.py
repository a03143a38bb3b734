"""CPU tests of the drop-in boundary: the reference's command line (cifar10cnn.py:245-273), the
ps / worker roles (:185-196) over gloo with two workers, the gradient-exchange semantics, the input
record reader (:54-62) and the learning-rate schedule (:161).  No GPU, no CUDA arithmetic."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_flags_match_the_reference():
    import cifar10cnn as cli
    f, unknown = cli.build_parser().parse_known_args([])
    assert (f.ps_hosts, f.worker_hosts, f.job_name, f.task_index) == ("", "", "", 0)
    assert f.data_dir == "/tmp/mnist_data" and f.log_dir == "/tmp/train_logs"
    f, unknown = cli.build_parser().parse_known_args(
        ["--ps_hosts=localhost:2222", "--worker_hosts=localhost:2223,localhost:2224", "--job_name=worker",
         "--task_index=1", "--some_future_flag=7"])
    assert f.worker_hosts.split(",") == ["localhost:2223", "localhost:2224"] and f.task_index == 1
    assert unknown == ["--some_future_flag=7"]                     # tolerated, like parse_known_args at :273
    assert (cli.BATCH_SIZE, cli.GENERATIONS, cli.OUTPUT_EVERY, cli.EVAL_EVERY) == (128, 20000, 200, 500)


def test_unknown_job_name_falls_through_silently():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "cifar10cnn.py"), "--job_name=evaluator"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout == ""


def test_worker_without_cuda_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "cifar10cnn.py"), "--job_name=worker",
                        "--worker_hosts=localhost:2223"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


def test_learning_rate_schedule_and_dead_generation_counter():
    from b200cnn import session
    gen = session.Variable(0)
    op = session.TrainOp(None, gen)
    assert op.learning_rate() == float(np.float32(0.1))
    gen.value = 499
    assert abs(op.learning_rate() - 0.09) < 1e-7
    gen.value = 500
    assert abs(op.learning_rate() - 0.081) < 1e-7


def test_record_reader(tmp_path):
    from b200cnn import data
    d = tmp_path / "cifar10data" / "cifar-10-batches-bin"
    d.mkdir(parents=True)
    rng = np.random.default_rng(0)
    rec = rng.integers(0, 256, (7, 3073), dtype=np.uint8)
    rec[:, 0] = np.arange(7)
    (d / "test_batch.bin").write_bytes(rec.tobytes())
    files = data.cifar_files(False, str(tmp_path / "cifar10data"))
    img, lab = data.read_records(files)
    assert img.shape == (7, 3, 32, 32) and lab.tolist() == list(range(7))
    np.testing.assert_array_equal(img[3].reshape(-1), rec[3, 1:])
    (d / "test_batch.bin").write_bytes(rec.tobytes()[:-1])
    with pytest.raises(ValueError):
        data.read_records(files)


def _ps_proc(port):
    from b200cnn import roles
    roles.serve_ps([f"127.0.0.1:{port}"], ["w0", "w1"], 0, poll_s=0.2, max_seconds=60)


def _worker_proc(rank, port, reduce, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from b200cnn import roles, exchange
    from oracle import ref_numpy as rn
    world, r = roles.init_worker([f"127.0.0.1:{port}"], ["w0", "w1"], rank, backend="gloo", timeout_s=60)
    assert (world, r) == (2, rank)
    ex = exchange.GradientExchange(reduce=reduce)
    params = {k: v.astype(np.float64) for k, v in rn.init_params(0).items()}
    flat = torch.cat([torch.as_tensor(params[k]).reshape(-1) for k in rn.PARAM_NAMES])
    if rank != 0:
        flat = flat + 1.0                                        # non-chief starts different ...
    ex.broadcast_params_(flat)                                   # ... the chief's variables win (:222)
    rng = np.random.default_rng(100 + rank)
    img = np.clip(np.rint(rng.normal(8, 2.5, (4, 3, 32, 32))), 0, 255).astype(np.uint8)
    lab = rng.integers(0, 10, (4,))
    g = rn.train_step(params, img, lab)[2]
    gflat = torch.cat([torch.as_tensor(g[k]).reshape(-1) for k in rn.PARAM_NAMES])
    own = gflat.clone()
    ex.allreduce_(gflat)
    flat = flat - 0.1 * ex.grad_scale * gflat                    # what b200_sgd_apply does on the GPU
    ok = ex.replicas_identical(flat.to(torch.float32))
    q.put((rank, own.numpy(), flat.numpy(), ok, ex.global_step_increment))
    dist.destroy_process_group()


@pytest.mark.parametrize("reduce", ["sum", "mean"])
def test_two_workers_and_ps_over_gloo(reduce):
    """README recipe shape: 1 ps + 2 workers on localhost.  One exchange round must equal the
    oracle's allreduce_step: SUM = two reference worker updates computed from one snapshot."""
    from oracle import ref_numpy as rn
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = ctx.Process(target=_ps_proc, args=(port,), daemon=True)
    ps.start()
    ws = [ctx.Process(target=_worker_proc, args=(r, port, reduce, q)) for r in range(2)]
    for w in ws:
        w.start()
    res = sorted([q.get(timeout=180) for _ in ws], key=lambda t: t[0])
    for w in ws:
        w.join(60)
        assert w.exitcode == 0
    ps.terminate()
    params = {k: v.astype(np.float64) for k, v in rn.init_params(0).items()}
    shapes = [params[k].shape for k in rn.PARAM_NAMES]
    sizes = [int(np.prod(s)) for s in shapes]

    def unflat(v):
        out, o = {}, 0
        for k, s, n in zip(rn.PARAM_NAMES, shapes, sizes):
            out[k] = v[o:o + n].reshape(s)
            o += n
        return out
    grads = [unflat(r[1]) for r in res]
    want = rn.allreduce_step(params, grads, 0.1, reduce)
    for r in res:
        got = unflat(r[2])
        for k in rn.PARAM_NAMES:
            np.testing.assert_allclose(got[k], want[k], rtol=1e-12, atol=1e-14)
        assert r[3] is True and r[4] == 2                        # replicas identical; global_step += world
    np.testing.assert_array_equal(res[0][2], res[1][2])


def test_checkpoint_names_and_roundtrip(tmp_path):
    """--log_dir contract (cifar10cnn.py:222): files keyed by the TF variable names, newest one restored."""
    from b200cnn import checkpoint
    from oracle import ref_numpy as rn
    names = checkpoint.variable_names()
    assert names[0] == "model_definition/conv1/conv1_kernel" and names[4] == "model_definition/full1/full_weight_1"
    assert names[-1] == "model_definition/full3/full_bias_3" and len(names) == 10
    p = rn.init_params(0)
    assert checkpoint.latest(str(tmp_path)) is None
    checkpoint.save(str(tmp_path), p, 200)
    p2 = {k: v + 1 for k, v in p.items()}
    path = checkpoint.save(str(tmp_path), p2, 1400)
    assert checkpoint.latest(str(tmp_path)) == path
    got, step = checkpoint.load(path)
    assert step == 1400
    for k in rn.PARAM_NAMES:
        np.testing.assert_array_equal(got[k], p2[k])
    z = np.load(path)
    assert z["model_definition/conv2/conv2_kernel"].shape == (5, 5, 64, 64) and int(z["Variable"]) == 0
