"""Multi-GPU tests (skipped on a box with fewer than two GPUs): one process per GPU, launched the way the
driver launches bench.py (torch.distributed.run on 127.0.0.1) or through the drop-in script's own roles."""
import os
import signal
import socket
import subprocess
import sys
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs at least two GPUs")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@needs2
@pytest.mark.parametrize("world", sorted({2, min(NGPU, 8)}) if NGPU >= 2 else [2])
def test_peer_memory_exchange_equals_nccl_and_oracle(world):
    """tests/multi_gpu_check.py under torchrun: for the same per-worker batches the fused peer-memory kernel, the NCCL
    allreduce + SGD kernel and the oracle's sum-update give the same parameters; replicas stay bit-identical through
    the plain and the overlapped exchange."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(r.stdout[-3000:])
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert "replicas identical: True / True" in r.stdout


def _run_cluster(tmp, steps):
    """README:11-13: one ps and two workers on localhost, through cifar10cnn.py's own flags."""
    ps, w0, w1 = free_port(), free_port(), free_port()
    common = [sys.executable, os.path.join(ROOT, "cifar10cnn.py"), f"--ps_hosts=127.0.0.1:{ps}",
              f"--worker_hosts=127.0.0.1:{w0},127.0.0.1:{w1}", f"--max_steps={steps}", f"--log_dir={tmp}"]
    procs = [subprocess.Popen(common + ["--job_name=ps", "--task_index=0"], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                              text=True, cwd=ROOT)]
    time.sleep(1.0)
    for i in range(2):
        procs.append(subprocess.Popen(common + ["--job_name=worker", f"--task_index={i}"], stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True, cwd=ROOT))
    outs = []
    try:
        for p in procs[1:]:
            o, e = p.communicate(timeout=600)
            outs.append((p.returncode, o, e))
    finally:
        procs[0].send_signal(signal.SIGTERM)          # the ps blocks forever, like server.join() (:192)
        try:
            procs[0].wait(timeout=20)
        except subprocess.TimeoutExpired:
            procs[0].kill()
        for p in procs[1:]:
            if p.poll() is None:
                p.kill()
    return outs


@needs2
def test_cli_one_ps_two_workers_is_reproducible(tmp_path):
    """The reference's launch recipe on two GPUs: both workers train through the captured fused step with the
    peer-memory exchange, global_step advances by two per round, and a second run from scratch ends in a
    bit-identical checkpoint (fixed-order reductions everywhere, counter-based input queue)."""
    from b200cnn import checkpoint
    finals = []
    for run in range(2):
        d = tmp_path / f"run{run}"
        outs = _run_cluster(str(d), steps=60)
        for rc, o, e in outs:
            assert rc == 0, e[-3000:]
            assert "Starting Training" in o.splitlines()          # (NCCL may print its version line first)
        path = checkpoint.latest(str(d))
        assert path and path.endswith("model.ckpt-60.npz")
        finals.append(checkpoint.load(path)[0])
    for k in finals[0]:
        np.testing.assert_array_equal(finals[0][k], finals[1][k], err_msg=k)
