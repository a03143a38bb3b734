"""GPU tests of the drop-in script and the session surface (cifar10cnn.py:198-242)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_single_worker_cli_runs_and_prints_the_reference_lines():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "cifar10cnn.py"), "--job_name=worker",
                        "--worker_hosts=localhost:2223", "--ps_hosts=localhost:2222", "--max_steps=500",
                        "--log_dir="],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.strip().splitlines()
    assert lines[0] == "Starting Training"
    assert any(l.startswith("global_step 200, task:0_step 199, training accuracy ") for l in lines)
    assert any(l.startswith(" --- Test Accuracy = ") and l.endswith("%.") for l in lines)   # Q3 fixed: no crash


def test_cli_checkpoint_and_resume(tmp_path):
    """chief writes model.ckpt-<step>.npz into --log_dir at the end and a second run resumes from it."""
    from b200cnn import checkpoint
    cmd = [sys.executable, os.path.join(ROOT, "cifar10cnn.py"), "--job_name=worker", "--worker_hosts=localhost:2223",
           f"--log_dir={tmp_path}", "--mode=fp32"]
    r = subprocess.run(cmd + ["--max_steps=3"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    first = checkpoint.latest(str(tmp_path))
    assert first and first.endswith("model.ckpt-3.npz")
    r = subprocess.run(cmd + ["--max_steps=5"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    assert checkpoint.latest(str(tmp_path)).endswith("model.ckpt-5.npz")          # resumed at 3, ran 2 more
    r = subprocess.run(cmd + ["--max_steps=5"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "Starting Training" in r.stdout                       # already done: no step


def test_session_towers_share_variables_and_step_matches_engine():
    import b200cnn
    from b200cnn import session
    import cifar10cnn as cli
    B = 16
    itr, ltr = session.input_pipeline(B, True, seed=5, synthetic=True)
    ite, lte = session.input_pipeline(B, False, seed=6, synthetic=True)
    scope = session.ModelScope(mode="fp32", seed=0)
    out_tr = cli.create_cnn(itr, B, scope)
    out_te = cli.create_cnn(ite, B, scope)
    assert out_tr.engine.params.data_ptr() == out_te.engine.params.data_ptr()       # reuse_variables (:209)
    loss = cli.cifar_loss(out_tr, ltr)
    acc_te = cli.batch_accuracy(out_te, lte)
    op = cli.train_step(loss, session.Variable(0))
    # reference engine fed the same batch stream
    ref = b200cnn.TrainStepEngine(batch=B, mode="fp32")
    ref.init_params(0)
    src = b200cnn.data.BatchSource(B, True, seed=5, synthetic=True)
    with session.MonitoredTrainingSession(last_step=3) as sess:
        n = 0
        while not sess.should_stop():
            _, lv = sess.run([op, loss])
            img, lab = src.next()
            ref.train_step(img, lab, 0.1)
            torch.cuda.synchronize()
            assert lv == float(ref.loss()) or (np.isnan(lv) and np.isnan(float(ref.loss())))
            n += 1
        assert n == 3 and sess.global_step == 3
        a = sess.run(acc_te)
        assert 0.0 <= a <= 1.0 or np.isnan(a)
    assert torch.equal(out_tr.engine.params, ref.params)


@pytest.mark.gpu
def test_graft_entry_smoke_passes():
    """__graft_entry__.smoke() (what the driver runs before the bench) on this build: both modes against the oracle."""
    import __graft_entry__ as entry
    entry.smoke()
