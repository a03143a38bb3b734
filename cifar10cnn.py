#!/usr/bin/env python
"""Drop-in launcher for the B200-native training step.

Same command line, roles, hyper-parameters and stdout lines as the reference script
(/root/reference/cifar10cnn.py: flags :245-273, roles :185-196, loop :221-242), e.g.

  python cifar10cnn.py --ps_hosts=localhost:2222 --worker_hosts=localhost:2223,localhost:2224 --job_name=ps --task_index=0
  python cifar10cnn.py --ps_hosts=localhost:2222 --worker_hosts=localhost:2223,localhost:2224 --job_name=worker --task_index=0
  python cifar10cnn.py --ps_hosts=localhost:2222 --worker_hosts=localhost:2223,localhost:2224 --job_name=worker --task_index=1

What differs underneath: worker i computes on GPU (i mod visible GPUs) with the CUDA library of
include/b200cnn.h; the parameter server only hosts the rendezvous (ps_hosts[0]) because the
push/apply it used to run is an allreduce + fused SGD kernel on the workers.  Deliberate fixes of
reference defects: the test-accuracy print evaluates the accuracy tensor
instead of crashing on an empty list; global_step is one shared counter; no download is
attempted -- cifar10data/cifar-10-batches-bin/ is used when present, synthetic records otherwise.
"""
import argparse
import sys

# hyper-parameters of the reference, cifar10cnn.py:11-23
OUTPUT_EVERY, EVAL_EVERY = 200, 500
BATCH_SIZE, GENERATIONS = 128, 20000
IMAGE_HEIGHT = IMAGE_WIDTH = 32
CROP_HEIGHT = CROP_WIDTH = 24
NUM_CHANNELS, NUM_TARGETS = 3, 10

FLAGS = None

# (flag, type, default, help) -- the six flags of :250-272 plus optional extras (unknown flags are
# tolerated, as with parse_known_args at :273)
_REFERENCE_FLAGS = (
    ("--ps_hosts", str, "", "Comma-separated list of hostname:port pairs"),
    ("--worker_hosts", str, "", "Comma-separated list of hostname:port pairs"),
    ("--job_name", str, "", "One of 'ps', 'worker'"),
    ("--task_index", int, 0, "Index of task within the job"),
    ("--data_dir", str, "/tmp/mnist_data", "Accepted for compatibility; unused, as in the reference"),
    ("--log_dir", str, "/tmp/train_logs", "Directory of train logs"),
)
_EXTRA_FLAGS = (
    ("--mode", str, None, "fp32 (validation) or bf16 (tensor cores); default: package default"),
    ("--reduce", str, "sum", "sum (N worker updates per round, the reference's semantics) or mean"),
    ("--max_steps", int, GENERATIONS, "stop when global_step reaches this (reference: 20000)"),
    ("--backend", str, "nccl", "torch.distributed backend for the worker group"),
    ("--staleness", int, 0, "asynchronous-ps fidelity: apply every gradient this many exchange rounds after the "
                            "snapshot it was computed on (-1: workers - 1, the expected staleness of the reference's "
                            "lock-free updates with that many workers); 0 = synchronous (default)"),
)


def build_parser():
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    for name, typ, default, text in _REFERENCE_FLAGS + _EXTRA_FLAGS:
        p.add_argument(name, type=typ, default=default, help=text)
    return p


# ---- the four functions of the reference, same names and argument meaning --------------------------
def create_cnn(input_images, batch_size, scope=None):
    """cifar10cnn.py:94-147.  Returns the logits node of a tower; towers created in the same
    scope share the ten variables (:204-210)."""
    import b200cnn
    scope = scope or _default_scope()
    return b200cnn.session.LogitsNode(input_images, scope.tower(batch_size))


def cifar_loss(logits, targets):
    """cifar10cnn.py:150-157."""
    import b200cnn
    return b200cnn.session.LossNode(logits, targets)


def train_step(loss_value, generation_num):
    """cifar10cnn.py:159-164: staircase-decayed SGD; minimize() = backward + apply + global_step."""
    import b200cnn
    return b200cnn.session.TrainOp(loss_value, generation_num)


def batch_accuracy(logits, targets):
    """cifar10cnn.py:166-176."""
    import b200cnn
    return b200cnn.session.AccuracyNode(logits, targets)


_scope = None


def _default_scope():
    global _scope
    if _scope is None:
        import b200cnn
        _scope = b200cnn.session.ModelScope(mode=getattr(FLAGS, "mode", None))
    return _scope


def run_worker(flags, ps_hosts, worker_hosts):
    import torch
    import b200cnn
    from b200cnn import roles, session

    if not torch.cuda.is_available():
        raise SystemExit("cifar10cnn.py worker: no CUDA device (this build has no CPU fallback)")
    dev = flags.task_index % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    roles.init_worker(ps_hosts, worker_hosts, flags.task_index, backend=flags.backend,
                      device_id=torch.device("cuda", dev))

    images_train, labels_train = session.input_pipeline(BATCH_SIZE, True, seed=1234 + flags.task_index)
    images_test, labels_test = session.input_pipeline(BATCH_SIZE, False, seed=4321 + flags.task_index)
    scope = session.ModelScope(mode=flags.mode, reduce=flags.reduce)
    model_output = create_cnn(images_train, BATCH_SIZE, scope)
    test_output = create_cnn(images_test, BATCH_SIZE, scope)
    loss = cifar_loss(model_output, labels_train)
    accuracy_test = batch_accuracy(test_output, labels_test)
    accuracy_train = batch_accuracy(model_output, labels_train)
    generation_num = session.Variable(0)
    train_op = train_step(loss, generation_num)

    import json
    import time
    staleness = len(worker_hosts) - 1 if flags.staleness < 0 else flags.staleness
    with session.MonitoredTrainingSession(is_chief=(flags.task_index == 0), last_step=flags.max_steps,
                                          reduce=flags.reduce, checkpoint_dir=flags.log_dir,
                                          staleness=staleness) as mon_sess:
        i = 0
        print('Starting Training')
        t_start, i_start = None, 0
        while not mon_sess.should_stop():
            if i == 20 and t_start is None:          # the first steps build and capture the step graph: not timed
                torch.cuda.synchronize()
                t_start, i_start = time.perf_counter(), i
            _, loss_value = mon_sess.run([train_op, loss])
            if loss_value is None:          # restored checkpoint already at the last step
                break
            if (i + 1) % OUTPUT_EVERY == 0:
                print('global_step %s, task:%d_step %d, training accuracy %g'
                      % (mon_sess.global_step, flags.task_index, i, mon_sess.run([accuracy_train])[0]))
            if (i + 1) % EVAL_EVERY == 0:
                print(' --- Test Accuracy = {:.2f}%.'.format(100. * mon_sess.run([accuracy_test])[0]))
            i += 1
        torch.cuda.synchronize()
        if t_start is not None and i > i_start:
            el = time.perf_counter() - t_start
            world = len(worker_hosts)
            # one JSON metrics line for harnesses (not in the reference's output): loop throughput of THIS launcher,
            # the two periodic accuracy passes included
            print(json.dumps({"b200cnn_metrics": {
                "task": flags.task_index, "workers": world, "local_steps": i - i_start, "seconds": el,
                "steps_per_sec": (i - i_start) / el, "images_per_sec_this_worker": (i - i_start) * BATCH_SIZE / el,
                "images_per_sec_all_workers": (i - i_start) * BATCH_SIZE * world / el,
                "data": "synthetic" if images_train.source.synthetic else "files", "mode": "bf16" if scope.engines[0].mode == 1 else "fp32"}}))
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


def main(flags):
    from b200cnn import roles
    ps_hosts = roles.parse_hosts(flags.ps_hosts)
    worker_hosts = roles.parse_hosts(flags.worker_hosts)
    if flags.job_name == "ps":
        roles.serve_ps(ps_hosts, worker_hosts, flags.task_index)          # blocks, like server.join() (:192)
    elif flags.job_name == "worker":
        run_worker(flags, ps_hosts, worker_hosts or ["localhost:0"])
    # any other job_name: fall through silently, as the reference does (:191-193)


if __name__ == "__main__":
    FLAGS, _unparsed = build_parser().parse_known_args()
    sys.exit(main(FLAGS))
