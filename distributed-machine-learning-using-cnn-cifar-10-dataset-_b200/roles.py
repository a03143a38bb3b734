"""Cluster roles of cifar10cnn.py:185-196 on top of torch.distributed.

  ps      (:191-192)  hosts the rendezvous store on ps_hosts[0] and blocks, like server.join().  The
                      variables it used to hold live replicated on the workers; the push/apply it
                      used to execute is the allreduce + fused SGD kernel (exchange.py).
  worker  (:193-242)  rank = task_index, world = len(worker_hosts), chief = task 0 (:222).
Any other job_name falls through silently, as in the reference."""
from __future__ import annotations

import datetime
import time

import torch.distributed as dist


def parse_hosts(s):
    return [h for h in s.split(",") if h]


def _host_port(hostport):
    host, _, port = hostport.rpartition(":")
    return (host or "127.0.0.1"), int(port)


def serve_ps(ps_hosts, worker_hosts, task_index=0, poll_s=1.0, max_seconds=None):
    """ps role.  Only ps task 0 hosts the store; further ps tasks just idle (there is nothing left
    to shard: 4.27 MB of variables are replicated).  Blocks forever unless max_seconds is given."""
    store = None
    if task_index == 0:
        host, port = _host_port(ps_hosts[0])
        store = dist.TCPStore(host, port, is_master=True, wait_for_workers=False,
                              timeout=datetime.timedelta(days=365))
        store.set("num_workers", str(len(worker_hosts)))
    t0 = time.time()
    while max_seconds is None or time.time() - t0 < max_seconds:
        time.sleep(poll_s)
    return store


def init_worker(ps_hosts, worker_hosts, task_index, backend="nccl", timeout_s=600, device_id=None):
    """worker role rendezvous: join the process group through the ps-hosted store.  With a single
    worker no process group is needed (and none is created)."""
    world = len(worker_hosts)
    if world <= 1:
        return 1, 0
    host, port = _host_port(ps_hosts[0])
    store = dist.TCPStore(host, port, is_master=False, timeout=datetime.timedelta(seconds=timeout_s))
    kw = {}
    if device_id is not None and backend == "nccl":
        kw["device_id"] = device_id
    dist.init_process_group(backend, store=dist.PrefixStore("b200cnn", store), rank=task_index, world_size=world,
                            timeout=datetime.timedelta(seconds=timeout_s), **kw)
    return world, task_index
