"""Tower mode: multi-GPU training on one node from a single command (north-star "multi-GPU tower batching").

The reference only alludes to towers (comment mnist_input.py:127-130, README link to mnist_multi_gpu_batching_train,
and the ``return_grad`` hook of ``train()``, mnist_input.py:246/:272).  On B200 the idiomatic shape of "k towers with
averaged gradients" is one process per GPU: each tower runs the full kernel sequence on its own sub-batch and the
tower gradients are averaged by the fused peer-memory all-reduce+update kernel (parallel/peer_comm.py) — the in-graph
"average_gradients" of the tower script becomes a single kernel over NVLink.  :func:`launch_towers` spawns the k
worker processes with a private loopback rendezvous and waits for them.
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch_towers(num_gpus: int, argv: list[str], *, script: str) -> int:
    ports = [_free_port() for _ in range(num_gpus)]
    worker_hosts = ",".join(f"127.0.0.1:{p}" for p in ports)
    procs = []
    for rank in range(num_gpus):
        env = dict(os.environ, LOCAL_RANK=str(rank), RANK=str(rank), WORLD_SIZE=str(num_gpus))
        cmd = [sys.executable, script, *argv, "--job_name=worker", f"--task_id={rank}", f"--worker_hosts={worker_hosts}"]
        procs.append(subprocess.Popen(cmd, env=env))
    rc = 0
    for p in procs:
        rc = p.wait() or rc
    if rc:
        for p in procs:
            if p.poll() is None:
                p.kill()
    return rc
