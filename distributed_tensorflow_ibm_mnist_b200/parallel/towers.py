"""Tower mode: multi-GPU training on one node from a single command (north-star "multi-GPU tower batching").

The reference only alludes to towers (comment mnist_input.py:127-130, README link to mnist_multi_gpu_batching_train,
and the ``return_grad`` hook of ``train()``, mnist_input.py:246/:272).  On B200 the idiomatic shape of "k towers with
averaged gradients" is one process per GPU: each tower runs the full kernel sequence on its own sub-batch and the
tower gradients are averaged by the fused peer-memory all-reduce+update kernel (parallel/peer_comm.py) — the in-graph
"average_gradients" of the tower script becomes a single kernel over NVLink.  Two forms: :func:`run_tower_threads` drives the k
GPUs from ONE process (one thread + one engine per GPU, ``--num_gpus=k``); :func:`launch_towers` spawns k worker processes
with a private loopback rendezvous (``--tower_procs``, and the only form on CPU) and supervises them.
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
import time


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(num_workers: int, argv: list[str], script: str, attempt: int) -> list[subprocess.Popen]:
    procs = []
    if num_workers <= 1:     # one supervised worker in local mode
        env = dict(os.environ, DTM_SUPERVISED="1", DTM_RESTART_COUNT=str(attempt))
        return [subprocess.Popen([sys.executable, script, *argv], env=env)]
    ports = [_free_port() for _ in range(num_workers)]
    worker_hosts = ",".join(f"127.0.0.1:{p}" for p in ports)
    for rank in range(num_workers):
        env = dict(os.environ, LOCAL_RANK=str(rank), RANK=str(rank), WORLD_SIZE=str(num_workers), DTM_SUPERVISED="1",
                   DTM_RESTART_COUNT=str(attempt))
        cmd = [sys.executable, script, *argv, "--job_name=worker", f"--task_id={rank}", f"--worker_hosts={worker_hosts}"]
        procs.append(subprocess.Popen(cmd, env=env))
    return procs


def _stop(procs: list[subprocess.Popen], grace_s: float = 5.0) -> None:
    """Terminate exactly the processes this supervisor started (never by name pattern)."""
    for p in procs:
        if p.poll() is None:
            p.terminate()
    t0 = time.time()
    for p in procs:
        try:
            p.wait(timeout=max(0.1, grace_s - (time.time() - t0)))
        except subprocess.TimeoutExpired:
            p.kill()
            p.wait()


def launch_towers(num_gpus: int, argv: list[str], *, script: str, max_restarts: int = 0, poll_s: float = 0.2, out=print) -> int:
    """Run ``num_gpus`` worker processes and supervise them: the process-level form of ``MonitoredTrainingSession``'s
    recoverable session (reference main.py:140-146).

    Synchronous replicas live and die together: when one worker exits with an error (a crash, a trapped kernel watchdog
    after a peer vanished, a lost connection) the survivors cannot make progress, so the supervisor stops them, and - up
    to ``max_restarts`` times - starts the whole set again.  Every new worker restores the newest checkpoint in
    ``--train_dir`` (``restore-or-init`` in main.py), so training continues from the last saved global step with
    bit-identical state.  ``DTM_RESTART_COUNT`` tells the workers which attempt they belong to."""
    attempt = 0
    while True:
        procs = _spawn(num_gpus, argv, script, attempt)
        failed = None
        while failed is None and any(p.poll() is None for p in procs):
            for k, p in enumerate(procs):
                rc = p.poll()
                if rc not in (None, 0):
                    failed = (k, rc)
                    break
            if failed is None:
                time.sleep(poll_s)
        if failed is None:
            failed = next(((k, p.returncode) for k, p in enumerate(procs) if p.returncode), None)
        if failed is None:
            return 0
        _stop(procs)
        if attempt >= max_restarts:
            out(f"[supervisor] worker {failed[0]} exited with status {failed[1]}; no restarts left ({attempt} used)")
            return failed[1] if failed[1] > 0 else 1
        attempt += 1
        out(f"[supervisor] worker {failed[0]} exited with status {failed[1]}; restarting all {max(1, num_gpus)} worker(s) from the "
            f"newest checkpoint (restart {attempt} of {max_restarts})")


def run_tower_threads(num_gpus: int, worker) -> int:
    """Single-process tower mode: ``num_gpus`` threads of THIS process, each driving one GPU with its own engine, streams and
    CUDA graphs; ``worker(role, plane)`` is the ordinary training function.  The towers' gradients meet in the same fused
    peer-memory kernels as in the multi-process form - the symmetric buffers come from the VMM allocator, every GPU maps its
    peers' allocations and the multicast object directly (no IPC handles, no sockets, no ports, no torch.distributed).
    Python's interpreter lock serialises the towers' host code, so the per-step launch work of k towers adds up: with many
    GPUs the process form (``--tower_procs``) keeps the GPUs busier."""
    import threading
    import traceback

    from ..cluster.spec import Role
    from .control_plane import ThreadPlane

    from ..ops import _lib

    _lib.load()
    print(f"single-process tower mode: {num_gpus} GPUs driven by pid {os.getpid()}", flush=True)
    planes = ThreadPlane.make(num_gpus)
    rcs = [1] * num_gpus

    def run(rank: int) -> None:
        try:
            rcs[rank] = int(worker(Role("worker", rank, rank == 0, rank, num_gpus), planes[rank]) or 0)
        except BaseException:  # noqa: BLE001 - report, then release the siblings that wait for this tower
            traceback.print_exc()
            rcs[rank] = 1
            planes[rank].abort()

    threads = [threading.Thread(target=run, args=(r,), name=f"tower-{r}") for r in range(num_gpus)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    return max(rcs)
