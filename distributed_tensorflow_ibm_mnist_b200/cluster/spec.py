"""Cluster description and process roles — the stand-in for ``tf.train.ClusterSpec`` / ``tf.train.Server``.

Reference behaviour (main.py:42-68): ``--ps_hosts`` and ``--worker_hosts`` are comma-separated host:port lists; with at
least one worker host a server is started for (``--job_name``, ``--task_id``) and a ``ps`` process blocks forever in
``server.join()``; without worker hosts the process runs in local mode and ``task_id`` is forced to 0.

Here the data plane is NVLink peer memory / NCCL (or gloo on CPU), so the cluster layer only has to provide
rendezvous: a ``torch.distributed.TCPStore``.  The first ps task hosts it when parameter servers are configured
(its only remaining duty — parameters live on the workers' GPUs and are updated by the fused all-reduce kernel),
otherwise worker 0 hosts it.  ``ps`` processes still block like ``server.join()`` but return once every worker has
reported completion, which fixes the reference's need to kill them by hand.
"""
from __future__ import annotations

import datetime
import time
from dataclasses import dataclass, field


def _split_hosts(s: str | None) -> list[str]:
    return [h.strip() for h in s.split(",") if h.strip()] if s else []


def _host_port(hp: str) -> tuple[str, int]:
    host, _, port = hp.rpartition(":")
    if not host or not port.isdigit():
        raise ValueError(f"expected host:port, got {hp!r}")
    return host, int(port)


@dataclass
class ClusterSpec:
    worker_hosts: list[str] = field(default_factory=list)
    ps_hosts: list[str] = field(default_factory=list)

    @classmethod
    def from_flags(cls, ps_hosts: str | None, worker_hosts: str | None) -> "ClusterSpec":
        return cls(_split_hosts(worker_hosts), _split_hosts(ps_hosts))

    def as_dict(self) -> dict:
        d = {}
        if self.worker_hosts:
            d["worker"] = list(self.worker_hosts)
        if self.ps_hosts:
            d["ps"] = list(self.ps_hosts)
        return d

    @property
    def is_local(self) -> bool:
        """No worker hosts -> in-process local mode (main.py:63-66); ps hosts alone do not make a cluster (main.py:56)."""
        return len(self.worker_hosts) == 0

    @property
    def num_workers(self) -> int:
        return max(1, len(self.worker_hosts))

    @property
    def num_ps(self) -> int:
        return len(self.ps_hosts)

    def rendezvous(self) -> tuple[str, int]:
        """(host, port) of the TCPStore: first ps task if any, else first worker."""
        return _host_port(self.ps_hosts[0] if self.ps_hosts else self.worker_hosts[0])


@dataclass
class Role:
    job_name: str     # "worker" | "ps" | "" (local)
    task_id: int
    is_chief: bool
    rank: int         # rank among workers (ps: -1)
    world_size: int   # number of workers


def resolve_role(cluster: ClusterSpec, job_name: str, task_id: int) -> Role:
    if cluster.is_local:
        return Role("", 0, True, 0, 1)
    if job_name not in ("ps", "worker"):
        raise ValueError('--job_name must be "ps" or "worker" when --worker_hosts is given')
    n = cluster.num_ps if job_name == "ps" else len(cluster.worker_hosts)
    if not 0 <= task_id < n:
        raise ValueError(f"--task_id {task_id} out of range for job {job_name!r} with {n} task(s)")
    if job_name == "ps":
        return Role("ps", task_id, False, -1, len(cluster.worker_hosts))
    return Role("worker", task_id, task_id == 0, task_id, len(cluster.worker_hosts))


_DONE_KEY = "dtm/workers_done"


def _store(cluster: ClusterSpec, is_master: bool, timeout_s: float = 300.0):
    import torch.distributed as dist

    host, port = cluster.rendezvous()
    return dist.TCPStore(host, port, None, is_master, timeout=datetime.timedelta(seconds=timeout_s), wait_for_workers=False)


def serve_parameter_server(cluster: ClusterSpec, role: Role, poll_s: float = 0.2, max_wait_s: float | None = None) -> None:
    """What a ``ps`` process does: ps task 0 hosts the rendezvous store; every ps task blocks until the workers are done."""
    store = _store(cluster, is_master=(role.task_id == 0))
    t0 = time.time()
    while True:
        try:
            done = int(store.add(_DONE_KEY, 0))
        except Exception:
            return  # the store went away: the job is over
        if done >= len(cluster.worker_hosts):
            return
        if max_wait_s is not None and time.time() - t0 > max_wait_s:
            return
        time.sleep(poll_s)


def init_worker_group(cluster: ClusterSpec, role: Role, backend: str):
    """Join the worker process group; returns (group, store)."""
    import torch.distributed as dist

    hosts_store = (not cluster.ps_hosts) and role.task_id == 0
    store = _store(cluster, is_master=hosts_store)
    dist.init_process_group(backend, store=dist.PrefixStore("pg", store), rank=role.rank, world_size=role.world_size)
    return dist.group.WORLD, store


def signal_worker_done(store) -> None:
    try:
        store.add(_DONE_KEY, 1)
    except Exception:
        pass
