"""Control plane of the peer-memory communicator: the few cold, host-side collectives it needs (rendezvous, handle
exchange, decisions that every rank must take together), behind one small interface with two implementations.

* :class:`DistPlane` — ranks are processes; ``torch.distributed`` (NCCL or gloo) carries objects and barriers, and file
  descriptors of GPU allocations travel over unix-domain sockets (``SCM_RIGHTS``).  This is what replaces the reference's
  gRPC session plumbing between worker processes (main.py:42-68, :140-146).
* :class:`ThreadPlane` — ranks are threads of ONE process, each driving its own GPU (single-process tower mode, the
  reference's multi-GPU tower idea: mnist_input.py:127-130, README.md:8).  Objects are exchanged through shared memory of
  the process, barriers are ``threading.Barrier``; no sockets, no ports, no ``torch.distributed``.

The hot path never comes here: gradients, weights and barrier flags move between GPUs inside our own kernels.
"""
from __future__ import annotations

import os
import secrets
import socket
import threading

import torch


class DistPlane:
    def __init__(self, group):
        import torch.distributed as dist

        self.dist, self.group = dist, group
        self.rank, self.world_size = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)
        self._token = None
        self._seq = 0
        self.same_process = False

    def barrier(self) -> None:
        self.dist.barrier(group=self.group)

    def all_gather_object(self, obj) -> list:
        out = [None] * self.world_size
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def broadcast_object(self, obj, src: int = 0):
        box = [obj]
        g = self.group
        self.dist.broadcast_object_list(box, src=self.dist.get_global_rank(g, src) if g is not self.dist.group.WORLD else src, group=g)
        return box[0]

    def _reduce_flag(self, flag: bool, reduce_any: bool) -> bool:
        # gloo: one int32 through all_reduce instead of a pickled object through all_gather_object - the trainer's hooks ask
        # this every hundred steps of 150 us, so the question has to be cheap.  NCCL keeps the object path (it reduces device
        # tensors only, and either form ends in a host read of the result there).
        if self.backend != "gloo":
            flags = self.all_gather_object(bool(flag))
            return any(flags) if reduce_any else all(flags)
        t = torch.tensor([1 if flag else 0], dtype=torch.int32)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if reduce_any else self.dist.ReduceOp.MIN, group=self.group)
        return bool(int(t.item()))

    def all_true(self, flag: bool) -> bool:
        return self._reduce_flag(flag, False)

    def any_true(self, flag: bool) -> bool:
        return self._reduce_flag(flag, True)

    def broadcast_tensor(self, t: torch.Tensor, src: int) -> None:
        g = self.group
        self.dist.broadcast(t, src=self.dist.get_global_rank(g, src) if g is not self.dist.group.WORLD else src, group=g)

    # ---- file descriptors between processes --------------------------------------------------------------------
    def exchange_fds(self, fds: list[int]) -> list[list[int]]:
        """Every rank offers ``fds`` (possibly none); returns, per rank, duplicates of that rank's descriptors valid in THIS
        process (the own entry is the list passed in).  Abstract unix sockets named by a job token; rank j serves, the others
        fetch, for j = 0..N-1."""
        if self._token is None:
            self._token = self.broadcast_object(secrets.token_hex(8) if self.rank == 0 else None)
        self._seq += 1
        name = lambda r: f"\0dtm-{self._token}-{self._seq}-{r}"
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name(self.rank))
        srv.listen(self.world_size)
        counts = self.all_gather_object(len(fds))      # also the "everybody listens" barrier

        def serve():
            for _ in range(self.world_size - 1):
                conn, _ = srv.accept()
                with conn:
                    if fds:
                        socket.send_fds(conn, [b"f"], list(fds))
                    else:
                        conn.sendall(b"n")

        th = threading.Thread(target=serve, daemon=True)
        th.start()
        out: list[list[int]] = []
        for j in range(self.world_size):
            if j == self.rank:
                out.append(list(fds))
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.connect(name(j))
                if counts[j]:
                    _, got, _, _ = socket.recv_fds(c, 16, counts[j])
                    if len(got) != counts[j]:
                        raise RuntimeError(f"expected {counts[j]} descriptors from rank {j}, got {len(got)}")
                    out.append(list(got))
                else:
                    c.recv(16)
                    out.append([])
        th.join(timeout=60)
        srv.close()
        self.barrier()
        return out


class _ThreadShared:
    def __init__(self, world_size: int):
        self.world_size = world_size
        self.barrier = threading.Barrier(world_size)
        self.slots: list = [None] * world_size
        self.lock = threading.Lock()


class ThreadPlane:
    """One instance per tower thread; all instances of a job share one :class:`_ThreadShared`."""

    backend = "threads"
    same_process = True

    def __init__(self, shared: _ThreadShared, rank: int):
        self.shared, self.rank, self.world_size = shared, rank, shared.world_size

    @staticmethod
    def make(world_size: int) -> list["ThreadPlane"]:
        sh = _ThreadShared(world_size)
        return [ThreadPlane(sh, r) for r in range(world_size)]

    def barrier(self) -> None:
        self.shared.barrier.wait()

    def all_gather_object(self, obj) -> list:
        self.shared.slots[self.rank] = obj
        self.shared.barrier.wait()
        out = list(self.shared.slots)
        self.shared.barrier.wait()          # nobody overwrites a slot before everybody has read it
        return out

    def broadcast_object(self, obj, src: int = 0):
        return self.all_gather_object(obj if self.rank == src else None)[src]

    def all_true(self, flag: bool) -> bool:
        return all(self.all_gather_object(bool(flag)))

    def any_true(self, flag: bool) -> bool:
        return any(self.all_gather_object(bool(flag)))

    def broadcast_tensor(self, t: torch.Tensor, src: int) -> None:
        # staged through host memory on purpose: a direct cross-device copy makes torch enable peer access for ALL of its
        # allocations, after which every cudaMalloc / cudaFree synchronises the sibling devices too
        views = self.all_gather_object(t.cpu() if self.rank == src else None)
        if self.rank != src:
            t.copy_(views[src])
            if t.is_cuda:
                torch.cuda.synchronize(t.device)
        self.shared.barrier.wait()          # the source keeps its tensor alive and unchanged until everybody has copied

    def abort(self) -> None:
        self.shared.barrier.abort()


def as_plane(group_or_plane):
    """Accept a ``torch.distributed`` process group or a plane."""
    if hasattr(group_or_plane, "all_gather_object") and hasattr(group_or_plane, "world_size"):
        return group_or_plane
    return DistPlane(group_or_plane)
