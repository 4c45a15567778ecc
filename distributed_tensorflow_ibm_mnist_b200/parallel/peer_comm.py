"""PeerComm — the NVLink peer-memory communicator (python side of csrc/comm/peer_comm.cu).

``torch.distributed`` is only the control plane here: it sets up the symmetric buffers once at start-up (either torch's
symmetric-memory rendezvous, which also yields an NVLS multicast address when the NVSwitch offers one, or our own CUDA
IPC handle exchange) and is used for rare, cold operations (gathering sharded optimizer state for a checkpoint).  Every hot collective — the
gradient all-reduce fused with the optimizer, the initial parameter / global_step broadcast, the stand-alone
all-reduce — is one of our own kernels reading and writing peer GPU memory directly.

Replaces the reference's parameter-server traffic: variable pulls and gradient pushes over gRPC (main.py:80-93,
mnist_input.py:262-264), chief initialisation + readiness polling (MonitoredTrainingSession, main.py:140-146).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_int, c_size_t, c_void_p

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import check, current_stream, ptr

MAX_RANKS = 8
P, I = c_void_p, c_int
_lib.register("dtm_symm_alloc", [c_size_t, P, P])
_lib.register("dtm_symm_open", [P, P])
_lib.register("dtm_symm_close", [P])
_lib.register("dtm_symm_free", [P])
_lib.register("dtm_fused_allreduce_update", [P, P, P, I, I, P, P, P, P, P, I, I, P])
_lib.register("dtm_allreduce_f32", [P, P, P, I, I, P, I, P, P])
_lib.register("dtm_broadcast_f32", [P, P, I, I, I, P, P])
_lib.register("dtm_set_comm_watchdog", [ctypes.c_longlong])


class PeerCtx(ctypes.Structure):
    """Mirror of ``dtm_peer_ctx``."""

    _fields_ = [("rank", c_int), ("world", c_int), ("grad", c_void_p * MAX_RANKS), ("w", c_void_p * MAX_RANKS),
                ("flags", c_void_p * MAX_RANKS), ("w_bf16", c_void_p * MAX_RANKS),
                ("mc_grad", c_void_p), ("mc_w", c_void_p), ("mc_w_bf16", c_void_p)]


class _CudaBuffer:
    """Exposes raw device memory through ``__cuda_array_interface__`` so torch can wrap it without copying."""

    def __init__(self, addr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (addr, False), "version": 3, "strides": None}


class SymmetricBuffer:
    """One allocation per rank, mapped into every rank: ``local`` tensor + ``ptrs[j]`` = address of rank j's copy.

    ``mc_ptr`` is the NVLS multicast address of the buffer (0 when the communicator runs without multicast)."""

    def __init__(self, comm: "PeerComm", nbytes: int):
        self.comm, self.nbytes = comm, nbytes
        self.mc_ptr = 0
        self._opened: list[int] = []
        self.base = 0
        self._symm = None
        if comm.nvls:
            self._init_multicast()
        else:
            self._init_ipc()
        self.ptrs_dev = torch.tensor(self.ptrs, dtype=torch.int64, device=comm.device)   # float* const* for the kernels

    def _init_ipc(self) -> None:
        lib, comm = _lib.load(), self.comm
        base = c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        check(lib.dtm_symm_alloc(self.nbytes, ctypes.byref(base), handle), "symm_alloc")
        self.base = base.value
        handles: list = [None] * comm.world_size
        dist.all_gather_object(handles, bytes(handle), group=comm.group)
        self.ptrs: list[int] = []
        for j, h in enumerate(handles):
            if j == comm.rank:
                self.ptrs.append(self.base)
                continue
            peer = c_void_p()
            hb = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            check(lib.dtm_symm_open(hb, ctypes.byref(peer)), f"symm_open(rank {j})")
            self.ptrs.append(peer.value)
            self._opened.append(peer.value)
        self.local_u8 = torch.as_tensor(_CudaBuffer(self.base, self.nbytes), device=comm.device)

    def _init_multicast(self) -> None:
        """Allocation + peer mapping + multicast binding through torch's symmetric-memory plumbing (cuMem* VMM
        allocations, handle exchange over the process group's store, cuMulticast* binding)."""
        import torch.distributed._symmetric_memory as symm
        comm = self.comm
        t = symm.empty(self.nbytes, dtype=torch.uint8, device=comm.device)
        h = symm.rendezvous(t, comm.group)
        ptrs = [int(p) for p in h.buffer_ptrs]
        mc = int(getattr(h, "multicast_ptr", 0) or 0)
        if len(ptrs) != comm.world_size or ptrs[comm.rank] != t.data_ptr() or mc == 0:
            raise RuntimeError("symmetric-memory rendezvous returned no usable multicast mapping")
        t.zero_()
        torch.cuda.synchronize(comm.device)
        self._symm = (t, h)
        self.local_u8, self.ptrs, self.mc_ptr = t, ptrs, mc

    def view(self, dtype: torch.dtype, numel: int | None = None) -> torch.Tensor:
        t = self.local_u8.view(dtype)
        return t if numel is None else t[:numel]

    def close(self) -> None:
        lib = _lib.load()
        for p in self._opened:
            lib.dtm_symm_close(c_void_p(p))
        self._opened = []
        if self.base:
            self.local_u8 = None
            lib.dtm_symm_free(c_void_p(self.base))
            self.base = 0
        self._symm = None


def _nvls_available(group, device) -> bool:
    """Collective decision: can every rank map a multicast buffer?  ``DTM_NVLS=0`` forces the P2P kernels."""
    mode = os.environ.get("DTM_NVLS", "auto").lower()
    if mode in ("0", "off", "false", "no"):
        return False
    ok = 1
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(4096, dtype=torch.uint8, device=device)
        h = symm.rendezvous(t, group)
        if not int(getattr(h, "multicast_ptr", 0) or 0):
            ok = 0
    except Exception:  # noqa: BLE001 - any failure means "use the IPC path"
        ok = 0
    flag = torch.tensor([ok], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    avail = bool(flag.item())
    if not avail and mode in ("1", "on", "true", "yes", "force"):
        raise RuntimeError("DTM_NVLS=1 but NVLS multicast memory is not available on every rank")
    return avail


class PeerComm:
    def __init__(self, group, device, *, overlap_ctas: int | None = None):
        self.group, self.device = group, torch.device(device)
        self.rank, self.world_size = dist.get_rank(group), dist.get_world_size(group)
        if self.world_size > MAX_RANKS:
            raise ValueError(f"PeerComm supports up to {MAX_RANKS} ranks on one NVSwitch node")
        torch.cuda.set_device(self.device)
        # DTM_COMM_WATCHDOG_MS: how long a kernel waits for a peer before it traps with "[dtm watchdog] peer barrier (a rank
        # did not arrive)" (default about two minutes; fault-injection tests use seconds)
        ms = os.environ.get("DTM_COMM_WATCHDOG_MS")
        if ms:
            khz = torch.cuda.get_device_properties(self.device).clock_rate if hasattr(torch.cuda.get_device_properties(self.device), "clock_rate") else 1_965_000
            check(_lib.load().dtm_set_comm_watchdog(int(float(ms) * khz)), "set_comm_watchdog")
        self.nvls = _nvls_available(group, self.device)   # in-switch reduction + multicast stores (multimem.*)
        # grid cap of a fused exchange that overlaps compute kernels.  NVLS kernel: 0 = one small CTA on every SM, co-resident
        # with the conv backward (2 GPUs: 222 -> 189 us per step against a 64-CTA cap); P2P kernel (512 threads, cannot
        # share an SM with conv2_dgrad): 64 CTAs.
        if overlap_ctas is None:
            overlap_ctas = 0 if self.nvls else 64
        self.overlap_ctas = int(os.environ.get("DTM_OVERLAP_CTAS", overlap_ctas))
        self.use_multicast = self.nvls                    # the fused update takes the NVLS path (A/B switch for tools)
        self._buffers: list[SymmetricBuffer] = []
        self.flags = self._alloc(MAX_RANKS * 4 * 16)          # epochs, padded
        self.ctrl = torch.zeros(16, dtype=torch.uint8, device=self.device)
        self.w_buf: SymmetricBuffer | None = None
        self.grad_buf: SymmetricBuffer | None = None
        self.bf16_buf: SymmetricBuffer | None = None
        self.ctx: PeerCtx | None = None
        self._scratch: SymmetricBuffer | None = None
        self._scratch_ctx: PeerCtx | None = None

    # ---- memory ------------------------------------------------------------------------------------
    def _alloc(self, nbytes: int) -> SymmetricBuffer:
        b = SymmetricBuffer(self, (nbytes + 255) // 256 * 256)
        self._buffers.append(b)
        return b

    def flat_allocator(self):
        """Allocator for :class:`FlatParams`: weights and gradients live in symmetric memory, the rest is rank-local."""
        def alloc(n: int, dtype: torch.dtype, name: str = ""):
            if name == "w":
                self.w_buf = self._alloc(n * 4)
                return self.w_buf.view(torch.float32, n)
            if name == "grad":
                self.grad_buf = self._alloc(n * 4)
                return self.grad_buf.view(torch.float32, n)
            if name == "w_bf16":
                self.bf16_buf = self._alloc(n * 2)
                return self.bf16_buf.view(torch.bfloat16, n)
            return torch.zeros(n, dtype=dtype, device=self.device)

        alloc.named = True
        return alloc

    def _make_ctx(self, grad: SymmetricBuffer | None, w: SymmetricBuffer | None, bf16: SymmetricBuffer | None = None) -> PeerCtx:
        c = PeerCtx()
        c.rank, c.world = self.rank, self.world_size
        for j in range(self.world_size):
            c.grad[j] = grad.ptrs[j] if grad else 0
            c.w[j] = w.ptrs[j] if w else 0
            c.w_bf16[j] = bf16.ptrs[j] if bf16 else 0
            c.flags[j] = self.flags.ptrs[j]
        if self.use_multicast and grad and w and bf16 and grad.mc_ptr and w.mc_ptr and bf16.mc_ptr:
            c.mc_grad, c.mc_w, c.mc_w_bf16 = grad.mc_ptr, w.mc_ptr, bf16.mc_ptr
        return c

    def _ensure_ctx(self) -> PeerCtx:
        if self.ctx is None:
            if self.w_buf is None or self.grad_buf is None or self.bf16_buf is None:
                raise RuntimeError("PeerComm: FlatParams must be created with comm.flat_allocator() first")
            self.ctx = self._make_ctx(self.grad_buf, self.w_buf, self.bf16_buf)
            dist.barrier(group=self.group)   # every rank has mapped every buffer before the first kernel touches a peer
        return self.ctx

    # ---- hot path ------------------------------------------------------------------------------------
    def fused_allreduce_update(self, cfg, state, flat, lo: int = 0, hi: int | None = None, *, finish: bool = True,
                               overlapped: bool = False) -> None:
        """Gradient all-reduce + optimizer + EMA + weight all-gather for flat elements [lo, hi) in one kernel."""
        ctx = self._ensure_ctx()
        hi = flat.total if hi is None else hi
        check(_lib.load().dtm_fused_allreduce_update(ctypes.byref(ctx), ctypes.byref(cfg), ptr(state), lo, hi, ptr(flat.slot0),
                                                     ptr(flat.slot1), ptr(flat.ema), ptr(flat.w_bf16), ptr(self.ctrl), int(finish),
                                                     self.overlap_ctas if overlapped else 0, current_stream()),
              "fused_allreduce_update")

    def broadcast_params(self, flat, state=None, root: int = 0) -> None:
        """Make the root's weights (and training state scalars) authoritative on every rank."""
        ctx = self._ensure_ctx()
        check(_lib.load().dtm_broadcast_f32(ctypes.byref(ctx), ptr(self.w_buf.ptrs_dev), 0, flat.total, root, ptr(self.ctrl),
                                            current_stream()), "broadcast")
        flat.w_bf16.copy_(flat.w)
        if flat.ema is not None:
            pass  # EMA shadows / optimizer slots are identical by construction (same init or same checkpoint files)
        if state is not None:
            # global_step & friends: 248 bytes through the same peer-memory broadcast (padded to a float4 multiple)
            n = (state.numel() + 15) // 16 * 4
            buf = self.scratch(n * 4)
            buf.view(torch.uint8)[:state.numel()].copy_(state)
            check(_lib.load().dtm_broadcast_f32(ctypes.byref(self._scratch_ctx), ptr(buf.ptrs_dev), 0, n, root, ptr(self.ctrl),
                                                current_stream()), "broadcast(state)")
            state.copy_(buf.view(torch.uint8)[:state.numel()])
        torch.cuda.synchronize(self.device)

    # ---- stand-alone all-reduce -------------------------------------------------------------------------
    def scratch(self, nbytes: int) -> SymmetricBuffer:
        """Symmetric scratch buffer (collective: all ranks must call with the same size sequence)."""
        if self._scratch is None or self._scratch.nbytes < nbytes:
            self._scratch = self._alloc(max(nbytes, 1 << 20))
            self._scratch_ctx = self._make_ctx(None, None)
            dist.barrier(group=self.group)
        return self._scratch

    def all_reduce(self, x: torch.Tensor, mode: str = "auto", nvls: bool | None = None) -> torch.Tensor:
        """Sum ``x`` (float32, numel % 4 == 0) over ranks through peer memory; returns a new tensor.

        ``mode``: "one_shot" (every rank reads everything), "two_shot" (reduce-scatter + all-gather) or "auto"
        (one-shot up to 256 KiB, where latency dominates; two-shot above).  ``nvls``: reduce inside the NVSwitch
        through the multicast mapping (default: whenever the communicator has one)."""
        if x.dtype != torch.float32 or x.numel() % 4:
            raise ValueError("all_reduce: float32 with a multiple of 4 elements")
        n = x.numel()
        if mode == "auto":
            mode = "one_shot" if n * 4 <= 256 * 1024 else "two_shot"
        buf = self.scratch(n * 4)
        buf.view(torch.float32, n).copy_(x.reshape(-1))
        out = torch.empty_like(x)
        self.all_reduce_inplace(buf, n, out, mode, nvls)
        return out

    def all_reduce_inplace(self, buf: SymmetricBuffer, n: int, out: torch.Tensor, mode: str, nvls: bool | None = None) -> None:
        use_mc = bool(buf.mc_ptr) if nvls is None else bool(nvls)
        if use_mc and not buf.mc_ptr:
            raise RuntimeError("all_reduce: this communicator has no multicast mapping (NVLS unavailable or DTM_NVLS=0)")
        check(_lib.load().dtm_allreduce_f32(ctypes.byref(self._scratch_ctx), ptr(buf.ptrs_dev), c_void_p(buf.mc_ptr if use_mc else None),
                                            0, n, ptr(out), {"one_shot": 0, "two_shot": 1}[mode], ptr(self.ctrl), current_stream()),
              "all_reduce")

    # ---- cold path: assembling sharded optimizer state for checkpoints ---------------------------------------
    def shard_range(self, lo: int, hi: int, rank: int) -> tuple[int, int]:
        per = ((hi - lo + self.world_size - 1) // self.world_size + 63) // 64 * 64
        s_lo = min(hi, lo + rank * per)
        return s_lo, min(hi, s_lo + per)

    def gather_sharded(self, t: torch.Tensor, buckets: list[tuple[int, int]]) -> torch.Tensor:
        """Full copy of a buffer whose elements are only valid on their owning rank (optimizer slots, EMA shadows)."""
        out = t.clone()
        for lo, hi in buckets:
            for r in range(self.world_size):
                a, b = self.shard_range(lo, hi, r)
                if b > a:
                    dist.broadcast(out[a:b], src=dist.get_global_rank(self.group, r) if self.group is not dist.group.WORLD else r,
                                   group=self.group)
        return out

    def barrier(self) -> None:
        dist.barrier(group=self.group)

    def close(self, collective: bool = True) -> None:
        """Unmap and free the symmetric buffers.  ``collective=False`` skips the rendezvous with the other ranks (error paths:
        a rank that is going down must not wait for peers that may already be gone)."""
        torch.cuda.synchronize(self.device)
        if collective:
            try:
                dist.barrier(group=self.group)
            except Exception:
                pass
        for b in self._buffers:
            b.close()
        self._buffers = []
