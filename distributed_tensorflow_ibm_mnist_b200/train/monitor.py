"""``monitor_cb`` stand-in: the chief's TensorBoard summaries.

The reference registers, through an IBM platform module that is not in its repository, histogram / norm / gradient
summaries for the layers conv1, conv2, local3, local4 and four scalars (main.py:95-109) and merges them into a
"train" and a "test" collection.  :class:`CMonitor` keeps the five registration methods and the two collection keys;
``collect()`` turns the registrations into a serialised ``Summary`` proto from an engine's current tensors.
On CUDA tensors the statistics run in the sm_100a stats kernels (csrc/stats), so a summary costs a few hundred bytes
of PCIe traffic instead of the full weight set.
"""
from __future__ import annotations

import numpy as np
import torch

from ..formats import events

DLMAO_TRAIN_SUMMARIES = "dlmao_train_summaries"
DLMAO_TEST_SUMMARIES = "dlmao_test_summaries"

_LIMITS_DEV: dict = {}
_SCRATCH: dict = {}


def _hist_proto_from_counts(stats, counts: np.ndarray) -> bytes:
    from ..formats import proto

    limits_all = events._BUCKET_LIMITS
    limits, buckets = [], []
    i, n = 0, len(counts)
    while i < n:
        end, count = limits_all[i], counts[i]
        i += 1
        if count <= 0:
            while i < n and counts[i] <= 0:
                end, count = limits_all[i], counts[i]
                i += 1
        limits.append(float(end))
        buckets.append(float(count))
    mn, mx, sm, sq, num = (float(stats[k]) for k in range(5))
    out = proto.f_double(1, mn) + proto.f_double(2, mx) + proto.f_double(3, num) + proto.f_double(4, sm) + proto.f_double(5, sq)
    return out + proto.f_packed_doubles(6, limits) + proto.f_packed_doubles(7, buckets)


def tensor_summary_stats(t: torch.Tensor) -> tuple[np.ndarray, np.ndarray]:
    """(stats[6] = min, max, sum, sumsq, count, l2norm ; histogram counts over TF's default buckets)."""
    if t.is_cuda and t.dtype in (torch.float32, torch.bfloat16):
        from ..ops import native

        dev = t.device
        if dev not in _LIMITS_DEV:
            _LIMITS_DEV[dev] = torch.from_numpy(events._BUCKET_LIMITS).to(dev)
        need = native.tensor_stats_scratch_bytes(t.numel())
        if dev not in _SCRATCH or _SCRATCH[dev].numel() < need:
            _SCRATCH[dev] = torch.zeros(max(need, 64 + 1024 * 32), dtype=torch.uint8, device=dev)
        out = torch.zeros(6, device=dev)
        counts = torch.zeros(_LIMITS_DEV[dev].numel(), device=dev)
        native.tensor_stats(t, out, _SCRATCH[dev])
        native.histogram(t, _LIMITS_DEV[dev], counts)
        return out.cpu().numpy().astype(np.float64), counts.cpu().numpy().astype(np.float64)
    v = t.detach().to(torch.float64).cpu().numpy().reshape(-1)
    stats = np.array([v.min(), v.max(), v.sum(), (v * v).sum(), v.size, np.sqrt((v * v).sum())])
    idx = np.minimum(np.searchsorted(events._BUCKET_LIMITS, v, side="right"), len(events._BUCKET_LIMITS) - 1)
    return stats, np.bincount(idx, minlength=len(events._BUCKET_LIMITS)).astype(np.float64)


class CMonitor:
    """``monitor_cb.CMonitor(log_dir, test_interval, max_steps)`` with the registration surface of main.py:97-107."""

    def __init__(self, log_dir: str, test_interval: int, max_steps: int):
        self.log_dir, self.test_interval, self.max_steps = log_dir, test_interval, max_steps
        self._hists: list[tuple[str, str, str]] = []      # (kind, tensor key, layer)
        self._norms: list[tuple[str, str, str]] = []
        self._gradients: str | None = None
        self._gw_ratio = False
        self._scalars: dict[str, list[tuple[str, str]]] = {DLMAO_TRAIN_SUMMARIES: [], DLMAO_TEST_SUMMARIES: []}

    # ---- registration (reference names) ----------------------------------------------------------------
    @staticmethod
    def _key(tensor_name) -> str:
        name = str(tensor_name)
        return name[:-2] if name.endswith(":0") else name

    def SummaryHist(self, kind: str, tensor, layer: str) -> None:  # noqa: N802
        self._hists.append((kind, self._key(tensor), layer))

    def SummaryNorm2(self, kind: str, tensor, layer: str) -> None:  # noqa: N802
        self._norms.append((kind, self._key(tensor), layer))

    def SummaryGradient(self, kind: str, loss=None) -> None:  # noqa: N802
        self._gradients = kind

    def SummaryGWRatio(self) -> None:  # noqa: N802
        self._gw_ratio = True

    def SummaryScalar(self, name: str, tensor) -> None:  # noqa: N802
        coll = DLMAO_TEST_SUMMARIES if "test" in name.lower() else DLMAO_TRAIN_SUMMARIES
        self._scalars[coll].append((name, str(tensor)))

    # ---- evaluation ----------------------------------------------------------------------------------
    def _lookup(self, key: str, tensors: dict) -> torch.Tensor | None:
        if key in tensors["params"]:
            return tensors["params"][key]
        parts = key.split("/")
        if len(parts) == 2 and parts[0] == parts[1]:          # "<layer>/<layer>" = the layer's activation
            return tensors["activations"].get(parts[0])
        return tensors["activations"].get(key)

    def collect(self, collection: str, tensors: dict, scalars: dict[str, float]) -> bytes:
        """Serialised Summary for one collection.

        ``tensors`` = {"params": {name: tensor}, "activations": {layer: tensor}, "gradients": {name: tensor}};
        ``scalars`` = {"loss": .., "accuracy": ..} of the batch the collection refers to.
        """
        out = b""
        for name, source in self._scalars[collection]:
            key = "accuracy" if "accuracy" in name.lower() else "loss"
            if key in scalars:
                out += events.scalar_value(name, scalars[key])
        if collection != DLMAO_TRAIN_SUMMARIES:
            return out
        for kind, key, layer in self._hists:
            t = self._lookup(key, tensors)
            if t is None:
                continue
            stats, counts = tensor_summary_stats(t)
            out += events.proto.f_bytes(1, events.proto.f_str(1, f"{layer}/{kind}") +
                                        events.proto.f_bytes(5, _hist_proto_from_counts(stats, counts)))
        norms: dict[str, float] = {}
        for kind, key, layer in self._norms:
            t = self._lookup(key, tensors)
            if t is None:
                continue
            stats, _ = tensor_summary_stats(t) if t.is_cuda else (None, None)
            norm = float(stats[5]) if stats is not None else float(t.detach().double().norm())
            norms[key] = norm
            out += events.scalar_value(f"{layer}/{kind}_norm2", norm)
        grads = tensors.get("gradients") or {}
        if self._gradients and grads:
            for name, g in grads.items():
                if not name.endswith("weights"):
                    continue
                stats, counts = tensor_summary_stats(g)
                layer = name.split("/")[0]
                out += events.proto.f_bytes(1, events.proto.f_str(1, f"{layer}/{self._gradients}_gradient") +
                                            events.proto.f_bytes(5, _hist_proto_from_counts(stats, counts)))
                if self._gw_ratio:
                    w = tensors["params"].get(name)
                    wn = norms.get(name) or (float(w.detach().double().norm()) if w is not None else 0.0)
                    out += events.scalar_value(f"{layer}/gradient_weight_ratio", float(stats[5]) / wn if wn else 0.0)
        return out
