"""Checkpoint save / restore in the tf.train.Saver layout (what ``MonitoredTrainingSession(checkpoint_dir=...)`` does
for the reference, main.py:140-146, and what ``saver.restore`` consumes, inference.py:86-91).

Variable inventory written:
  global_step (int64 scalar) · the model variables · ``<var>/ExponentialMovingAverage`` · optimizer slots
  (``<var>/Momentum`` | ``<var>/Adam``, ``<var>/Adam_1``, ``beta1_power``, ``beta2_power``) · the loss averages
  ``<layer>/weight_loss/avg``, ``cross_entropy/avg``, ``total_loss/avg``.
Data shards mirror TensorFlow's default sharded Saver: one shard per parameter-server task, variables assigned
round-robin in creation order like ``replica_device_setter`` (main.py:80-82), derived variables co-located.
"""
from __future__ import annotations

import glob
import os
import time

import numpy as np
import torch

from ..formats import checkpoint_state as cs
from ..formats import graph_def as gd
from ..formats import tensor_bundle as tb
from ..models.presets import ModelSpec

EMA_SUFFIX = "/ExponentialMovingAverage"
SLOT_NAMES = {"sgd": [], "momentum": ["Momentum"], "adam": ["Adam", "Adam_1"]}


def loss_avg_names(spec: ModelSpec) -> list[tuple[str, int]]:
    """(checkpoint name, index into the engine's loss_avg vector) for the seven-ish loss moving averages."""
    from ..ops.native import MAX_SEGMENTS

    out = [(f"{v.layer}/weight_loss/avg", i) for i, v in enumerate(spec.variables) if v.wd is not None]
    out += [("cross_entropy/avg", MAX_SEGMENTS), ("total_loss/avg", MAX_SEGMENTS + 1)]
    return out


def _engine_loss_avg(engine, spec: ModelSpec) -> list[float]:
    """Loss averages as a MAX_SEGMENTS+2 vector regardless of the engine flavour."""
    from ..ops.native import MAX_SEGMENTS

    vec = [0.0] * (MAX_SEGMENTS + 2)
    if hasattr(engine, "read_state"):
        st = engine.read_state()
        return [float(v) for v in st.loss_avg]
    la = engine.scalar_state()["loss_avg"]
    n = len(spec.variables)
    for i in range(n):
        vec[i] = la[i]
    vec[MAX_SEGMENTS], vec[MAX_SEGMENTS + 1] = la[n], la[n + 1]
    return vec


def _scalars(engine) -> dict:
    if hasattr(engine, "read_state"):
        st = engine.read_state()
        return {"global_step": int(st.global_step), "beta1_power": float(st.beta1_power), "beta2_power": float(st.beta2_power)}
    s = engine.scalar_state()
    return {"global_step": int(s["global_step"]), "beta1_power": float(s["beta1_power"]), "beta2_power": float(s["beta2_power"])}


def collect_variables(engine, spec: ModelSpec, optimizer: str) -> dict[str, np.ndarray]:
    """Everything a TF training graph of this model would save, keyed by TensorFlow variable name."""
    out: dict[str, np.ndarray] = {}
    sc = _scalars(engine)
    out["global_step"] = np.array(sc["global_step"], dtype=np.int64)
    for k, v in engine.params().items():
        out[k] = v.numpy().astype(np.float32)
    for k, v in engine.ema_params().items():
        out[k + EMA_SUFFIX] = v.numpy().astype(np.float32)
    for slot_name, d in zip(SLOT_NAMES[optimizer.lower()], engine.slots()):
        for k, v in d.items():
            out[f"{k}/{slot_name}"] = v.numpy().astype(np.float32)
    if optimizer.lower() == "adam":
        out["beta1_power"] = np.array(sc["beta1_power"], dtype=np.float32)
        out["beta2_power"] = np.array(sc["beta2_power"], dtype=np.float32)
    la = _engine_loss_avg(engine, spec)
    for name, idx in loss_avg_names(spec):
        out[name] = np.array(la[idx], dtype=np.float32)
    return out


def variable_inventory(spec: ModelSpec, optimizer: str) -> dict[str, tuple]:
    """``name -> (dtype, shape)`` of everything :func:`collect_variables` returns, derived from the model spec alone (no
    engine, no collective): what ``graph.pbtxt`` lists when the chief starts."""
    out: dict[str, tuple] = {"global_step": (np.int64, ())}
    for v in spec.variables:
        out[v.name] = (np.float32, tuple(v.shape))
    for v in spec.variables:
        out[v.name + EMA_SUFFIX] = (np.float32, tuple(v.shape))
    for slot_name in SLOT_NAMES[optimizer.lower()]:
        for v in spec.variables:
            out[f"{v.name}/{slot_name}"] = (np.float32, tuple(v.shape))
    if optimizer.lower() == "adam":
        out["beta1_power"] = (np.float32, ())
        out["beta2_power"] = (np.float32, ())
    for name, _ in loss_avg_names(spec):
        out[name] = (np.float32, ())
    return out


def write_graph(train_dir: str, spec: ModelSpec, optimizer: str) -> str:
    """``<train_dir>/graph.pbtxt`` — written once by the chief at session start, like MonitoredTrainingSession."""
    return gd.write_graph_pbtxt(train_dir, variable_inventory(spec, optimizer))


def shard_assignment(names: list[str], spec: ModelSpec, num_shards: int) -> dict[str, int]:
    """Round-robin placement over ps tasks in variable creation order (global_step first), like replica_device_setter."""
    if num_shards <= 1:
        return {n: 0 for n in names}
    order = ["global_step"] + [v.name for v in spec.variables]
    primary = {n: i % num_shards for i, n in enumerate(order)}
    out = {}
    nxt = len(order)
    for n in names:
        if n in primary:
            out[n] = primary[n]
            continue
        base = n.rsplit("/", 1)[0]
        if base in primary:           # EMA shadow / optimizer slot: co-located with its variable
            out[n] = primary[base]
        else:                         # free-standing scalars (beta powers, loss averages): next round-robin slot
            out[n] = nxt % num_shards
            nxt += 1
    return out


def checkpoint_prefix(train_dir: str, step: int) -> str:
    return os.path.join(train_dir, f"model.ckpt-{step}")


def save_checkpoint(train_dir: str, engine, spec: ModelSpec, optimizer: str, *, num_shards: int = 1, max_to_keep: int = 5,
                    write: bool = True) -> str:
    """Write ``model.ckpt-<global_step>`` + update the ``checkpoint`` state file; returns the prefix.

    With a sharded engine (fused multi-GPU update) collecting the variables is a collective: every rank calls this,
    only the chief passes ``write=True``."""
    variables = collect_variables(engine, spec, optimizer)
    step = int(variables["global_step"])
    prefix = checkpoint_prefix(train_dir, step)
    if not write:
        return prefix
    tb.write_bundle(prefix, variables, shard_assignment(list(variables), spec, num_shards), num_shards)
    gd.write_meta(prefix, gd.describe(variables), max_to_keep=max_to_keep, sharded=True)   # Saver(write_meta_graph=True)
    st = cs.read_state(train_dir) or cs.CheckpointState()
    rel = os.path.basename(prefix)
    paths = [p for p in st.all_model_checkpoint_paths if p != rel] + [rel]
    while max_to_keep and len(paths) > max_to_keep:
        old = paths.pop(0)
        old_prefix = old if os.path.isabs(old) else os.path.join(train_dir, old)
        for f in glob.glob(glob.escape(old_prefix) + ".*"):
            try:
                os.remove(f)
            except OSError:
                pass
    cs.write_state(train_dir, cs.CheckpointState(rel, paths))
    return prefix


def resolve_checkpoint(path: str) -> str | None:
    """Accept a checkpoint directory (uses its ``checkpoint`` state file) or a prefix; returns a prefix or None."""
    if os.path.isdir(path):
        return cs.latest_checkpoint(path)
    if path.endswith(".index"):
        path = path[: -len(".index")]
    return path if tb.checkpoint_exists(path) else None


def restore_checkpoint(path: str, engine, spec: ModelSpec, optimizer: str, *, model_only: bool = False) -> int | None:
    """Restore the newest checkpoint under ``path`` into ``engine``; returns its global_step or None if there is none.

    ``model_only`` loads just the model variables by name — the reference's inference script does exactly that
    (``tf.train.Saver()`` over the ten variables, inference.py:86).  A checkpoint whose conv1 kernel has a different
    number of input channels than the model is adapted (reference quirk B1: 3-channel training records vs. 1-channel
    inference images): 3 -> 1 sums the channel kernels (equivalent for replicated grey input), 1 -> 3 replicates /3.
    """
    prefix = resolve_checkpoint(path)
    if prefix is None:
        return None
    rd = tb.BundleReader(prefix)
    model = {}
    for v in spec.variables:
        if v.name not in rd:
            raise KeyError(f"checkpoint {prefix} lacks variable {v.name!r}")
        arr = rd.get(v.name)
        if v.name == "conv1/weights" and arr.shape != v.shape and arr.shape[:2] == (5, 5) and arr.shape[3] == 32:
            if arr.shape[2] == 3 and v.shape[2] == 1:
                arr = arr.sum(axis=2, keepdims=True)
            elif arr.shape[2] == 1 and v.shape[2] == 3:
                arr = np.repeat(arr, 3, axis=2) / 3.0
        if tuple(arr.shape) != tuple(v.shape):
            raise ValueError(f"{v.name}: checkpoint shape {arr.shape} != model shape {v.shape}")
        model[v.name] = torch.from_numpy(arr.astype(np.float32))
    step = int(rd.get("global_step")) if "global_step" in rd else 0
    if model_only:
        engine.load_params(model, global_step=step)
        return step
    ema = {v.name: torch.from_numpy(rd.get(v.name + EMA_SUFFIX)) for v in spec.variables
           if v.name + EMA_SUFFIX in rd and tuple(rd.entry(v.name + EMA_SUFFIX).shape) == tuple(v.shape)}
    slots = []
    for slot_name in SLOT_NAMES[optimizer.lower()]:
        slots.append({v.name: torch.from_numpy(rd.get(f"{v.name}/{slot_name}")) for v in spec.variables
                      if f"{v.name}/{slot_name}" in rd and tuple(rd.entry(f"{v.name}/{slot_name}").shape) == tuple(v.shape)})
    from ..ops.native import MAX_SEGMENTS

    la = [0.0] * (MAX_SEGMENTS + 2)
    for name, idx in loss_avg_names(spec):
        if name in rd:
            la[idx] = float(rd.get(name))
    if not hasattr(engine, "read_state"):   # TorchEngine keeps a compact vector
        n = len(spec.variables)
        la = la[:n] + [la[MAX_SEGMENTS], la[MAX_SEGMENTS + 1]]
    engine.load_params(model, ema=ema or None, slots=slots or None, global_step=step, loss_avg=la)
    return step


class CheckpointSaver:
    """Time-based saver with the defaults of MonitoredTrainingSession: every 600 s and at the end (chief only)."""

    def __init__(self, train_dir: str, spec: ModelSpec, optimizer: str, *, save_secs: float = 600.0, num_shards: int = 1,
                 max_to_keep: int = 5):
        self.train_dir, self.spec, self.optimizer = train_dir, spec, optimizer
        self.save_secs, self.num_shards, self.max_to_keep = save_secs, num_shards, max_to_keep
        self._last = time.time()
        self.saved: list[str] = []

    def due(self) -> bool:
        return self.save_secs is not None and time.time() - self._last >= self.save_secs

    def save(self, engine, write: bool = True) -> str:
        self._last = time.time()
        p = save_checkpoint(self.train_dir, engine, self.spec, self.optimizer, num_shards=self.num_shards,
                            max_to_keep=self.max_to_keep, write=write)
        if write:
            self.saved.append(p)
        return p

    def maybe_save(self, engine, force: bool = False) -> str | None:
        if not force and not self.due():
            return None
        return self.save(engine)
