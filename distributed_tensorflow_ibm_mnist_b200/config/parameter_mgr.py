"""``tf_parameter_mgr`` stand-in: hyper-parameters and dataset locations.

In the reference these come from an IBM DLI platform module that is not part of the repository; only its call sites
exist (main.py:38-40, :77, :137; mnist_input.py:10, :22-23, :47, :50, :261; inference.py:63).  Here they are read from,
in order of precedence: explicit ``configure(...)`` calls / CLI flags, the JSON file named by ``$DTM_PARAMS`` (or
``--params_file``), built-in defaults.  The getter names are the reference's.
"""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass, field


@dataclass
class HyperParams:
    max_steps: int = 1000
    test_interval: int = 100
    train_batch_size: int = 128
    base_learning_rate: float = 0.1
    learning_rate_decay: float = 0.1
    optimizer: str = "sgd"            # sgd | momentum | adam
    momentum: float = 0.9
    train_data: list[str] = field(default_factory=list)   # TFRecord files
    test_data: list[str] = field(default_factory=list)
    val_data: list[str] = field(default_factory=list)
    model: str = "reference"          # reference | baseline | wide
    channels: int | None = None       # None -> preset default (3 for reference, 1 otherwise)
    synthetic: bool = True            # no data files -> synthetic MNIST-shaped batches
    seed: int = 0
    save_checkpoint_secs: float = 600.0
    max_to_keep: int = 5


_PARAMS = HyperParams()
_LOADED = False


def _load_file_once() -> None:
    global _LOADED
    if _LOADED:
        return
    _LOADED = True
    path = os.environ.get("DTM_PARAMS")
    if path and os.path.exists(path):
        load_file(path)


def load_file(path: str) -> None:
    with open(path, "r", encoding="utf-8") as f:
        data = json.load(f)
    configure(**data)


def configure(**kw) -> HyperParams:
    """Override hyper-parameters (unknown keys raise). Returns the live object."""
    global _LOADED
    _LOADED = True if not os.environ.get("DTM_PARAMS") else _LOADED
    for k, v in kw.items():
        if not hasattr(_PARAMS, k):
            raise KeyError(f"unknown hyper-parameter {k!r}; known: {', '.join(asdict(_PARAMS))}")
        if k in ("train_data", "test_data", "val_data") and isinstance(v, str):
            v = [p for p in v.split(",") if p]
        setattr(_PARAMS, k, v)
    return _PARAMS


def params() -> HyperParams:
    _load_file_once()
    return _PARAMS


def reset() -> None:
    global _PARAMS, _LOADED
    _PARAMS = HyperParams()
    _LOADED = False


# ---- the reference's getter surface (SURVEY C-X1) ------------------------------------------------------
def getMaxSteps() -> int:  # noqa: N802
    return int(params().max_steps)


def getTestInterval() -> int:  # noqa: N802
    return int(params().test_interval)


def getTrainBatchSize() -> int:  # noqa: N802
    return int(params().train_batch_size)


def getLearningRateDecay() -> float:  # noqa: N802
    return float(params().learning_rate_decay)


def getBaseLearningRate() -> float:  # noqa: N802
    return float(params().base_learning_rate)


def getTrainData() -> list[str]:  # noqa: N802
    return list(params().train_data)


def getTestData() -> list[str]:  # noqa: N802
    return list(params().test_data)


def getValData() -> list[str]:  # noqa: N802
    return list(params().val_data)


@dataclass
class OptimizerSpec:
    """What ``getOptimizer(lr)`` hands to the trainer: the optimizer family and its constants.

    ``learning_rate`` may be a float or a callable step -> lr (the staircase schedule of mnist_input.py:252-256)."""
    kind: str
    learning_rate: object
    momentum: float = 0.9
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8

    def slot_names(self) -> list[str]:
        """TensorFlow slot-variable suffixes (checkpoint names ``<var>/<slot>``)."""
        return {"sgd": [], "momentum": ["Momentum"], "adam": ["Adam", "Adam_1"]}[self.kind]


def getOptimizer(lr) -> OptimizerSpec:  # noqa: N802
    p = params()
    kind = p.optimizer.lower()
    if kind not in ("sgd", "momentum", "adam"):
        raise ValueError(f"unknown optimizer {p.optimizer!r} (sgd | momentum | adam)")
    return OptimizerSpec(kind, lr, momentum=p.momentum)
