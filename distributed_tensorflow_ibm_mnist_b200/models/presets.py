"""Model presets and the variable inventory of the MNIST CNN.

``reference`` reproduces the graph of reference mnist_input.py:118-208 exactly (variable scopes/names, shapes,
initialisers, weight-decay factors, LRN placement); ``baseline`` is the conv32/conv64/fc1024/dropout/fc10 network
of the north-star configs (BASELINE.json), ``wide`` its communication-bound fc-8192 variant.  Variable names are the
TensorFlow checkpoint keys, so checkpoints round-trip with the reference's Saver.
"""
from __future__ import annotations

from dataclasses import dataclass, field

IMAGE_SIZE = 28
NUM_CLASSES = 10
NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 50000   # mnist_input.py:18
NUM_EPOCHS_PER_DECAY = 350.0               # mnist_input.py:21
MOVING_AVERAGE_DECAY = 0.9999              # mnist_input.py:20
FLAT_ALIGN = 64                            # flat-buffer alignment of every variable (elements)


@dataclass(frozen=True)
class VarSpec:
    name: str                 # TensorFlow variable name, e.g. "conv1/weights"
    shape: tuple[int, ...]
    stddev: float | None      # truncated-normal stddev (None -> constant init)
    const: float = 0.0
    wd: float | None = None   # weight-decay factor of _variable_with_weight_decay (None for biases)
    atomic_grad: bool = True  # gradient accumulated with red.add (cleared by the update kernel) vs. stored by a GEMM

    @property
    def size(self) -> int:
        n = 1
        for d in self.shape:
            n *= d
        return n

    @property
    def layer(self) -> str:
        return self.name.split("/")[0]

    @property
    def tensor_core_only(self) -> bool:
        """The big fc matrices are read by the GEMM kernels only (as bf16): with several GPUs their fp32 master stays on
        the owning rank and only bf16 travels over NVLink."""
        return not self.atomic_grad


@dataclass(frozen=True)
class ModelSpec:
    preset: str
    channels: int
    lrn: bool
    fc1: int
    fc2: int                  # 0 -> no second hidden layer
    keep_prob: float          # dropout keep probability on the last hidden activation (1.0 -> none)
    variables: tuple[VarSpec, ...] = field(default_factory=tuple)

    @property
    def head_in(self) -> int:
        return self.fc2 if self.fc2 else self.fc1

    @property
    def image_pixels(self) -> int:
        return IMAGE_SIZE * IMAGE_SIZE * self.channels

    def var(self, name: str) -> VarSpec:
        for v in self.variables:
            if v.name == name:
                return v
        raise KeyError(name)

    @property
    def num_params(self) -> int:
        return sum(v.size for v in self.variables)

    @property
    def layers(self) -> list[str]:
        out: list[str] = []
        for v in self.variables:
            if v.layer not in out:
                out.append(v.layer)
        return out


def make_spec(preset: str = "baseline", channels: int | None = None, fc: int | None = None) -> ModelSpec:
    preset = preset.lower()
    if preset not in ("reference", "baseline", "wide"):
        raise ValueError(f"unknown model preset {preset!r} (reference | baseline | wide)")
    if channels is None:
        channels = 3 if preset == "reference" else 1
    if channels not in (1, 3):
        raise ValueError("channels must be 1 or 3")
    fc1 = fc or (8192 if preset == "wide" else 1024)
    if fc1 % 64 != 0:
        raise ValueError("fc width must be a multiple of 64")
    fc2 = 192 if preset == "reference" else 0
    v = [
        VarSpec("conv1/weights", (5, 5, channels, 32), 5e-2, wd=0.0),
        VarSpec("conv1/biases", (32,), None, 0.0),
        VarSpec("conv2/weights", (5, 5, 32, 64), 5e-2, wd=0.0),
        VarSpec("conv2/biases", (64,), None, 0.1),
        VarSpec("local3/weights", (3136, fc1), 0.04, wd=0.004 if preset == "reference" else 0.0, atomic_grad=False),
        VarSpec("local3/biases", (fc1,), None, 0.1),
    ]
    if fc2:
        v += [
            VarSpec("local4/weights", (fc1, fc2), 0.04, wd=0.004, atomic_grad=False),
            VarSpec("local4/biases", (fc2,), None, 0.1),
        ]
    head_in = fc2 or fc1
    v += [
        VarSpec("softmax_linear/weights", (head_in, NUM_CLASSES), 1.0 / head_in, wd=0.0),
        VarSpec("softmax_linear/biases", (NUM_CLASSES,), None, 0.0),
    ]
    return ModelSpec(preset, channels, preset == "reference", fc1, fc2, 1.0 if preset == "reference" else 0.5, tuple(v))


def decay_steps(batch_size: int) -> int:
    """``int(NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN / batch_size * NUM_EPOCHS_PER_DECAY)`` — mnist_input.py:248-249."""
    return int(NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN / batch_size * NUM_EPOCHS_PER_DECAY)
