"""Flat parameter space of the native engine.

All variables of the model live back to back in one fp32 buffer (each offset aligned to 64 elements), mirrored by
equally shaped buffers for the gradient, the optimizer slots, the EMA(0.9999) shadows and the bf16 copy that the
tensor-core kernels read.  One buffer = one message: the gradient exchange that the reference performs as ten
per-variable gRPC pushes and pulls becomes a single fused all-reduce+update over this buffer.
Names, shapes and order are the TensorFlow checkpoint inventory (presets.py), so ``to_dict``/``load_dict`` are the
bridge to the tf.train.Saver-compatible checkpoint files.
"""
from __future__ import annotations

import torch

from .presets import FLAT_ALIGN, MOVING_AVERAGE_DECAY, ModelSpec


def _align(n: int, a: int = FLAT_ALIGN) -> int:
    return (n + a - 1) // a * a


class FlatParams:
    def __init__(self, spec: ModelSpec, device="cuda", *, with_slots: int = 2, with_ema: bool = True, allocator=None):
        self.spec = spec
        self.device = torch.device(device)
        self.offsets: dict[str, int] = {}
        self.index: dict[str, int] = {}
        off = 0
        for i, v in enumerate(spec.variables):
            self.offsets[v.name] = off
            self.index[v.name] = i
            off += _align(v.size)
        self.total = off
        self.blob = None
        if allocator is None and with_slots >= 2 and with_ema:
            # one contiguous allocation [w | grad | slot0 | slot1 | ema | w_bf16]: a single address range that the update
            # kernels can keep in the persisting L2 (native.set_update_l2_window)
            self.blob = torch.zeros(self.total * 22, dtype=torch.uint8, device=self.device)
            order = {"w": 0, "grad": 1, "slot0": 2, "slot1": 3, "ema": 4}

            def allocator(n, dtype, name=""):   # noqa: E306
                if name == "w_bf16":
                    return self.blob[self.total * 20:].view(torch.bfloat16)
                k = order[name]
                return self.blob[self.total * 4 * k:self.total * 4 * (k + 1)].view(torch.float32)
        alloc = allocator or (lambda n, dtype, name="": torch.zeros(n, dtype=dtype, device=self.device))
        self.w = alloc(self.total, torch.float32, "w")
        self.grad = alloc(self.total, torch.float32, "grad")
        self.slot0 = alloc(self.total, torch.float32, "slot0") if with_slots >= 1 else None
        self.slot1 = alloc(self.total, torch.float32, "slot1") if with_slots >= 2 else None
        self.ema = alloc(self.total, torch.float32, "ema") if with_ema else None
        self.w_bf16 = alloc(self.total, torch.bfloat16, "w_bf16")

    # ---- views ------------------------------------------------------------------------------------
    def _view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        v = self.spec.var(name)
        o = self.offsets[name]
        return buf[o:o + v.size].view(v.shape)

    def weight(self, name: str) -> torch.Tensor:
        return self._view(self.w, name)

    def weight_bf16(self, name: str) -> torch.Tensor:
        return self._view(self.w_bf16, name)

    def gradient(self, name: str) -> torch.Tensor:
        return self._view(self.grad, name)

    def _dict(self, buf) -> dict[str, torch.Tensor]:
        return {v.name: self._view(buf, v.name).clone() for v in self.spec.variables}

    def to_dict(self):
        return self._dict(self.w)

    def grad_dict(self):
        return self._dict(self.grad)

    def ema_dict(self):
        return self._dict(self.ema)

    def bf16_dict(self):
        return self._dict(self.w_bf16)

    def slot_dict(self, k: int):
        return self._dict(self.slot0 if k == 0 else self.slot1)

    # ---- loading ----------------------------------------------------------------------------------
    def _load(self, buf: torch.Tensor, tensors, strict=True) -> None:
        for v in self.spec.variables:
            if v.name not in tensors:
                if strict:
                    raise KeyError(f"missing variable {v.name}")
                continue
            t = torch.as_tensor(tensors[v.name])
            if tuple(t.shape) != tuple(v.shape):
                raise ValueError(f"{v.name}: shape {tuple(t.shape)} != expected {tuple(v.shape)}")
            self._view(buf, v.name).copy_(t.to(buf.dtype))

    def load_dict(self, tensors, *, reset_ema: bool = True) -> None:
        """Set the master weights (and refresh the bf16 shadow; EMA shadows restart from the weights like TF's)."""
        self._load(self.w, tensors)
        self.w_bf16.copy_(self.w)
        if reset_ema and self.ema is not None:
            self.ema.copy_(self.w)

    def load_grad_dict(self, tensors) -> None:
        self._load(self.grad, tensors)

    def load_ema_dict(self, tensors) -> None:
        self._load(self.ema, tensors, strict=False)

    def load_slot_dict(self, k: int, tensors) -> None:
        self._load(self.slot0 if k == 0 else self.slot1, tensors, strict=False)

    # ---- kernel configuration ------------------------------------------------------------------------
    def update_config(self, *, optimizer="adam", base_lr=1e-3, decay_rate=0.1, decay_steps=0, momentum=0.9, beta1=0.9,
                      beta2=0.999, epsilon=1e-8, ema_decay=MOVING_AVERAGE_DECAY):
        from ..ops import native

        kinds = {"sgd": native.OPT_SGD, "momentum": native.OPT_MOMENTUM, "adam": native.OPT_ADAM}
        if optimizer.lower() not in kinds:
            raise ValueError(f"unknown optimizer {optimizer!r} (sgd | momentum | adam)")
        if len(self.spec.variables) > native.MAX_SEGMENTS:
            raise ValueError("too many variables for the update kernel")
        cfg = native.UpdateConfig()
        cfg.optimizer = kinds[optimizer.lower()]
        cfg.n_segments = len(self.spec.variables)
        cfg.total = self.total
        cfg.decay_steps = int(decay_steps)
        cfg.base_lr, cfg.decay_rate, cfg.momentum = base_lr, decay_rate, momentum
        cfg.beta1, cfg.beta2, cfg.epsilon, cfg.ema_decay = beta1, beta2, epsilon, ema_decay
        for i, v in enumerate(self.spec.variables):
            cfg.seg_offset[i] = self.offsets[v.name]
            cfg.seg_wd[i] = float(v.wd or 0.0)
            cfg.seg_zero_grad[i] = 1 if v.atomic_grad else 0
            cfg.seg_gather_bf16[i] = 1 if v.tensor_core_only else 0
        cfg.seg_offset[len(self.spec.variables)] = self.total
        cfg.pack_seg = -1
        return cfg
