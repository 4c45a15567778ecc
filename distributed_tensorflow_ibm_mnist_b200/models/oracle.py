"""Plain-PyTorch statement of the model math: the CPU execution path and the numerical oracle of every kernel.

Mirrors reference mnist_input.py: ``inference`` (:118-208), ``loss`` (:210-235), ``accuracy`` (:237-243),
``train`` (:246-273).  Works on any torch device in fp32; with ``emulate_bf16=True`` tensors are rounded to
bf16 at the same points where the sm_100a path stores bf16, which lets GPU tests use tight tolerances.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .presets import MOVING_AVERAGE_DECAY, ModelSpec

LRN_RADIUS, LRN_BIAS, LRN_ALPHA, LRN_BETA = 4, 1.0, 0.001 / 9.0, 0.75


def tf_lrn(x: torch.Tensor) -> torch.Tensor:
    """``tf.nn.lrn(x, 4, bias=1.0, alpha=0.001/9, beta=0.75)`` on an NCHW tensor.

    TensorFlow does *not* divide alpha by the window size (torch's ``local_response_norm`` does).
    """
    sq = (x * x).unsqueeze(1)
    n = 2 * LRN_RADIUS + 1
    sq = F.pad(sq, (0, 0, 0, 0, LRN_RADIUS, LRN_RADIUS))
    s = F.avg_pool3d(sq, (n, 1, 1), stride=1).squeeze(1) * n
    return x * (LRN_BIAS + LRN_ALPHA * s) ** (-LRN_BETA)


def truncated_normal_(t: torch.Tensor, stddev: float, generator: torch.Generator | None = None) -> torch.Tensor:
    """TF ``truncated_normal_initializer``: N(0, stddev) re-drawn outside two standard deviations."""
    with torch.no_grad():
        t.normal_(0.0, 1.0, generator=generator)
        bad = t.abs() > 2.0
        while bool(bad.any()):
            fresh = torch.empty_like(t).normal_(0.0, 1.0, generator=generator)
            t[bad] = fresh[bad]
            bad = t.abs() > 2.0
        t.mul_(stddev)
    return t


def init_params(spec: ModelSpec, seed: int = 0, device="cpu") -> dict[str, torch.Tensor]:
    """Fresh variables with the reference initialisers (mnist_input.py:137-141, :157-162, :181-183, :189-191, :201-204)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    out: dict[str, torch.Tensor] = {}
    for v in spec.variables:
        t = torch.empty(v.shape, dtype=torch.float32)
        if v.stddev is None:
            t.fill_(v.const)
        else:
            truncated_normal_(t, v.stddev, g)
        out[v.name] = t.to(device)
    return out


def _q(x: torch.Tensor, on: bool) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32) if on else x


class _RoundGradToBf16(torch.autograd.Function):
    """Identity whose backward rounds the incoming gradient to bf16 - placed where the sm_100a backward pass stores a
    gradient tensor as bf16 (dH, dP2 / dZ2, dPool1), so that the oracle's gradients carry the same quantisation."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(g.dtype)


def _qg(x: torch.Tensor, on: bool) -> torch.Tensor:
    return _RoundGradToBf16.apply(x) if (on and x.requires_grad) else x


def normalise_u8(images_u8: torch.Tensor) -> torch.Tensor:
    """``tf.cast(image, tf.float32) * (1. / 255) - 0.5`` — mnist_input.py:39."""
    return images_u8.to(torch.float32) * (1.0 / 255.0) - 0.5


def forward(params: dict[str, torch.Tensor], images: torch.Tensor, spec: ModelSpec, *, dropout_mask: torch.Tensor | None = None,
            emulate_bf16: bool = False, return_activations: bool = False):
    """Logits of the CNN. ``images``: float [B, 28*28*C] already normalised (HWC order, like the TFRecord payload).

    ``dropout_mask`` (bool [B, hidden], True = keep) enables dropout on the last hidden activation (training only).
    """
    q = emulate_bf16
    B = images.shape[0]
    x = _q(images, q).view(B, 28, 28, spec.channels).permute(0, 3, 1, 2)
    acts: dict[str, torch.Tensor] = {}
    w1 = _q(params["conv1/weights"], q).permute(3, 2, 0, 1)
    a1 = F.relu(F.conv2d(x, w1, params["conv1/biases"], padding=2))
    acts["conv1"] = a1
    p1 = _qg(F.max_pool2d(a1, 2), q)          # conv2_dgrad stores the gradient of the pooled activation as bf16
    if spec.lrn:
        p1 = tf_lrn(_q(p1, q))
    p1 = _q(p1, q)
    w2 = _q(params["conv2/weights"], q).permute(3, 2, 0, 1)
    a2 = F.relu(_qg(F.conv2d(p1, w2, params["conv2/biases"], padding=2), q))   # dZ2 planes are bf16
    acts["conv2"] = a2
    if spec.lrn:
        a2 = tf_lrn(_q(a2, q))
    p2 = _qg(_q(F.max_pool2d(a2, 2), q), q)   # dP2 is bf16
    flat = p2.permute(0, 2, 3, 1).reshape(B, -1)
    h = F.relu(_qg(flat @ _q(params["local3/weights"], q) + params["local3/biases"], q))   # dH of local3 is bf16
    acts["local3"] = h
    if spec.fc2:
        h = _q(h, q)
        h = F.relu(_qg(h @ _q(params["local4/weights"], q) + params["local4/biases"], q))
        acts["local4"] = h
    if dropout_mask is not None and spec.keep_prob < 1.0:
        h = h * dropout_mask.to(h.dtype) / spec.keep_prob
    h = _q(h, q)
    logits = h @ params["softmax_linear/weights"] + params["softmax_linear/biases"]
    acts["softmax_linear"] = logits
    if return_activations:
        return logits, acts
    return logits


def weight_losses(params: dict[str, torch.Tensor], spec: ModelSpec) -> dict[str, torch.Tensor]:
    """``l2_loss(var) * wd`` for every variable created with a weight-decay factor (mnist_input.py:113)."""
    return {v.name: 0.5 * v.wd * (params[v.name] ** 2).sum() for v in spec.variables if v.wd is not None}


def loss_fn(logits: torch.Tensor, labels: torch.Tensor, params, spec: ModelSpec):
    """(total_loss, cross_entropy_mean): cross entropy mean + all weight-decay terms — mnist_input.py:224-231."""
    xent = F.cross_entropy(logits, labels.long())
    total = xent
    for wl in weight_losses(params, spec).values():
        total = total + wl
    return total, xent


def accuracy_fn(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """``in_top_k(logits, labels, 1)`` mean: correct unless another class scores strictly higher (mnist_input.py:238-240)."""
    target = logits.gather(1, labels.long().view(-1, 1))
    return ((logits > target).sum(dim=1) == 0).float().mean()


def learning_rate(base_lr: float, decay_rate: float, decay_steps: int, step: int) -> float:
    """Staircase ``exponential_decay`` — mnist_input.py:252-256."""
    return base_lr * decay_rate ** math.floor(step / decay_steps) if decay_steps > 0 else base_lr


def ema_keep(step: int, decay: float = MOVING_AVERAGE_DECAY) -> float:
    """``min(decay, (1 + num_updates) / (10 + num_updates))`` of tf.train.ExponentialMovingAverage."""
    return min(decay, (1.0 + step) / (10.0 + step))


class ReferenceOptimizer:
    """TF-1.x optimizer semantics on a dict of tensors: SGD / Momentum / Adam, EMA(0.9999) shadows, global_step.

    This is the oracle the fused update kernels are tested against and the optimizer of the CPU execution path.
    """

    def __init__(self, params: dict[str, torch.Tensor], spec: ModelSpec, *, optimizer="adam", base_lr=1e-3, decay_rate=0.1,
                 decay_steps=0, momentum=0.9, beta1=0.9, beta2=0.999, epsilon=1e-8, ema_decay=MOVING_AVERAGE_DECAY):
        self.spec, self.params = spec, params
        self.kind = optimizer.lower()
        if self.kind not in ("sgd", "momentum", "adam"):
            raise ValueError(f"unknown optimizer {optimizer!r} (sgd | momentum | adam)")
        self.base_lr, self.decay_rate, self.decay_steps = base_lr, decay_rate, decay_steps
        self.momentum, self.beta1, self.beta2, self.epsilon, self.ema_decay = momentum, beta1, beta2, epsilon, ema_decay
        self.global_step = 0
        self.beta1_power, self.beta2_power = beta1, beta2
        z = lambda: {k: torch.zeros_like(v) for k, v in params.items()}
        self.slot0 = z() if self.kind in ("momentum", "adam") else {}
        self.slot1 = z() if self.kind == "adam" else {}
        self.ema = {k: v.detach().clone() for k, v in params.items()}  # TF initialises shadows to the variable value

    def step(self, grads: dict[str, torch.Tensor]) -> float:
        lr = learning_rate(self.base_lr, self.decay_rate, self.decay_steps, self.global_step)
        keep = ema_keep(self.global_step, self.ema_decay)
        lr_t = lr
        if self.kind == "adam":
            lr_t = lr * math.sqrt(1.0 - self.beta2_power) / (1.0 - self.beta1_power)
        with torch.no_grad():
            for name, w in self.params.items():
                g = grads[name]
                if self.kind == "adam":
                    m, v = self.slot0[name], self.slot1[name]
                    m.mul_(self.beta1).add_(g, alpha=1.0 - self.beta1)
                    v.mul_(self.beta2).addcmul_(g, g, value=1.0 - self.beta2)
                    w.sub_(lr_t * m / (v.sqrt() + self.epsilon))
                elif self.kind == "momentum":
                    a = self.slot0[name]
                    a.mul_(self.momentum).add_(g)
                    w.sub_(lr * a)
                else:
                    w.sub_(lr * g)
                s = self.ema[name]
                s.sub_((1.0 - keep) * (s - w))
        if self.kind == "adam":
            self.beta1_power *= self.beta1
            self.beta2_power *= self.beta2
        self.global_step += 1
        return lr
