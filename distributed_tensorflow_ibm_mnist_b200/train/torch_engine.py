"""CPU / reference execution engine: the same model math in plain PyTorch autograd.

This is the path used (a) on machines without a B200 — the `single worker on CPU, batch 128` and the
`1 ps + 2 workers on gloo` configurations of BASELINE.json — and (b) as the oracle that the sm_100a engine is
tested against.  It exposes the same interface as :class:`NativeEngine`; gradients are synchronised across workers
with ``torch.distributed.all_reduce`` when a process group is given (synchronous data parallelism replaces the
reference's asynchronous parameter-server updates).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..models import oracle
from ..models.presets import ModelSpec
from ..ops.philox import dropout_keep_mask


class TorchEngine:
    def __init__(self, spec: ModelSpec, batch_size: int, *, device="cpu", optimizer="adam", base_lr=1e-3, decay_rate=0.1,
                 decay_steps=0, momentum=0.9, seed=0, group=None, **_unused):
        self.spec, self.B = spec, batch_size
        self.device = torch.device(device)
        self.seed = seed
        self.group = group
        self.world = dist.get_world_size(group) if group is not None else 1
        self.optimizer = optimizer
        self._opt_kw = dict(optimizer=optimizer, base_lr=base_lr, decay_rate=decay_rate, decay_steps=decay_steps, momentum=momentum)
        self.w: dict[str, torch.Tensor] = {}
        self.opt: oracle.ReferenceOptimizer | None = None
        self.x = [None, None]
        self.y = [None, None]
        self.loss_avg = [0.0] * (len(spec.variables) + 2)   # per-variable weight_loss..., cross_entropy, total_loss
        self._last = {"cross_entropy": float("nan"), "total_loss": float("nan"), "accuracy": float("nan"), "lr": base_lr}
        self._last_wl: dict[str, float] = {}
        self._grads: dict[str, torch.Tensor] = {}
        self._acts: dict[str, torch.Tensor] = {}
        self.nan = False
        self.launches_per_step = 0

    # ---- parameters / state ---------------------------------------------------------------------------
    def load_params(self, tensors, *, ema=None, slots=None, global_step=0, loss_avg=None) -> None:
        self.w = {v.name: torch.as_tensor(tensors[v.name]).to(self.device, torch.float32).clone() for v in self.spec.variables}
        self.opt = oracle.ReferenceOptimizer(self.w, self.spec, **self._opt_kw)
        self.opt.global_step = int(global_step)
        self.opt.beta1_power = self.opt.beta1 ** (global_step + 1)
        self.opt.beta2_power = self.opt.beta2 ** (global_step + 1)
        if ema:
            for k, v in ema.items():
                if k in self.opt.ema:
                    self.opt.ema[k].copy_(torch.as_tensor(v))
        if slots:
            for store, d in zip((self.opt.slot0, self.opt.slot1), slots):
                for k, v in (d or {}).items():
                    if k in store:
                        store[k].copy_(torch.as_tensor(v))
        if loss_avg:
            self.loss_avg = list(loss_avg) + [0.0] * (len(self.loss_avg) - len(loss_avg))

    def params(self):
        return {k: v.detach().cpu().clone() for k, v in self.w.items()}

    def ema_params(self):
        return {k: v.detach().cpu().clone() for k, v in self.opt.ema.items()}

    def slots(self):
        out = []
        if self.opt.slot0:
            out.append({k: v.cpu().clone() for k, v in self.opt.slot0.items()})
        if self.opt.slot1:
            out.append({k: v.cpu().clone() for k, v in self.opt.slot1.items()})
        return out

    def gradients(self):
        return {k: v.cpu().clone() for k, v in self._grads.items()}

    def activations(self):
        return {k: v.detach().cpu() for k, v in self._acts.items()}

    # ---- stepping -----------------------------------------------------------------------------------
    def stage_batch(self, slot: int, images_u8: torch.Tensor, labels: torch.Tensor, stream=None) -> None:
        self.x[slot] = images_u8.reshape(self.B, -1).to(self.device)
        self.y[slot] = labels.reshape(self.B).to(self.device)

    def _forward_loss(self, slot: int, training: bool):
        x = oracle.normalise_u8(self.x[slot]) if self.x[slot].dtype == torch.uint8 else self.x[slot].float()
        mask = None
        if training and self.spec.keep_prob < 1.0:
            mask = torch.from_numpy(dropout_keep_mask(self.B, self.spec.head_in, seed=self.seed, step=self.opt.global_step,
                                                      keep_prob=self.spec.keep_prob)).to(self.device)
        logits, acts = oracle.forward(self.w, x, self.spec, dropout_mask=mask, return_activations=True)
        total, xent = oracle.loss_fn(logits, self.y[slot], self.w, self.spec)
        return logits, acts, total, xent

    def train_step(self, slot: int = 0) -> None:
        for v in self.w.values():
            v.requires_grad_(True)
            v.grad = None
        logits, acts, total, xent = self._forward_loss(slot, training=True)
        total.backward()
        grads = {k: v.grad.detach() for k, v in self.w.items()}
        for v in self.w.values():
            v.requires_grad_(False)
        if self.group is not None and self.world > 1:
            flat = torch.cat([g.reshape(-1) for g in grads.values()])
            dist.all_reduce(flat, group=self.group)
            flat /= self.world
            off = 0
            for k, g in grads.items():
                grads[k] = flat[off:off + g.numel()].view_as(g)
                off += g.numel()
        wl = {k: float(v) for k, v in oracle.weight_losses(self.w, self.spec).items()}
        acc = float(oracle.accuracy_fn(logits.detach(), self.y[slot]))
        lr = self.opt.step(grads)
        self._grads, self._acts = grads, acts
        xent_f, total_f = float(xent.detach()), float(total.detach())
        if not (abs(total_f) < float("inf")):
            self.nan = True
        # loss moving averages, decay 0.9 (mnist_input.py:288-290)
        values = [wl.get(v.name, 0.0) for v in self.spec.variables] + [xent_f, total_f]
        self.loss_avg = [a - 0.1 * (a - b) for a, b in zip(self.loss_avg, values)]
        self._last = {"cross_entropy": xent_f, "total_loss": total_f, "accuracy": acc, "lr": lr}
        self._last_wl = wl

    def evaluate(self, slot: int = 0) -> tuple[float, float]:
        """(total loss, accuracy) of the staged batch.  The reference feeds ONE tensor - ``total_loss``, cross-entropy plus
        the weight-decay terms - to both the 'train loss' and the 'test loss' scalar (main.py:103-106), so the test value
        includes the l2 terms as well; otherwise the two TensorBoard curves are not comparable."""
        with torch.no_grad():
            logits, _, total, xent = self._forward_loss(slot, training=False)
            return float(total), float(oracle.accuracy_fn(logits, self.y[slot]))

    def predict(self, images: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            x = oracle.normalise_u8(images) if images.dtype == torch.uint8 else images.float()
            return torch.softmax(oracle.forward(self.w, x.to(self.device).reshape(x.shape[0], -1), self.spec), dim=1).cpu()

    def metrics(self) -> dict:
        return {"global_step": self.opt.global_step, **self._last, "nan": self.nan}

    def scalar_state(self) -> dict:
        """Scalar checkpoint variables: global_step, Adam beta powers, loss averages."""
        return {"global_step": self.opt.global_step, "beta1_power": self.opt.beta1_power, "beta2_power": self.opt.beta2_power,
                "loss_avg": list(self.loss_avg)}
