#!/usr/bin/env python
"""Stock PyTorch baseline of the MNIST CNN: nn.Module + cuDNN/cuBLAS + torch.optim + DDP/NCCL.

This is the *comparison* arm recorded in BASELINE.md (the reference itself — TF-1.x/py2 with three absent IBM
modules, reference main.py:1,:34-36 — cannot run on this box).  Nothing from the framework's own engine is used
here: no custom kernels, no fused communication.  Launch with torchrun for N > 1.

    python baseline/torch_ddp_baseline.py --preset baseline --batch 512 --steps 50 --warmup 10 [--amp]
"""
from __future__ import annotations

import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class TFLocalResponseNorm(nn.Module):
    """tf.nn.lrn semantics (alpha is NOT divided by the window size, unlike torch's LocalResponseNorm)."""

    def __init__(self, radius=4, bias=1.0, alpha=0.001 / 9.0, beta=0.75):
        super().__init__()
        self.radius, self.bias, self.alpha, self.beta = radius, bias, alpha, beta

    def forward(self, x):  # NCHW
        sq = (x * x).unsqueeze(1)
        n = 2 * self.radius + 1
        sq = F.pad(sq, (0, 0, 0, 0, self.radius, self.radius))
        s = F.avg_pool3d(sq, (n, 1, 1), stride=1).squeeze(1) * n
        return x / (self.bias + self.alpha * s) ** self.beta


class MnistCNN(nn.Module):
    def __init__(self, preset="baseline", channels=1, fc=1024):
        super().__init__()
        self.preset = preset
        self.conv1 = nn.Conv2d(channels, 32, 5, padding=2)
        self.conv2 = nn.Conv2d(32, 64, 5, padding=2)
        self.fc1 = nn.Linear(3136, fc)
        if preset == "reference":
            self.lrn = TFLocalResponseNorm()
            self.fc2 = nn.Linear(fc, 192)
            self.out = nn.Linear(192, 10)
        else:
            self.drop = nn.Dropout(0.5)
            self.out = nn.Linear(fc, 10)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        if self.preset == "reference":
            x = self.lrn(x)
        x = F.relu(self.conv2(x))
        if self.preset == "reference":
            x = self.lrn(x)
        x = F.max_pool2d(x, 2)
        x = x.permute(0, 2, 3, 1).flatten(1)  # NHWC flatten order, like the TF model
        x = F.relu(self.fc1(x))
        if self.preset == "reference":
            x = F.relu(self.fc2(x))
        else:
            x = self.drop(x)
        return self.out(x)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="baseline", choices=["baseline", "reference", "wide"])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--amp", action="store_true", help="bf16 autocast (same compute dtype as the framework)")
    ap.add_argument("--channels_last", action="store_true")
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--e2e", action="store_true", help="include pinned H2D of inputs and D2H of the loss every step")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl")
    dev = torch.device("cuda", local)
    torch.backends.cudnn.benchmark = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True

    channels = 3 if args.preset == "reference" else 1
    fc = 8192 if args.preset == "wide" else 1024
    model = MnistCNN(args.preset, channels, fc).to(dev)
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
    if args.optimizer == "adam":
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    else:
        opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9 if args.optimizer == "momentum" else 0.0)

    B = args.batch
    host_x = torch.randint(0, 256, (4, B, 28 * 28 * channels), dtype=torch.uint8).pin_memory()
    host_y = torch.randint(0, 10, (4, B), dtype=torch.int32).pin_memory()
    dev_x = host_x.to(dev)
    dev_y = host_y.to(dev)

    def step(i):
        if args.e2e:
            xb = host_x[i % 4].to(dev, non_blocking=True)
            yb = host_y[i % 4].to(dev, non_blocking=True)
        else:
            xb, yb = dev_x[i % 4], dev_y[i % 4]
        x = (xb.float() * (1.0 / 255) - 0.5).view(B, 28, 28, channels).permute(0, 3, 1, 2)
        if args.channels_last:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=args.amp):
            logits = model(x)
            loss = F.cross_entropy(logits.float(), yb.long())
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        if args.e2e:
            return loss.item()
        return loss

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    if rank == 0:
        print(json.dumps({
            "impl": "torch_ddp_baseline", "preset": args.preset, "amp_bf16": args.amp, "e2e": args.e2e,
            "channels_last": args.channels_last, "n_gpus": world, "batch_per_gpu": B, "steps": args.steps,
            "ms_per_step": ms / args.steps, "images_per_sec": world * B * args.steps / (ms / 1e3),
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
