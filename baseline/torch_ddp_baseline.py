#!/usr/bin/env python
"""Stock PyTorch baseline of the MNIST CNN: nn.Module + cuDNN/cuBLAS + torch.optim + DDP/NCCL.

This is the *comparison* arm recorded in BASELINE.md (the reference itself — TF-1.x/py2 with three absent IBM
modules, reference main.py:1,:34-36 — cannot run on this box).  Nothing from the framework's own engine is used
here: no custom kernels, no fused communication.  Launch with torchrun for N > 1.

    python baseline/torch_ddp_baseline.py --preset baseline --batch 512 --steps 50 --warmup 10 \
        --variants eager,eager_amp,graph,graph_amp,graph_amp_cl

Variants (one JSON line each, rank 0):
  eager          torch.optim.Adam (default foreach implementation), fp32 parameters, TF32 tensor-core math allowed
  eager_amp      + bf16 autocast (the framework's operand dtype)
  graph          the whole step (forward, backward, DDP's NCCL all-reduce, Adam(fused=True, capturable=True)) captured in
                 ONE CUDA graph and replayed - the strongest stock configuration: no host launch cost at all
  graph_amp      + bf16 autocast          graph_amp_cl   + channels_last activations/weights
Every variant is timed twice: device-resident inputs (CUDA events, max over ranks) and end to end (pinned H2D of each
batch + D2H read of each step's loss inside the timed region), mirroring bench.py's two numbers.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import traceback

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


def run_variant(variant: str, args, dev, world: int, rank: int, local: int) -> dict:
    graph = variant.startswith("graph")
    amp = "_amp" in variant
    cl = variant.endswith("_cl")
    channels = 3 if args.preset == "reference" else 1
    fc = 8192 if args.preset == "wide" else 1024
    B = args.batch
    torch.manual_seed(1234)
    model = MnistCNN(args.preset, channels, fc).to(dev)
    if cl:
        model = model.to(memory_format=torch.channels_last)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):      # DDP built and warmed on the stream the graph will be captured from
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        if args.optimizer == "adam":
            opt = torch.optim.Adam(model.parameters(), lr=1e-3, **({"fused": True, "capturable": True} if graph else {}))
        else:
            opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9 if args.optimizer == "momentum" else 0.0)

    pool = max(4, (160 * 1024 * 1024) // (B * 28 * 28 * channels) + 1) if args.big_pool else 4
    g = torch.Generator().manual_seed(rank)
    host_x = torch.randint(0, 256, (16, B, 28 * 28 * channels), dtype=torch.uint8, generator=g).pin_memory()
    host_y = torch.randint(0, 10, (16, B), dtype=torch.int32, generator=g).pin_memory()
    dev_x = torch.randint(0, 256, (pool, B, 28 * 28 * channels), dtype=torch.uint8, device=dev)
    dev_y = torch.randint(0, 10, (pool, B), dtype=torch.int32, device=dev)
    static_x = torch.zeros(B, 28 * 28 * channels, dtype=torch.uint8, device=dev)
    static_y = torch.zeros(B, dtype=torch.int32, device=dev)
    static_loss = torch.zeros((), device=dev)

    def body(xb, yb):
        x = (xb.float() * (1.0 / 255) - 0.5).view(B, 28, 28, channels).permute(0, 3, 1, 2)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            logits = model(x)
            loss = F.cross_entropy(logits.float(), yb.long())
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss

    cuda_graph = None
    if graph:
        with torch.cuda.stream(side):
            for _ in range(11):           # DDP needs >= 11 eager iterations before whole-network capture (torch CUDA-graphs notes)
                body(static_x, static_y)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        cuda_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(cuda_graph, stream=side):
            static_loss.copy_(body(static_x, static_y))
        torch.cuda.synchronize()

    def step(xb, yb):
        if cuda_graph is not None:
            static_x.copy_(xb, non_blocking=True)
            static_y.copy_(yb, non_blocking=True)
            cuda_graph.replay()
            return static_loss
        return body(xb.to(dev, non_blocking=True), yb.to(dev, non_blocking=True))

    def timed(e2e: bool) -> float:
        def one(i):
            if e2e:
                return float(step(host_x[i % 16], host_y[i % 16]).item())
            return step(dev_x[i % pool], dev_y[i % pool])

        prime = min(3, args.warmup)
        for i in range(args.warmup - prime):
            one(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        for i in range(prime):
            one(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            one(i)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms = timed(False)
    ms_e2e = timed(True)
    return {"impl": "torch_ddp_baseline", "variant": variant, "preset": args.preset, "amp_bf16": amp, "cuda_graph": graph,
            "channels_last": cl, "n_gpus": world, "batch_per_gpu": B, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "images_per_sec": world * B * args.steps / (ms / 1e3),
            "e2e_ms_per_step": ms_e2e / args.steps, "e2e_images_per_sec": world * B * args.steps / (ms_e2e / 1e3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="baseline", choices=["baseline", "reference", "wide"])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--variants", default="eager,eager_amp,graph,graph_amp,graph_amp_cl")
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--big_pool", type=int, default=1, help="cycle device inputs through a pool larger than L2")
    # flags of the round-1 script, mapped onto variants
    ap.add_argument("--amp", action="store_true")
    ap.add_argument("--channels_last", action="store_true")
    ap.add_argument("--e2e", action="store_true")
    args = ap.parse_args()
    if args.amp:
        args.variants = "eager_amp"

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")    # whole-network capture of DDP (torch CUDA-graphs notes)
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True

    for variant in filter(None, args.variants.split(",")):
        try:
            row = run_variant(variant, args, dev, world, rank, local)
        except Exception as e:  # noqa: BLE001 - a variant that cannot run is reported, the others still count
            row = {"impl": "torch_ddp_baseline", "variant": variant, "preset": args.preset, "n_gpus": world,
                   "error": f"{type(e).__name__}: {e}".splitlines()[0][:300]}
            traceback.print_exc(file=sys.stderr)
            # a failed capture can leave this process (or the communicator) unusable for the next variant
            if rank == 0:
                print(json.dumps(row), flush=True)
            break
        if rank == 0:
            print(json.dumps(row), flush=True)
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
