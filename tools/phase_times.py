#!/usr/bin/env python
"""Per-phase timeline of one training step on the main stream (CUDA events between kernel groups, no CUDA graph).

    python tools/phase_times.py [--preset baseline] [--batch 512] [--steps 30]

Prints the mean time between consecutive marks; the joins show how much of the side-stream work was NOT hidden.
"""
import argparse
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--preset", default="baseline")
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--steps", type=int, default=30)
args = ap.parse_args()
spec = presets.make_spec(args.preset)
eng = NativeEngine(spec, args.batch, optimizer="adam", use_graph=False)
eng.load_params(oracle.init_params(spec))
x = torch.randint(0, 256, (args.batch, spec.image_pixels), dtype=torch.uint8).cuda()
y = torch.randint(0, 10, (args.batch,), dtype=torch.int32).cuda()
eng.stage_batch(0, x, y)
for _ in range(5):
    eng.train_step(0)
torch.cuda.synchronize()
acc = collections.OrderedDict()
total = 0.0
for _ in range(args.steps):
    eng.trace = []
    eng.train_step(0)
    torch.cuda.synchronize()
    marks = eng.trace
    for (la, ea), (lb, eb) in zip(marks[:-1], marks[1:]):
        acc.setdefault(lb, []).append(ea.elapsed_time(eb) * 1e3)
    total += marks[0][1].elapsed_time(marks[-1][1]) * 1e3
for k, v in acc.items():
    print(f"{sum(v) / len(v):8.1f} us  -> {k}")
print(f"{total / args.steps:8.1f} us  whole step (eager launches, main-stream span)")
