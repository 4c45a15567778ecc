#!/usr/bin/env python
"""Marginal cost of every kernel *inside the replayed CUDA graph*: rebuild the engine with one kernel left out (DTM_SKIP)
or with all kernels on one stream (DTM_SERIAL) and time graph replays.  Numerical results of ablated runs are garbage;
only the step time matters.  `python tools/ablate.py [--preset baseline] [--batch 512]`"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KERNELS = ["conv1_fwd", "conv2_fwd", "fc1_fwd", "head_fwd", "head_bwd", "dW3", "dP2", "update0", "fused_unpool", "conv2_dgrad",
           "conv2_wgrad", "conv1_wgrad", "update1"]


def run_one(preset, batch, skip="", serial=False):
    env = dict(os.environ, DTM_SKIP=skip, DTM_SERIAL="1" if serial else "0")
    code = f"""
import sys, torch
sys.path.insert(0, {ROOT!r})
from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine
spec = presets.make_spec({preset!r}); B = {batch}
eng = NativeEngine(spec, B, optimizer='adam', use_graph=True); eng.load_params(oracle.init_params(spec))
x = torch.randint(0, 256, (B, spec.image_pixels), dtype=torch.uint8).cuda(); y = torch.randint(0, 10, (B,), dtype=torch.int32).cuda()
eng.stage_batch(0, x, y)
for _ in range(20): eng.train_step(0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(300): eng.train_step(0)
e1.record(); torch.cuda.synchronize()
print('STEP_US', e0.elapsed_time(e1) / 300 * 1e3)
"""
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    for line in r.stdout.splitlines():
        if line.startswith("STEP_US"):
            return float(line.split()[1])
    return float("nan")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="baseline")
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    base = run_one(args.preset, args.batch)
    serial = run_one(args.preset, args.batch, serial=True)
    for mode in ("0", "1"):
        os.environ["DTM_PDL"] = mode
        print(f"DTM_PDL={mode}: {run_one(args.preset, args.batch):.1f} us", flush=True)
    os.environ.pop("DTM_PDL", None)
    out = {"full_step_us": base, "serial_step_us": serial, "marginal_us": {}}
    print(f"full step {base:.1f} us ; all kernels on one stream {serial:.1f} us", flush=True)
    for k in (args.only.split(",") if args.only else KERNELS):
        t = run_one(args.preset, args.batch, skip=k)
        out["marginal_us"][k] = base - t
        print(f"  without {k:12s}: {t:7.1f} us  (marginal {base - t:6.1f})", flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "ablate.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
