#!/usr/bin/env python
"""Run the GPU test-suite once; re-run every failed test in its own process.

A device-side trap (e.g. the pipeline watchdog) poisons the CUDA context of the whole pytest process, so later
tests fail by contagion.  The isolated second pass separates real failures from collateral ones.
Writes gpurun_out/gpu_tests.log and gpurun_out/gpu_tests_isolated.log.
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
args = sys.argv[1:] or ["tests"]
r = subprocess.run([sys.executable, "-m", "pytest", *args, "-m", "gpu", "-q", "-rf", "--timeout=300"], cwd=ROOT,
                   capture_output=True, text=True)
open(os.path.join(OUT, "gpu_tests.log"), "w").write(r.stdout + r.stderr)
print(r.stdout[-6000:])
failed = re.findall(r"^FAILED (\S+)", r.stdout, flags=re.M)
if failed:
    with open(os.path.join(OUT, "gpu_tests_isolated.log"), "w") as f:
        for node in failed[:25]:
            rr = subprocess.run([sys.executable, "-m", "pytest", node, "-q", "-x", "--timeout=300"], cwd=ROOT, capture_output=True, text=True)
            verdict = "PASS-isolated" if rr.returncode == 0 else "FAIL-isolated"
            f.write(f"===== {node}: {verdict}\n{rr.stdout[-3500:]}\n{rr.stderr[-1500:]}\n")
            print(f"===== {node}: {verdict}")
            if rr.returncode != 0:
                print(rr.stdout[-2500:])
sys.exit(r.returncode)
