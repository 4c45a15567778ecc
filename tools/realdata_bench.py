#!/usr/bin/env python
"""Training throughput from REAL TFRecord files on one GPU: C++ loader -> pinned batch ring -> copy stream -> NativeEngine.

    python tools/realdata_bench.py [--steps 2000] [--warmup 100] [--threads 2] [--prefetch 4] [--out file.json]

Writes an MNIST-shaped TFRecord file (60 000 records), then times `DeviceFeeder.step()` - H2D copy of every batch from
pinned memory included - with CUDA events around the timed steps and a device->host read of the loss at the end, for the
TFRecord source and, as the yard-stick, for the synthetic pinned pool bench.py uses.  The host side must keep up with a
140 us step: the JSON reports both rates and the loader's CPU budget.
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_tensorflow_ibm_mnist_b200.data.readers import TFRecordBatches  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.formats import tfrecord  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine  # noqa: E402


def run(engine, source, steps, warmup):
    feeder = DeviceFeeder(engine, source)
    for _ in range(warmup):
        feeder.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        feeder.step()
    e1.record()
    host_s = time.perf_counter() - t0            # how long the host needed to ENQUEUE the steps
    torch.cuda.synchronize()
    loss = engine.metrics()["cross_entropy"]
    ms = e0.elapsed_time(e1)
    return {"images_per_s": round(engine.B * steps / (ms * 1e-3)), "us_per_step": round(ms * 1e3 / steps, 1),
            "host_enqueue_us_per_step": round(host_s * 1e6 / steps, 1), "final_xent": round(float(loss), 4)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--threads", type=int, nargs="+", default=[2])
    ap.add_argument("--prefetch", type=int, nargs="+", default=[4, 0])
    ap.add_argument("--records", type=int, default=60000)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    spec = presets.make_spec("baseline")
    B = args.batch
    res = {"batch": B, "steps": args.steps, "warmup": args.warmup, "cpus": os.cpu_count(), "rows": []}
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "train.tfrecords")
        pool = SyntheticMnist(B, spec.image_pixels, pool_batches=max(1, args.records // B), seed=1, separable=True, pin=False)
        imgs, labels = [], []
        for _ in range(max(1, args.records // B)):
            x, y = pool.next()
            imgs += [bytes(r.numpy().tobytes()) for r in x.view(B, -1)]
            labels += [int(v) for v in y]
        tfrecord.write_mnist_tfrecord(path, imgs, labels)
        eng = NativeEngine(spec, B, optimizer="adam", base_lr=1e-3, seed=3, use_graph=True)
        eng.load_params(oracle.init_params(spec, seed=0))
        syn = SyntheticMnist(B, spec.image_pixels, pool_batches=16, seed=2, separable=True, pin=True)
        res["rows"].append({"source": "synthetic pinned pool", **run(eng, syn, args.steps, args.warmup)})
        print(json.dumps(res["rows"][-1]), flush=True)
        for ring in args.prefetch:
            for t in args.threads:
                src = TFRecordBatches([path], B, spec.image_pixels, seed=4, num_threads=t, pin=True, prefetch_batches=ring)
                row = {"source": "tfrecord", "reader_threads": t, "prefetch_batches": ring, **run(eng, src, args.steps, args.warmup)}
                src.close()
                res["rows"].append(row)
                print(json.dumps(row), flush=True)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
