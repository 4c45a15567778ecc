#!/usr/bin/env python
"""Throughput of the native TFRecord input pipeline (csrc/io/record_loader.cpp) on this host, images per second.

    python tools/loader_bench.py [--threads 1 2 4] [--records 60000] [--batch 512] [--seconds 2.0]

Writes one MNIST-shaped TFRecord file (28 x 28 x 1 uint8 `image_raw` + int64 `label`, the layout convert_mnist.py
produces) into a temporary directory, then times `next_packed()` - the call the trainer's feeder makes - after the shuffle
queue has filled, with the loader's batch ring (a C++ thread assembles batches ahead) and without it.  Host-only: no GPU needed.
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_tensorflow_ibm_mnist_b200.data.readers import TFRecordBatches  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.formats.tfrecord import TFRecordWriter, encode_example  # noqa: E402


def _steal_jiffies() -> int:
    """CPU time the hypervisor gave to other guests (all cores), from /proc/stat; explains run-to-run spread on shared hosts."""
    try:
        with open("/proc/stat") as f:
            return int(f.readline().split()[8])
    except (OSError, IndexError, ValueError):
        return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, nargs="+", default=[1, 2, 4])
    ap.add_argument("--records", type=int, default=60000)
    ap.add_argument("--files", type=int, default=1)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--seconds", type=float, default=3.0)
    ap.add_argument("--channels", type=int, default=1)
    ap.add_argument("--prefetch", type=int, nargs="+", default=[4, 0],
                    help="batch-ring sizes to measure (0 = the caller's thread dequeues)")
    args = ap.parse_args()
    image_bytes = 28 * 28 * args.channels
    rng = np.random.default_rng(0)
    with tempfile.TemporaryDirectory() as d:
        names = []
        per = args.records // args.files
        for f in range(args.files):
            names.append(os.path.join(d, "train-%d.tfrecords" % f))
            with TFRecordWriter(names[-1]) as w:
                for i in range(per):
                    w.write(encode_example({"image_raw": rng.integers(0, 256, image_bytes, dtype=np.uint8).tobytes(),
                                            "label": int(i % 10), "height": 28, "width": 28, "depth": args.channels}))
        rows = []
        for t, ring in [(t, r) for r in args.prefetch for t in args.threads]:
            it = TFRecordBatches(names, args.batch, image_bytes, num_threads=t, seed=1, prefetch_batches=ring)
            for _ in range(8):
                it.next_packed()
            rates = []                                   # the host is shared: report the median and the best of five windows
            steal0, cpu0, total = _steal_jiffies(), time.thread_time(), 0
            for _ in range(5):
                n, t0 = 0, time.perf_counter()
                while time.perf_counter() - t0 < args.seconds / 5:
                    it.next_packed()
                    n += args.batch
                rates.append(n / (time.perf_counter() - t0))
                total += n
            cpu_ns = (time.thread_time() - cpu0) / total * 1e9
            steal = (_steal_jiffies() - steal0) / 100.0 / args.seconds
            it.close()
            rates.sort()
            rows.append({"threads": t, "prefetch_batches": ring, "images_per_s": round(rates[2]), "best_window": round(rates[-1]), "worst_window": round(rates[0]),
                         "caller_cpu_ns_per_image": round(cpu_ns), "stolen_cpu_seconds_per_second": round(steal, 2),
                         "native": type(it).__name__ == "NativeRecordBatches"})
            print(json.dumps(rows[-1]), flush=True)
    return rows


if __name__ == "__main__":
    main()
