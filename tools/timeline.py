#!/usr/bin/env python
"""Kernel timeline of replayed training steps (CUPTI activity records through torch.profiler).

For finding gaps and overlaps inside the CUDA graph - NOT for reporting throughput (tracing perturbs timing slightly).
Prints, for one steady-state step, every kernel with its start offset, duration and the idle gap before it on the
device, and writes the raw records to gpurun_out/timeline.json.   python tools/timeline.py [--preset baseline] [--batch 512]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="baseline")
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=6)
    args = ap.parse_args()
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine
    from torch.profiler import ProfilerActivity, profile

    spec = presets.make_spec(args.preset)
    B = args.batch
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    comm = None
    if world > 1:       # under torchrun: the multi-GPU step (rank 0 reports its own device's timeline)
        import torch.distributed as dist
        from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = PeerComm(dist.group.WORLD, torch.device("cuda", local))
    eng = NativeEngine(spec, B, optimizer="adam", use_graph=True, comm=comm, device=torch.device("cuda", local))
    eng.load_params(oracle.init_params(spec))
    if comm is not None:
        eng.broadcast_params()
    x = torch.randint(0, 256, (B, spec.image_pixels), dtype=torch.uint8).cuda()
    y = torch.randint(0, 10, (B,), dtype=torch.int32).cuda()
    eng.stage_batch(0, x, y)
    for _ in range(30):
        eng.train_step(0)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(args.steps):
            eng.train_step(0)
        torch.cuda.synchronize()
    if rank != 0:
        torch.cuda.synchronize()
        dist.barrier()
        return 0
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "Memcpy" not in e.name and "Memset" not in e.name]
    recs = sorted(({"name": e.name, "start_us": e.time_range.start, "dur_us": e.time_range.end - e.time_range.start} for e in evs),
                  key=lambda r: r["start_us"])
    if not recs:
        print("no CUDA kernel records (CUPTI unavailable?)")
        return 1
    # split into steps at every conv1 forward kernel
    starts = [i for i, r in enumerate(recs) if "conv1_tc_kernel" in r["name"] or "conv1_sw_kernel" in r["name"]]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(recs, open(os.path.join(ROOT, "gpurun_out", "timeline.json"), "w"))
    if len(starts) < 4:
        print("too few steps recorded", len(starts))
        return 1
    a, b = starts[-3], starts[-2]
    step = recs[a:b]
    t0 = step[0]["start_us"]
    print(f"step length (conv1 start -> next conv1 start): {recs[b]['start_us'] - t0:.1f} us")
    busy_end = t0
    for r in step:
        gap = r["start_us"] - busy_end
        name = r["name"].replace("void dtm::", "").replace("dtm::", "")[:46]
        print(f"  +{r['start_us'] - t0:7.1f} us  dur {r['dur_us']:6.1f}  {'gap %5.1f' % gap if gap > 0.2 else 'overlap  ' if gap < -0.2 else '         '}  {name}")
        busy_end = max(busy_end, r["start_us"] + r["dur_us"])
    print(f"  end of last kernel +{busy_end - t0:.1f} us ; idle before the next step's first kernel {recs[b]['start_us'] - busy_end:.1f} us")
    if world > 1:
        dist.barrier()
    return 0


if __name__ == "__main__":
    sys.exit(main())
