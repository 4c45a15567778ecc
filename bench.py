#!/usr/bin/env python
"""Headline benchmark: MNIST-CNN training throughput (images/sec) on N B200 GPUs of one node.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Config (BASELINE.json): `baseline` preset (conv5x5-32, conv5x5-64, fc1024, dropout, fc10), batch 512 per GPU (weak
scaling), Adam, bf16 tensor-core operands with fp32 accumulation and fp32 master weights, synthetic MNIST-shaped
data, random-init weights.  Prints ONE JSON line on rank 0.

  value   device-timed (CUDA events, max over ranks) steady-state throughput; inputs come from a device-resident
          pool larger than L2, one fresh batch per step.
  e2e     the same steps through the public feeder API: every step copies its batch from pinned host memory (H2D)
          and the host reads that step's loss (D2H) inside the timed region.

`--impl reference` would run the unmodified reference; it cannot be installed on this box, so that
arm prints {"impl": "reference", "unavailable": ...} and exits 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# The reference publishes no number (BASELINE.json "published": {}).  BASELINE.md section 4 records the
# fallback baseline the north-star names - stock PyTorch DDP + NCCL + cuDNN/cuBLAS, same CNN, same box, batch 512/GPU - and
# vs_baseline is quoted against that table where it has an entry (images/sec, better of fp32/TF32 and bf16 autocast).
BASELINE_IMAGES_PER_SEC = {
    ("baseline", 1): 483170.0, ("baseline", 2): 677172.0, ("baseline", 4): 1433660.0, ("baseline", 8): 2296187.0,
    ("reference", 1): 170616.0, ("reference", 2): 342580.0, ("reference", 4): 683660.0,
    ("wide", 1): 315920.0, ("wide", 2): 571068.0, ("wide", 4): 1125617.0,
}


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (B200_PROFILING.md recipe)."""

    def __init__(self, index: int, period_s: float = 0.05):
        self.index, self.period = index, period_s
        self.samples: list[int] = []
        self.reasons: set[str] = set()
        self.max_mhz = 0
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self) -> dict:
        return {"sm_mhz": int(statistics.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz or None,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def reference_arm() -> None:
    if int(os.environ.get("RANK", "0")) != 0:     # launched with torchrun for N > 1: ONE line, from rank 0
        return
    print(json.dumps({
        "impl": "reference",
        "unavailable": "reference is 5 loose TF-1.x/py2 scripts: pip reports 'neither setup.py nor pyproject.toml', "
                       "tensorflow is absent from /opt/wheelhouse, and tf_parameter_mgr/monitor_cb/inference_helper exist nowhere",
    }))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--preset", default="baseline", choices=["baseline", "reference", "wide"])
    ap.add_argument("--batch", type=int, default=512, help="per-GPU batch")
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm()
        return
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist

    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("for --gpus N > 1 launch with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm

        comm = PeerComm(dist.group.WORLD, dev)

    spec = presets.make_spec(args.preset)
    B = args.batch
    eng = NativeEngine(spec, B, device=dev, optimizer=args.optimizer, base_lr=1e-3, decay_rate=0.1,
                       decay_steps=presets.decay_steps(B), seed=1234, comm=comm, use_graph=not args.no_graph)
    eng.load_params(oracle.init_params(spec, seed=0))
    if comm is not None:
        eng.broadcast_params()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-timed steady state: one fresh batch per step out of a device pool larger than L2 (126 MB) -----
    pool_batches = max(8, (160 * 1024 * 1024) // (B * spec.image_pixels) + 1)
    host_pool = SyntheticMnist(B, spec.image_pixels, pool_batches=pool_batches, seed=rank, pin=False)

    class DevicePool:
        """The same packed batches, resident on the device (160 MB > L2): the feeder's copy stream moves one of them into
        the engine's input slot per step, beside the previous step's kernels - no host memory in this loop."""
        label_offset = host_pool.label_offset

        def __init__(self):
            self.packed = host_pool.packed.to(dev)
            self.i = 0

        def next_packed(self):
            self.i += 1
            return self.packed[(self.i - 1) % pool_batches]

    dev_feeder = DeviceFeeder(eng, DevicePool())

    def device_step(i: int) -> None:
        dev_feeder.step()

    for i in range(args.warmup):
        device_step(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        for i in range(args.steps):
            device_step(args.warmup + i)
        e1.record()
        barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    ms_per_step = ms / args.steps
    value = world * B * args.steps / (ms / 1e3)
    clocks = clk.summary()
    launches = eng.launches_per_step * args.steps
    st = eng.metrics()

    # ---- end to end through the public feeder: pinned H2D of every batch + D2H of every step's loss ---------
    e2e = None
    if not args.no_e2e:
        src = SyntheticMnist(B, spec.image_pixels, pool_batches=16, seed=100 + rank, pin=True)
        feeder = DeviceFeeder(eng, src)
        host_metrics = torch.zeros(4, dtype=torch.float32).pin_memory()   # last_xent, last_total, last_acc, last_lr
        state_f32 = eng.state.view(torch.float32)
        d2h_done, step_done = torch.cuda.Event(), torch.cuda.Event()
        d2h_stream = torch.cuda.Stream(device=dev)
        metrics_dev = state_f32[8:12]
        losses = []

        def e2e_step(first: bool) -> None:
            feeder.step()
            step_done.record(torch.cuda.current_stream())
            if not first:
                d2h_done.synchronize()                 # host reads the previous step's loss ...
                losses.append(float(host_metrics[0]))
            # ... while this step's loss is on its way: the D2H copy waits for this step on its own stream, so the next
            # step's kernels are not queued behind it (the values it reads are only rewritten at the END of the next step)
            d2h_stream.wait_event(step_done)
            with torch.cuda.stream(d2h_stream):
                host_metrics.copy_(metrics_dev, non_blocking=True)
                d2h_done.record(d2h_stream)

        for i in range(args.warmup):
            e2e_step(i == 0)
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(args.steps):
            e2e_step(False)
        d2h_done.synchronize()
        losses.append(float(host_metrics[0]))
        t1.record()
        barrier()
        ms_e2e = max_over_ranks(t0.elapsed_time(t1))
        e2e = {"value": world * B * args.steps / (ms_e2e / 1e3), "unit": "images/sec", "ms_per_step": ms_e2e / args.steps,
               "h2d_bytes_per_step": feeder.h2d_bytes_per_step, "d2h_bytes_per_step": 16,
               "host_losses_read": len(losses), "last_loss": losses[-1]}

    if rank == 0:
        out = {
            "metric": "MNIST-CNN train images/sec", "value": value, "unit": "images/sec", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": (value / BASELINE_IMAGES_PER_SEC[(args.preset, world)]) if (args.preset, world) in BASELINE_IMAGES_PER_SEC and args.batch == 512 else None,
            "baseline_source": "BASELINE.md section 4: PyTorch DDP+NCCL arm of this repo (the reference publishes nothing)",
            "dtype": "bf16", "data": "synthetic MNIST-shaped uint8 28x28x%d, random-init weights" % spec.channels,
            "config": {"model": "MNIST CNN (%s preset: conv5x5-32, conv5x5-64, fc%d, fc10)" % (spec.preset, spec.fc1),
                       "global_batch": world * B, "batch_per_gpu": B, "seq_len": None, "parallelism": "dp%d" % world,
                       "optimizer": args.optimizer, "cuda_graph": not args.no_graph,
                       "l2_policy": "inputs cycled from a %d MB device pool (> 126 MB L2), one fresh batch per step"
                                    % (pool_batches * B * spec.image_pixels // (1024 * 1024))},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "launches_per_step": eng.launches_per_step,
            "final_global_step": st["global_step"], "final_loss": st["cross_entropy"], "nan": st["nan"], "impl": "native",
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
