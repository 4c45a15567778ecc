#!/usr/bin/env python
"""Headline benchmark: MNIST-CNN training throughput (images/sec) on N B200 GPUs of one node.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Config (BASELINE.json): `baseline` preset (conv5x5-32, conv5x5-64, fc1024, dropout, fc10), batch 512 per GPU (weak
scaling), Adam, bf16 tensor-core operands with fp32 accumulation and fp32 master weights, synthetic MNIST-shaped
data, random-init weights.  Prints ONE JSON line on rank 0.

  value     device-timed (CUDA events, max over ranks) steady-state throughput; inputs come from a device-resident
            pool larger than L2, one fresh batch per step.
  e2e       the same steps through the public feeder API: every step copies its batch from pinned host memory (H2D)
            and the host reads that step's loss (D2H) inside the timed region.
  baseline_measured   the comparison arm (baseline/torch_ddp_baseline.py: stock torch.nn + cuDNN/cuBLAS + DDP/NCCL,
            eager and whole-step CUDA-graph variants) run by this same process tree on the same GPUs with the same
            N / batch / steps / warm-up; `vs_baseline` = value / the FASTEST of those variants.
  presets   device-timed throughput of the other two model presets (`reference` = the reference's real graph,
            `wide` = the communication-bound fc-8192 model) and, for N > 1, the stand-alone time of the fused
            gradient exchange kernel with its own-bytes NVLink roofline.

Timing protocol (the reference's loop being timed is main.py:147-154): graph capture and W warm-up steps, the last
three of them AFTER the cross-rank host barrier so that the device queues are primed and - every step contains a
cross-GPU barrier - the ranks are aligned on the device when the start event is recorded; nothing but step launches
sits between that barrier and the stop event; the clock sampler thread (NVML) is created before the barrier.

`--impl reference` would run the unmodified reference; it cannot be installed on this box, so that
arm prints {"impl": "reference", "unavailable": ...} and exits 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Fallback only (used when the in-run comparison arm could not be measured, e.g. --no-baseline): BASELINE.md section 4,
# stock PyTorch DDP + NCCL + cuDNN/cuBLAS, same CNN, same box, batch 512/GPU, images/sec.
BASELINE_MD_IMAGES_PER_SEC = {      # round 2: best variant (whole step in a CUDA graph, bf16 autocast, channels_last)
    ("baseline", 1): 579253.0, ("baseline", 2): 1073680.0, ("baseline", 4): 2128059.0, ("baseline", 8): 4084252.0,
}
NVLINK_BYTES_PER_SEC = 900e9      # per GPU and direction (task statement / B200_PROFILING.md)


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU about every millisecond (B200_PROFILING.md recipe).  NVML is
    initialised in the constructor - long before the timed region - and every sample carries a host timestamp, so a
    window's clocks are the samples between the host times that bracket its launches and its final synchronize."""

    def __init__(self, index: int, period_s: float = 0.001):
        self.index, self.period = index, period_s
        self.samples: list[tuple[float, int, int]] = []      # (host time, SM MHz, reason mask)
        self.max_mhz = 0
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self._sample()      # first call pays NVML's lazy set-up
        except Exception:
            self.nv = None

    def _sample(self) -> None:
        nv = self.nv
        mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        try:
            mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        self.samples.append((time.perf_counter(), mhz, mask))

    def _loop(self):
        while not self._stop.is_set():
            try:
                self._sample()
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.nv is not None and self._thread is None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self, t0: float | None = None, t1: float | None = None) -> dict:
        names = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake": 0x80}
        rows = [s for s in self.samples if (t0 is None or s[0] >= t0) and (t1 is None or s[0] <= t1)]
        mask = 0
        for _, _, m in rows:
            mask |= m
        return {"sm_mhz": int(statistics.median(r[1] for r in rows)) if rows else None,
                "sm_min_mhz": min((r[1] for r in rows), default=None), "sm_max_mhz": self.max_mhz or None,
                "reasons": sorted(k for k, bit in names.items() if mask & bit), "samples": len(rows)}


def reference_arm() -> None:
    if int(os.environ.get("RANK", "0")) != 0:     # launched with torchrun for N > 1: ONE line, from rank 0
        return
    print(json.dumps({
        "impl": "reference",
        "unavailable": "reference is 5 loose TF-1.x/py2 scripts: pip reports 'neither setup.py nor pyproject.toml', "
                       "tensorflow is absent from /opt/wheelhouse, and tf_parameter_mgr/monitor_cb/inference_helper exist nowhere",
    }))


def run_baseline_arm(args, preset: str, variants: str, timeout_s: int = 900) -> dict:
    """Run the stock-PyTorch comparison arm on the same GPUs with the same N / batch / steps / warm-up: every rank of this
    job starts one rank of the DDP job (own rendezvous port) and waits for it; rank 0 returns the parsed rows."""
    env = os.environ.copy()
    env.pop("TORCHELASTIC_USE_AGENT_STORE", None)       # the child job brings up its own store on its own port
    env["MASTER_ADDR"] = env.get("MASTER_ADDR", "127.0.0.1")
    env["MASTER_PORT"] = str(int(env.get("MASTER_PORT", "29500")) + 23)
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "torch_ddp_baseline.py"), "--preset", preset, "--batch", str(args.batch),
           "--steps", str(args.steps), "--warmup", str(args.warmup), "--optimizer", args.optimizer, "--variants", variants]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout_s)
    except subprocess.TimeoutExpired:
        return {"error": f"comparison arm timed out after {timeout_s} s"}
    rows = {}
    for line in r.stdout.splitlines():
        if line.startswith("{"):
            try:
                row = json.loads(line)
                rows[row.get("variant", "?")] = row
            except ValueError:
                pass
    if r.returncode != 0 and not rows:
        return {"error": f"comparison arm exited {r.returncode}: {r.stderr.strip().splitlines()[-1][:300] if r.stderr.strip() else ''}"}
    return rows


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
    ap.add_argument("--no-baseline", action="store_true", help="skip the in-run stock-PyTorch comparison arm")
    ap.add_argument("--no-presets", action="store_true", help="skip the other presets and the stand-alone exchange timing")
    ap.add_argument("--baseline-variants", default="eager,eager_amp,graph,graph_amp,graph_amp_cl")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm()
        return
    if args.warmup < 3:
        args.warmup = 3

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("for --gpus N > 1 launch with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
    clk = ClockSampler(local).start()        # NVML set-up happens here, not inside any timed window

    import torch
    import torch.distributed as dist

    from distributed_tensorflow_ibm_mnist_b200.data.synthetic import DeviceFeeder, SyntheticMnist
    from distributed_tensorflow_ibm_mnist_b200.models import oracle, presets
    from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        from distributed_tensorflow_ibm_mnist_b200.parallel.peer_comm import PeerComm
    B = args.batch

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

    def timed_steps(step, tail=None):
        """W warm-up steps (the last 3 after the host barrier: primed queues, device-aligned ranks), then EXACTLY K timed
        steps between two CUDA events; returns (max-over-ranks ms, host window) - no host work other than the step
        launches happens between the barrier and the stop event."""
        prime = 3
        for _ in range(args.warmup - prime):
            step()
        barrier()
        for _ in range(prime):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        e0.record()
        for _ in range(args.steps):
            step()
        if tail is not None:
            tail()
        e1.record()
        torch.cuda.synchronize()
        h1 = time.perf_counter()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)), (h0, h1)

    def measure(preset: str, *, with_e2e: bool, with_exchange: bool) -> dict:
        spec = presets.make_spec(preset)
        comm = PeerComm(dist.group.WORLD, dev) if world > 1 else None
        eng = NativeEngine(spec, B, device=dev, optimizer=args.optimizer, base_lr=1e-3, decay_rate=0.1,
                           decay_steps=presets.decay_steps(B), seed=1234, comm=comm, use_graph=not args.no_graph)
        eng.load_params(oracle.init_params(spec, seed=0))
        if comm is not None:
            eng.broadcast_params()

        # ---- device-timed steady state: one fresh batch per step out of a device pool larger than L2 (126 MB) -----
        pool_batches = max(8, (160 * 1024 * 1024) // (B * spec.image_pixels) + 1)
        host_pool = SyntheticMnist(B, spec.image_pixels, pool_batches=pool_batches, seed=rank, pin=False)

        class DevicePool:
            """The same packed batches, resident on the device (160 MB > L2): the feeder's copy stream moves one of them
            into the engine's input slot per step, beside the previous step's kernels - no host memory in this loop."""
            label_offset = host_pool.label_offset

            def __init__(self):
                self.packed = host_pool.packed.to(dev)
                self.i = 0

            def next_packed(self):
                self.i += 1
                return self.packed[(self.i - 1) % pool_batches]

        dev_feeder = DeviceFeeder(eng, DevicePool())
        for _ in range(2):              # set-up, not warm-up: both input slots capture their CUDA graph here
            dev_feeder.step()
        ms, window = timed_steps(dev_feeder.step)
        st = eng.metrics()
        out = {"preset": preset, "value": world * B * args.steps / (ms / 1e3), "ms_per_step": ms / args.steps,
               "launches_per_step": eng.launches_per_step, "window": window, "spec": spec,
               "pool_mb": pool_batches * B * spec.image_pixels // (1024 * 1024),
               "final_global_step": st["global_step"], "final_loss": st["cross_entropy"], "nan": st["nan"]}

        # ---- end to end through the public feeder: pinned H2D of every batch + D2H of every step's loss ---------
        if with_e2e:
            src = SyntheticMnist(B, spec.image_pixels, pool_batches=16, seed=100 + rank, pin=True)
            feeder = DeviceFeeder(eng, src)
            host_metrics = torch.zeros(4, dtype=torch.float32).pin_memory()   # last_xent, last_total, last_acc, last_lr
            metrics_dev = eng.state.view(torch.float32)[8:12]
            d2h_done, step_done = torch.cuda.Event(), torch.cuda.Event()
            d2h_stream = torch.cuda.Stream(device=dev)
            losses: list[float] = []
            started = [False]

            def e2e_step() -> None:
                feeder.step()
                step_done.record(torch.cuda.current_stream())
                if started[0]:
                    d2h_done.synchronize()                 # host reads the previous step's loss ...
                    losses.append(float(host_metrics[0]))
                started[0] = True
                # ... while this step's loss is on its way: the D2H copy waits for this step on its own stream, so the next
                # step's kernels are not queued behind it (the values it reads are only rewritten at the END of the next step)
                d2h_stream.wait_event(step_done)
                with torch.cuda.stream(d2h_stream):
                    host_metrics.copy_(metrics_dev, non_blocking=True)
                    d2h_done.record(d2h_stream)

            def e2e_tail() -> None:
                d2h_done.synchronize()
                losses.append(float(host_metrics[0]))

            ms_e2e, _ = timed_steps(e2e_step, tail=e2e_tail)
            out["e2e"] = {"value": world * B * args.steps / (ms_e2e / 1e3), "unit": "images/sec", "ms_per_step": ms_e2e / args.steps,
                          "h2d_bytes_per_step": feeder.h2d_bytes_per_step, "d2h_bytes_per_step": 16,
                          "host_losses_read": len(losses), "last_loss": losses[-1]}

        # ---- the fused gradient exchange alone (all-reduce + Adam + EMA + weight all-gather, ONE kernel) -------------
        if with_exchange and comm is not None:
            f, cfg = eng.flat, eng.cfg
            torch.cuda.synchronize()
            iters = 30
            for _ in range(3):
                comm.fused_allreduce_update(cfg, eng.state, f, 0, f.total, finish=True)
            barrier()
            for _ in range(3):
                comm.fused_allreduce_update(cfg, eng.state, f, 0, f.total, finish=True)
            x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x0.record()
            for _ in range(iters):
                comm.fused_allreduce_update(cfg, eng.state, f, 0, f.total, finish=True)
            x1.record()
            torch.cuda.synchronize()
            us = max_over_ranks(x0.elapsed_time(x1)) / iters * 1e3
            # bytes this kernel moves per GPU in its busier direction: gradients are fp32; the all-gather carries bf16 for
            # the tensor-core-only variables and fp32 for the rest
            s_grad = f.total * 4.0
            s_gather = sum((2.0 if v.tensor_core_only else 4.0) * ((v.size + 63) // 64 * 64) for v in spec.variables)
            n = world
            nvls = comm.use_multicast
            if nvls:   # multimem.ld_reduce pulls every rank's copy through the switch; multimem.st is replicated by it
                own = max(s_grad + s_gather / n, s_grad / n + s_gather)
            else:
                own = (n - 1) / n * (s_grad + s_gather)
            roof_us = own / NVLINK_BYTES_PER_SEC * 1e6
            out["fused_exchange"] = {"us": round(us, 2), "transport": "nvls" if nvls else "p2p", "gradient_bytes": int(s_grad),
                                     "own_bytes_busier_direction": int(own), "own_bytes_roofline_us": round(roof_us, 2),
                                     "fraction_of_roofline": round(roof_us / us, 3)}
        if comm is not None:
            torch.cuda.synchronize()
            comm.close()
        del eng
        torch.cuda.empty_cache()
        return out

    head = measure(args.preset, with_e2e=not args.no_e2e, with_exchange=not args.no_presets)
    clocks = clk.summary(*head["window"])
    extra: dict = {}
    if not args.no_presets:
        for p in ("reference", "wide", "baseline"):
            if p == args.preset:
                continue
            try:
                m = measure(p, with_e2e=False, with_exchange=(p == "wide"))
                extra[p] = {"images_per_sec": m["value"], "ms_per_step": m["ms_per_step"], "launches_per_step": m["launches_per_step"]}
                if "fused_exchange" in m:
                    extra[p]["fused_exchange"] = m["fused_exchange"]
            except Exception as e:  # noqa: BLE001 - the headline must survive a failure of the side measurements
                extra[p] = {"error": f"{type(e).__name__}: {e}"[:300]}
    clk.stop()

    baseline_rows: dict = {}
    if not args.no_baseline:
        barrier()
        baseline_rows = run_baseline_arm(args, args.preset, args.baseline_variants)
        barrier()

    if rank == 0:
        spec = head["spec"]
        value = head["value"]
        good = {k: v for k, v in baseline_rows.items() if isinstance(v, dict) and "images_per_sec" in v}
        warnings = []
        if good:
            best = max(good, key=lambda k: good[k]["images_per_sec"])
            base_value, base_src = good[best]["images_per_sec"], (
                "measured in this run: baseline/torch_ddp_baseline.py variant '%s' (fastest of %s); the reference publishes nothing"
                % (best, ",".join(sorted(good))))
            best_e2e = max(v["e2e_images_per_sec"] for v in good.values())
        else:
            base_value = BASELINE_MD_IMAGES_PER_SEC.get((args.preset, world)) if B == 512 else None
            base_src = "BASELINE.md section 4 table (comparison arm not measured in this run)"
            best_e2e = None
            if not args.no_baseline:
                warnings.append("comparison arm failed: %s" % json.dumps(baseline_rows)[:300])
        e2e = head.get("e2e")
        if e2e is not None:
            if best_e2e:
                e2e["vs_baseline_e2e"] = e2e["value"] / best_e2e
            if abs(e2e["value"] - value) / value > 0.10:
                warnings.append("value and e2e differ by %.0f %%" % (100.0 * abs(e2e["value"] - value) / value))
        out = {
            "metric": "MNIST-CNN train images/sec", "value": value, "unit": "images/sec", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (value / base_value) if base_value else None, "baseline_source": base_src,
            "baseline_measured": baseline_rows or None,
            "dtype": "bf16", "data": "synthetic MNIST-shaped uint8 28x28x%d, random-init weights" % spec.channels,
            "config": {"model": "MNIST CNN (%s preset: conv5x5-32, conv5x5-64, fc%d, fc10)" % (spec.preset, spec.fc1),
                       "global_batch": world * B, "batch_per_gpu": B, "seq_len": None, "parallelism": "dp%d" % world,
                       "optimizer": args.optimizer, "cuda_graph": not args.no_graph,
                       "l2_policy": "inputs cycled from a %d MB device pool (> 126 MB L2), one fresh batch per step" % head["pool_mb"],
                       "timing": "2 graph-capture steps + W warm-up steps (last 3 after the host barrier), then exactly K steps "
                                 "between CUDA events, max over ranks"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": head["launches_per_step"] * args.steps,
            "launches_per_step": head["launches_per_step"],
            "fused_exchange": head.get("fused_exchange"), "presets": extra or None,
            "final_global_step": head["final_global_step"], "final_loss": head["final_loss"], "nan": head["nan"],
            "warnings": warnings, "impl": "native",
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
