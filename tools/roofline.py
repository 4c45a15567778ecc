#!/usr/bin/env python
"""Roofline table of the training step: algorithmic FLOPs and bytes of every kernel against the MEASURED peaks.

    python tools/roofline.py [--batch 512] [--ncu profiles/step_kernels_r1_final.json] > profiles/roofline_r1.md

Durations come from the ncu summary (`tools/summarize_ncu.py --json`, gpu__time_duration per kernel, cold caches) - the
replayed graph runs the same kernels warm and overlapped, so the step total is quoted from the bench JSON instead.
Peaks: MEASURED_PEAKS.json (driver-written: copy bandwidth and cuBLAS bf16 throughput on this box); fallback = the
profiling recipe's nominal numbers.  frac = max(time at the tensor peak, time at the HBM peak) / measured time, i.e. how
far the kernel is from the faster of its two roofs.  Pure arithmetic - runs on the CPU box.
"""
import argparse
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"] * 1e9, d.get("bf16_tflops_sustained", d["bf16_tflops"]) * 1e12, "MEASURED_PEAKS.json"
    return 7.7e12, 2.25e15, "nominal (MEASURED_PEAKS.json absent)"


def kernels(B: int, fc1: int = 1024):
    """name -> (ncu kernel-name substring, grid hint, FLOPs, minimal HBM bytes) for the baseline preset."""
    px1, px2 = 28 * 28, 14 * 14
    p1 = B * px2 * 32 * 2            # pooled conv1 activation, bf16 (plane layout adds 30 % padding: 16x16 grid)
    p1_planes = B * 4 * 4096
    dz2_planes = B * 8 * 4096
    w3 = 3136 * fc1
    n_params = 800 + 32 + 51200 + 64 + w3 + fc1 + fc1 * 10 + 10
    bucket0 = w3 + fc1 + fc1 * 10 + 10
    upd = lambda n: n * (4 * 5 + 4 * 4 + 2)          # read w,g,m,v,ema ; write w,m,v,ema ; write bf16
    return [
        ("conv1 (fwd+pool)", "conv1_", None, 2 * B * px1 * 25 * 32, B * px1 + p1_planes + B * px2 * 32),
        ("conv2_fwd (+pool)", "conv2_fwd_kernel", None, 2 * B * px2 * 800 * 64, p1_planes + 102400 + B * 49 * 64 * 3),
        ("fc1 GEMM (+bias,ReLU,dropout)", "gemm_bf16_kernel<0, 1, 8>", None, 2 * B * 3136 * fc1, B * 3136 * 2 + w3 * 2 + B * fc1 * 2),
        ("head_forward (+dH)", "head_forward_kernel", None, 2 * 2 * B * fc1 * 10, B * fc1 * 2 * 2 + fc1 * 10 * 4),
        ("dP2 GEMM (+mask, un-pool)", "gemm_bf16_kernel<0, 0, 4>", 196, 2 * B * 3136 * fc1, B * fc1 * 2 + w3 * 2 + B * 3136 * 2 * 2 + dz2_planes),
        ("head_backward", "head_backward_kernel", None, 2 * B * fc1 * 10, B * fc1 * 2 * 2 + fc1 * 10 * 4),
        ("dW3 GEMM", "gemm_bf16_kernel<1, 1, 4>", 400, 2 * B * 3136 * fc1, B * 3136 * 2 + B * fc1 * 2 + w3 * 4),
        ("colsum (conv2 bias grad)", "colsum_bf16_kernel", None, B * 3136, B * 3136 * 2),
        ("fused_update fc/head bucket", "fused_update", 148, 12 * bucket0, upd(bucket0)),
        ("conv2_dgrad (+ReLU mask)", "conv2_dgrad_kernel", None, 2 * B * px2 * 1600 * 32, dz2_planes + p1_planes + 102400 + p1),
        ("conv2_wgrad", "conv2_wgrad_kernel", None, 2 * B * px2 * 800 * 64, p1_planes + dz2_planes + 51200 * 4),
        ("conv1_wgrad", "conv1_wgrad_kernel", None, 2 * B * px2 * 25 * 32, B * px1 + p1 + B * px2 * 32),
        # round 2: both weight gradients in one kernel (each y-half of the grid reads every image's planes)
        ("conv_wgrad_fused (conv2 + conv1)", "conv_wgrad_fused_kernel", None, 2 * B * px2 * 800 * 64 + 2 * B * px1 * 25 * 32,
         2 * (p1_planes + dz2_planes) + B * px1 + p1 + B * px2 * 32 + 51200 * 4),
        ("fused_update conv bucket + bookkeeping", "fused_update", 51, 12 * (n_params - bucket0), upd(n_params - bucket0)),
    ]


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--ncu", default=os.path.join(ROOT, "profiles", "step_kernels_r1_final.json"))
    ap.add_argument("--bench", default=os.path.join(ROOT, "profiles", "bench_1gpu_baseline_r1_final.json"))
    args = ap.parse_args()
    hbm, tc, src = peaks()
    rows = json.load(open(args.ncu))
    print(f"# Roofline of the training step (baseline preset, batch {args.batch}) against {src}\n")
    print(f"Peaks: HBM copy {hbm / 1e9:.1f} GB/s, bf16 tensor {tc / 1e12:.1f} TFLOP/s (sustained cuBLAS).  Durations: ncu "
          "`gpu__time_duration` per kernel with cold caches.  `t_roof` = the larger of FLOPs/peak and bytes/peak.\n")
    print("| kernel | FLOPs | min. HBM bytes | t at tensor peak | t at HBM peak | measured | t_roof / measured | ncu DRAM read | ncu tensor-pipe % |")
    print("|---|---|---|---|---|---|---|---|---|")
    tot_f = tot_b = tot_t = 0.0
    used = set()
    for name, pat, grid, flops, byts in kernels(args.batch):
        cand = [i for i, r in enumerate(rows) if pat in r["kernel"] and i not in used and (grid is None or int(r.get("grid", 0)) == grid)]
        if not cand:
            continue            # kernel not part of this round's schedule
        i = cand[0]
        # the report holds two consecutive steps: mark both launches of this kernel as consumed
        for j in cand[:1]:
            used.add(j)
        r = rows[i]
        t = r["dur_us"]
        t_tc, t_hbm = flops / tc * 1e6, byts / hbm * 1e6
        roof = max(t_tc, t_hbm)
        tot_f += flops; tot_b += byts; tot_t += t
        print(f"| {name} | {flops / 1e9:.2f} G | {byts / 1e6:.1f} MB | {t_tc:.2f} µs | {t_hbm:.2f} µs | {t:.1f} µs | {100 * roof / t:.0f} % | "
              f"{r.get('dram_rd_MB', '')} MB | {r.get('tensor_pct', '')} |")
    roof_tot = max(tot_f / tc, tot_b / hbm) * 1e6
    print(f"| **sum** | {tot_f / 1e9:.1f} G | {tot_b / 1e6:.0f} MB | {tot_f / tc * 1e6:.1f} µs | {tot_b / hbm * 1e6:.1f} µs | {tot_t:.0f} µs (serialised, cold) | {100 * roof_tot / tot_t:.0f} % | | |")
    if os.path.exists(args.bench):
        b = json.load(open(args.bench))
        step = b["ms_per_step"] * 1e3
        print(f"\nReplayed graph (`{os.path.basename(args.bench)}`): **{step:.1f} µs per step** = {b['value'] / 1e6:.2f} M images/s; "
              f"the step's combined roof (all FLOPs at the tensor peak + all bytes at the HBM peak, no overlap) is "
              f"{(tot_f / tc + tot_b / hbm) * 1e6:.1f} µs, i.e. the step runs at {100 * (tot_f / tc + tot_b / hbm) * 1e6 / step:.0f} % of it.")
    print("\nReading (round 1 text, still valid): with 3.5 images per SM every kernel is dominated by fixed latencies (launch, prologue, first loads, "
          "epilogue tail) - none is within 2x of either roof; the optimizer updates are the closest (memory bound by "
          "construction; the fc/head one is deliberately capped to one CTA per SM so that it runs beside conv2_dgrad - alone "
          "and uncapped it takes 24.5 µs = 76 % of its HBM roof).  The DRAM-read column shows that most operand traffic is "
          "served by the 126 MB L2.")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
