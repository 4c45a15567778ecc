#!/usr/bin/env python
"""Condense an ncu report (`ncu -i x.ncu-rep --page raw --csv`) into one row per launch.

Usage:  ncu -i gpurun_out/step.ncu-rep --page raw --csv | python tools/summarize_ncu.py [--json out.json] [--md out.md]

Columns: duration, DRAM read/write bytes, DRAM % of peak, tensor-pipe % of peak, achieved occupancy (warps), registers,
dynamic+static shared memory, grid x block, executed instructions, L2 hit rate and the top warp-stall reason.
Only the metric names are looked up; nothing is recomputed, so the numbers are ncu's own.
"""
import argparse
import csv
import json
import re
import sys

COLS = [
    ("dur_us", r"^gpu__time_duration\.sum$"),
    ("dram_rd_MB", r"^dram__bytes_read\.sum$"),
    ("dram_wr_MB", r"^dram__bytes_write\.sum$"),
    ("dram_pct", r"^gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed$"),
    ("tensor_pct", r"^sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_active$"),
    ("tensor_pct_elapsed", r"^sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_elapsed$"),
    ("sm_pct", r"^sm__throughput\.avg\.pct_of_peak_sustained_elapsed$"),
    ("l2_hit_pct", r"^lts__t_sector_hit_rate\.pct$"),
    ("warps_active_pct", r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$"),
    ("issue_active_pct", r"^smsp__issue_active\.avg\.pct_of_peak_sustained_active$"),
    ("regs", r"^launch__registers_per_thread$"),
    ("smem_dyn_KB", r"^launch__shared_mem_per_block_dynamic$"),
    ("smem_static_KB", r"^launch__shared_mem_per_block_static$"),
    ("grid", r"^launch__grid_size$"),
    ("block", r"^launch__block_size$"),
    ("inst", r"^smsp__inst_executed\.sum$"),
    ("local_spill_B", r"^smsp__inst_executed_op_local_(ld|st)\.sum$"),
]
UNIT_SCALE = {"Kbyte": 1 / 1024.0, "Mbyte": 1.0, "Gbyte": 1024.0, "byte": 1 / 1048576.0,
              "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json")
    ap.add_argument("--md")
    args = ap.parse_args()
    rows = [r for r in csv.reader(sys.stdin) if r]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units, data = rows[start], rows[start + 1], rows[start + 2:]
    name_col = hdr.index("Kernel Name")
    stall_cols = [i for i, h in enumerate(hdr) if re.match(r"^smsp__average_warps?_issue_stalled_.*_per_issue_active\.ratio$", h)
                  or re.match(r"^smsp__average_warp_latency_issue_stalled_.*\.ratio$", h)]
    out = []
    for r in data:
        rec = {"kernel": re.sub(r"\(.*", "", r[name_col])}
        for key, pat in COLS:
            idx = [i for i, h in enumerate(hdr) if re.match(pat, h)]
            if not idx:
                continue
            vals = [num(r[i]) for i in idx]
            if any(v is None for v in vals):
                continue
            v = sum(vals)
            u = units[idx[0]]
            if key.endswith("_MB") or key.endswith("_us"):
                v *= UNIT_SCALE.get(u, 1.0)
            if key.endswith("_KB"):
                v *= {"byte": 1 / 1024.0, "Kbyte": 1.0, "Mbyte": 1024.0}.get(u, 1.0)
            rec[key] = round(v, 3)
        stalls = sorted(((num(r[i]) or 0.0, hdr[i]) for i in stall_cols), reverse=True)[:2]
        rec["top_stalls"] = [re.sub(r".*issue_stalled_(.*?)(_per_issue_active)?\.ratio", r"\1", h) + f"={v:.1f}" for v, h in stalls if v > 0]
        out.append(rec)
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)
    keys = ["kernel", "dur_us", "dram_rd_MB", "dram_wr_MB", "dram_pct", "tensor_pct", "sm_pct", "l2_hit_pct", "warps_active_pct",
            "issue_active_pct", "regs", "smem_dyn_KB", "grid", "block", "top_stalls"]
    lines = ["| " + " | ".join(keys) + " |", "|" + "---|" * len(keys)]
    for rec in out:
        lines.append("| " + " | ".join(str(rec.get(k, "")) if k != "top_stalls" else ", ".join(rec.get(k, [])) for k in keys) + " |")
    text = "\n".join(lines)
    if args.md:
        open(args.md, "w").write(text + "\n")
    print(text)


if __name__ == "__main__":
    main()
