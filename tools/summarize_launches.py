#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: mean duration per kernel over the last K launches."""
import collections
import csv
import sys

path = sys.argv[1]
last = int(sys.argv[2]) if len(sys.argv) > 2 else 4
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    try:
        v = float(row["Metric Value"].replace(",", ""))
    except (ValueError, KeyError):
        continue
    unit = row.get("Metric Unit", "ns")
    v = v / 1000 if unit == "ns" else v * 1000 if unit == "ms" else v
    agg.setdefault(row["Kernel Name"][:64], []).append(v)
total = 0.0
for k, v in agg.items():
    per = sum(v[-last:]) / min(last, len(v))
    if k.startswith(("dtm::", "void dtm::")):
        total += per
    print(f"{per:9.1f} us  n={len(v):3d}  {k}")
print(f"{total:9.1f} us  sum over dtm kernels (serialised, cold caches)")
