#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: python tools/launch_summary.py file.csv"""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    agg.setdefault(row["Kernel Name"][:80], []).append(float(row["Metric Value"].replace(",", "")))
for k, v in agg.items():
    print(f"{len(v):4d} x avg {sum(v) / len(v) / 1e3:9.1f} us  min {min(v) / 1e3:8.1f} us  {k}")
