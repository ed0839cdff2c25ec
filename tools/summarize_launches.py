#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
usage: tools/summarize_launches.py gpurun_out/launches.csv [title] > profiles/<name>.md"""
import collections
import csv
import re
import sys

path = sys.argv[1]
title = sys.argv[2] if len(sys.argv) > 2 else path
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    key = re.sub(r"\(.*", "", r["Kernel Name"])[:80]
    agg[key][0] += 1
    agg[key][1] += v
    tot += v
print(f"# {title}\n")
print(f"{sum(a[0] for a in agg.values())} launches, {tot / 1e3:.2f} ms of device time "
      "(ncu per-launch times: cold cache, serialised -- compare SHARES, not absolutes)\n")
print("| share | total us | launches | avg us | kernel |\n|---:|---:|---:|---:|---|")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {t / tot * 100:.2f}% | {t:.1f} | {n} | {t / n:.1f} | `{k}` |")
