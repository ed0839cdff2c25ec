#!/usr/bin/env python
"""Joins tools/hbm_ncu.py's case list with the ncu raw CSV of the same run -> markdown + json for profiles/.
A case may launch several kernels (AdamW: one per group); rows are matched in launch order by the counts below."""
import csv, json, sys
cases = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
rows = list(csv.reader(open(sys.argv[2])))
hdr, units, data = rows[0], rows[1], rows[2:]
def col(k): return hdr.index(k)
def val(r, k):
    v = float(r[col(k)].replace(",", "")); u = units[col(k)]
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3}.get(u, 1)
    return v * scale
out = []
for r in data:
    t = val(r, "gpu__time_duration.sum")
    traffic = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    out.append(dict(kernel=r[col("Kernel Name")][:60], us=round(t, 1), dram_bytes=int(traffic), dram_gbs=round(traffic / t / 1e3),
                    dram_pct=round(val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                    l2_pct=round(val(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                    issue_pct=round(val(r, "sm__issue_active.avg.pct_of_peak_sustained_elapsed"), 1),
                    warps_pct=round(val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
                    grid=r[col("launch__grid_size")], block=r[col("launch__block_size")], regs=int(val(r, "launch__registers_per_thread"))))
json.dump(dict(cases=cases, launches=out), open(sys.argv[3], "w"), indent=1)
print("| kernel | us | DRAM MB | DRAM GB/s | DRAM % | L2 % | issue % | warps active % | grid x block | regs |\n|---|---:|---:|---:|---:|---:|---:|---:|---|---:|")
for o in out:
    print(f"| {o['kernel']} | {o['us']} | {o['dram_bytes']/1e6:.0f} | {o['dram_gbs']} | {o['dram_pct']} | {o['l2_pct']} | {o['issue_pct']} | {o['warps_pct']} | {o['grid']}x{o['block']} | {o['regs']} |")
