#!/usr/bin/env python
"""Joins tools/gemm_ncu.py's case list with the ncu raw CSV of the same run -> markdown + json for profiles/."""
import csv, json, sys
cases = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
rows = list(csv.reader(open(sys.argv[2])))
hdr, units, data = rows[0], rows[1], rows[2:]
def col(k): return hdr.index(k)
def val(r, k):
    v = float(r[col(k)].replace(",", "")); u = units[col(k)]
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3}.get(u, 1)
    return v * scale
assert len(cases) == len(data), (len(cases), len(data))
out = []
for c, r in zip(cases, data):
    t = val(r, "gpu__time_duration.sum")
    traffic = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    out.append(dict(name=c["name"], m=c["m"], n=c["n"], k=c["k"], us=round(t, 1), tflops=round(c["flops"] / t / 1e6),
                    tensor_pipe_pct=round(val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), 1),
                    dram_bytes=int(traffic), algorithmic_bytes=c["algorithmic_bytes"],
                    traffic_over_algorithmic=round(traffic / c["algorithmic_bytes"], 2),
                    dram_pct=round(val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                    l2_pct=round(val(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                    sm_mhz=round(val(r, "sm__cycles_active.avg") / t), regs=int(val(r, "launch__registers_per_thread"))))
json.dump(out, open(sys.argv[3], "w"), indent=1)
print("| GEMM (M,N,K) | us | TFLOP/s | tensor pipe % | DRAM MB | algorithmic MB | traffic/alg | DRAM % | L2 % | SM MHz |\n|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
for o in out:
    print(f"| {o['name']} ({o['m']},{o['n']},{o['k']}) | {o['us']} | {o['tflops']} | {o['tensor_pipe_pct']} | {o['dram_bytes']/1e6:.0f} | {o['algorithmic_bytes']/1e6:.0f} | {o['traffic_over_algorithmic']} | {o['dram_pct']} | {o['l2_pct']} | {o['sm_mhz']} |")
