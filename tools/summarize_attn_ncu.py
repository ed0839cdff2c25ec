#!/usr/bin/env python
"""ncu --set full capture of the attention kernels -> markdown table for profiles/ (tensor pipe, XU (MUFU) pipe, FMA / ALU
pipes, issue slots, DRAM bytes, registers).  usage: ncu -i rep --page raw --csv | python tools/summarize_attn_ncu.py B T H d"""
import csv, sys
B, T, H, d = (int(a) for a in sys.argv[1:5])
rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
def val(r, k):
    if k not in hdr: return float("nan")
    v = float(r[hdr.index(k)].replace(",", "")); u = units[hdr.index(k)]
    return v * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3}.get(u, 1)
fl_fwd = 4.0 * T * T * d * B * H * 0.5
print("| kernel | us | TFLOP/s (causal, algorithmic) | tensor pipe % | XU (MUFU) pipe % | FMA pipe % | ALU pipe % | issue slots % | warps active % | DRAM MB | regs | SM MHz |")
print("|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
for r in data:
    name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")
    t = val(r, "gpu__time_duration.sum")
    fl = fl_fwd if "fwd" in name else (2.5 * fl_fwd if "bwd_tc" in name else 0)
    print(f"| `{name}` | {t:.1f} | {fl / t / 1e6:.0f} | {val(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active'):.1f} | {val(r, 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active'):.1f} | {val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | {(val(r, 'dram__bytes_read.sum') + val(r, 'dram__bytes_write.sum')) / 1e6:.0f} | "
          f"{int(val(r, 'launch__registers_per_thread'))} | {val(r, 'sm__cycles_active.avg') / t:.0f} |")
