#!/usr/bin/env python
"""Top stall sites by CUDA source line from `ncu -i rep --page source --csv --print-source cuda,sass --kernel-name regex:K`.
usage: python tools/ncu_source_lines.py file.csv [top_n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname, hdr, out = "", None, []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].strip():
        out.append((fname, r))
i_s, i_ex = hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
def num(x):
    try: return int(x)
    except ValueError: return 0
tot = sum(num(r[i_s]) for _, r in out)
print("total samples", tot, "warp instructions", sum(num(r[i_ex]) for _, r in out))
agg = {}
for _, r in out:
    for i in stalls:
        agg[hdr[i]] = agg.get(hdr[i], 0) + num(r[i])
print("by reason:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for f, r in sorted(out, key=lambda fr: -num(fr[1][i_s]))[:top_n]:
    rs = sorted(((hdr[i][6:], num(r[i])) for i in stalls), key=lambda kv: -kv[1])[:2]
    print(f"{num(r[i_s]):6d} {100 * num(r[i_s]) / max(tot, 1):5.1f}% ex={r[i_ex]:>9s} {f}:{r[0]:<5s} {r[1].strip()[:95]:95s} {rs}")
