#!/usr/bin/env python
"""prints the ab_probe.py logs given on the command line side by side (us per kernel, ratio to the first)"""
import sys
cols = []
for f in sys.argv[1:]:
    d = {}
    for line in open(f):
        if line.startswith("#") or " us " not in line: continue
        name = line[11:43].strip(); d[name] = float(line[43:52])
    cols.append(d)
names = list(cols[0])
print(f"{'kernel':32s}" + "".join(f"{f.split('/')[-1][:14]:>16s}" for f in sys.argv[1:]))
for n in names:
    base = cols[0][n]
    print(f"{n:32s}" + "".join(f"{c.get(n, float('nan')):9.1f} {c.get(n, float('nan')) / base:5.2f}x" for c in cols))
