#!/usr/bin/env python
"""Aggregates an `ncu --page source --print-source cuda,sass --csv` dump by CUDA source line:
share of warp-stall samples and of executed instructions per line.  usage: ncu_lines.py dump.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
lines = {}
fname = ""
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if len(r) >= 3 and r[0] == "Line No":
        hdr = r
        ix_inst = hdr.index("Instructions Executed")
        ix_samp = hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or not r or not r[0].isdigit():
        continue
    def f(x):
        try:
            return float(x)
        except ValueError:
            return 0.0
    key = (fname, int(r[0]))
    s, i, src = lines.get(key, (0.0, 0.0, r[1]))
    lines[key] = (s + f(r[ix_samp]), i + f(r[ix_inst]), src)
ts = sum(v[0] for v in lines.values())
ti = sum(v[1] for v in lines.values())
print(f"total samples {ts:.0f}, total warp instructions {ti:.0f}")
for (fn, ln), (s, i, src) in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{fn}:{ln:<4d} samples {100 * s / ts:5.1f}%  inst {100 * i / ti:5.1f}%  {src.strip()[:100]}")
