#!/usr/bin/env python
"""Opcode-level summary of an `ncu -i X.ncu-rep --page source --csv` (SASS view) dump: share of executed warp
instructions and of warp-stall samples per opcode, plus the hottest individual instructions.
usage: ncu_sass.py dump.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = None
sass = []
for r in rows:
    if len(r) > 3 and r[0] == "Address":
        hdr = r
        ix_s, ix_i = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
        continue
    if hdr and r and r[0].startswith("0x"):
        def f(x):
            try:
                return float(x)
            except ValueError:
                return 0.0
        sass.append((r[1].strip(), f(r[ix_s]), f(r[ix_i])))
ts = sum(s for _, s, _ in sass)
ti = sum(i for _, _, i in sass)
print(f"{len(sass)} SASS instructions, {ti:.0f} warp instructions executed, {ts:.0f} stall samples")
op, ops = collections.Counter(), collections.Counter()
for t, s, i in sass:
    w = t.split()
    o = w[1] if w[0].startswith("@") and len(w) > 1 else w[0]
    op[o] += i
    ops[o] += s
for o, c in op.most_common(top):
    print(f"{o:30s} inst {100 * c / ti:5.1f}%  samples {100 * ops[o] / ts:5.1f}%")
print("-- hottest instructions by samples")
for n, (t, s, i) in sorted(enumerate(sass), key=lambda kv: -kv[1][1])[:top]:
    print(f"#{n:<5d} samples {100 * s / ts:5.1f}%  inst {100 * i / ti:5.2f}%  {t[:90]}")
