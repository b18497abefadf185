#!/usr/bin/env python3
"""Reads `ncu --page source --csv` for one kernel and prints executed warp instructions per
code region (regions split at RET/EXIT), with a short opcode signature, to attribute time to
the out-of-line device functions."""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, iex, ith, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), \
    hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
regions, cur = [], []
for r in rows[2:]:
    if len(r) <= iex or not r[iex].isdigit():
        continue
    cur.append(r)
    op = r[isrc].split()[0] if r[isrc].split() else ""
    if op.startswith("@"):
        op = r[isrc].split()[1]
    if op.startswith("RET") or op.startswith("EXIT"):
        regions.append(cur)
        cur = []
if cur:
    regions.append(cur)
tot = sum(int(r[iex]) for reg in regions for r in reg)
tots = sum(int(r[ismp]) for reg in regions for r in reg)
print("total warp instructions", tot, "samples", tots)
for reg in regions:
    ex = sum(int(r[iex]) for r in reg)
    th = sum(int(r[ith]) for r in reg)
    smp = sum(int(r[ismp]) for r in reg)
    if ex == 0:
        continue
    ops = Counter()
    for r in reg:
        t = r[isrc].split()
        o = t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "")
        ops[o.split(".")[0]] += 1
    sig = " ".join(f"{k}{v}" for k, v in ops.most_common(7))
    print(f"{reg[0][ia][-5:]} n={len(reg):5d} exec={ex:11d} {100*ex/tot:5.1f}%  samples {100*smp/max(tots,1):5.1f}%  lanes={th/max(ex,1):5.1f}  {sig}")
