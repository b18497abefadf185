#!/usr/bin/env python3
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count, average and share."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    n = r[ki].split("(")[0].split("::")[-1][:40]
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    agg.setdefault(n, []).append(v)
tot = sum(sum(v) for v in agg.values())
for n, v in agg.items():
    print(f"{n:42s} n={len(v):3d} avg={sum(v)/len(v):9.1f} us  share={100*sum(v)/tot:5.1f}%")
