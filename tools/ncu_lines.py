#!/usr/bin/env python3
"""Executed warp instructions and stall samples per CUDA source line of one profiled kernel, from
`ncu -i X.ncu-rep --page source --csv --print-source sass,cuda` (needs -lineinfo and --import-source on).
  python tools/ncu_lines.py X.ncu-rep [top_n] [kernel_regex]"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"]
if len(sys.argv) > 3:
    cmd += ["-k", "regex:" + sys.argv[3]]
txt = subprocess.run(cmd, capture_output=True, text=True).stdout
cur_file, hdr = "?", None
per_line = defaultdict(lambda: [0, 0, ""])
per_file = defaultdict(lambda: [0, 0])
for r in csv.reader(io.StringIO(txt)):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        iex, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= iex or not r[0].isdigit():
        continue
    ex = int(r[iex]) if r[iex].isdigit() else 0
    smp = int(r[ismp]) if r[ismp].isdigit() else 0
    k = (cur_file, int(r[0]))
    per_line[k][0] += ex
    per_line[k][1] += smp
    per_line[k][2] = r[1].strip()[:90]
    per_file[cur_file][0] += ex
    per_file[cur_file][1] += smp
tot = sum(v[0] for v in per_file.values()) or 1
tots = sum(v[1] for v in per_file.values()) or 1
print("total warp instructions %d, samples %d" % (tot, tots))
for f, v in sorted(per_file.items(), key=lambda kv: -kv[1][0]):
    print("  %-22s %6.2f %% instr  %6.2f %% samples" % (f, 100 * v[0] / tot, 100 * v[1] / tots))
print()
for (f, ln), v in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-18s %5d  %5.2f %% instr %5.2f %% smp  %s" % (f, ln, 100 * v[0] / tot, 100 * v[1] / tots, v[2]))
