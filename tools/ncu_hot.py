#!/usr/bin/env python3
"""Hot code footprint of each profiled kernel: SASS instructions executed more than `thr` times, grouped into address
ranges, from `ncu -i X.ncu-rep --page source --csv --print-source sass` (what has to stay in the instruction cache).
  python tools/ncu_hot.py X.ncu-rep [kernel_regex] [thr]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"]
if len(sys.argv) > 2 and sys.argv[2]:
    cmd += ["-k", "regex:" + sys.argv[2]]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 50000
txt = subprocess.run(cmd, capture_output=True, text=True).stdout
kern, cur, hdr = [], None, None
for r in csv.reader(io.StringIO(txt)):
    if r and r[0] in ("Kernel Name", "Function Name"):
        cur = []
        kern.append((r[1][:70], cur))
        continue
    if r and r[0] == "Address":
        hdr = r
        isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr and cur is not None and len(r) > iex and r[0].startswith("0x"):
        cur.append((int(r[0], 16), int(r[iex]) if r[iex].isdigit() else 0, r[isrc], int(r[ismp]) if r[ismp].isdigit() else 0))
seen = set()
for name, execs in kern:
    if not execs or name in seen:
        continue
    seen.add(name)
    base, tot = execs[0][0], len(execs)
    hot = [(a - base, e, s, m) for a, e, s, m in execs if e > thr]
    print("%s\n  total %.1f KB, hot %.1f KB, %.0f M warp instructions" % (name, tot * 16 / 1024, len(hot) * 16 / 1024, sum(e for _, e, _, _ in execs) / 1e6))
    ranges, start = [], None
    for a, e, s, m in hot:
        if start is None:
            start, prev, cnt, ex, sm = a, a, 1, e, m
        elif a - prev <= 16 * 8:
            prev, cnt, ex, sm = a, cnt + 1, ex + e, sm + m
        else:
            ranges.append((start, prev, cnt, ex, sm))
            start, prev, cnt, ex, sm = a, a, 1, e, m
    if start is not None:
        ranges.append((start, prev, cnt, ex, sm))
    for r in ranges:
        print("    %#7x-%#7x %4d instr %5.1f KB  execs %7.1f M  samples %6d" % (r[0], r[1], r[2], r[2] * 16 / 1024, r[3] / 1e6, r[4]))
