#!/usr/bin/env python
"""Executed warp instructions of the top kernel per barrier-delimited segment, with the opcode mix of each segment.
usage: ncu_instmix.py rep.ncu-rep [units]   (units = warp-tiles of the launch, to print per-warp-tile counts)"""
import collections
import csv
import io
import subprocess
import sys

src = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
units = float(sys.argv[2]) if len(sys.argv) > 2 else None
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
h = rows[hi]
ia, ie, it = h.index("Source"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
seg = 0
warp, lanes, ops = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for r in rows[hi + 1:]:
    if len(r) <= ie or not r[ie].isdigit():
        continue
    n = int(r[ie])
    t = r[ia].split()
    op = t[1] if t[0].startswith("@") else t[0]
    warp[seg] += n
    lanes[seg] += int(r[it] or 0)
    ops[seg][op] += n
    if "BAR.SYNC" in r[ia] or "EXIT" in r[ia]:
        seg += 1
tot = sum(warp.values())
print(f"executed warp instructions: {tot}")
for s, n in warp.items():
    if n < tot * 0.01:
        continue
    per = f", {n / units:8.1f} per warp-tile" if units else ""
    print(f"segment {s:2d}: {n / tot * 100:5.1f}%{per}, {lanes[s] / n:4.1f} active lanes | " +
          ", ".join(f"{k} {v / n * 100:.0f}%" for k, v in ops[s].most_common(8)))
