#!/usr/bin/env python
"""Opcode mix of one barrier-delimited segment of the top kernel. usage: ncu_segments.py rep.ncu-rep SEGMENT [--dump]"""
import csv, subprocess, sys, io
rep, want = sys.argv[1], int(sys.argv[2])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
h = rows[hi]; ia = h.index('Source'); ie = h.index('Instructions Executed')
body = [r for r in rows[hi + 1:] if len(r) > ie and r[ie].isdigit()]
seg = 0; ops = {}; tot = 0; lines = []
for r in body:
    if seg == want:
        t = r[ia].split(); op = (t[1] if t[0].startswith('@') else t[0])
        ops[op] = ops.get(op, 0) + int(r[ie]); tot += int(r[ie]); lines.append((r[ie], r[ia]))
    if 'BAR.SYNC' in r[ia] or 'EXIT' in r[ia]: seg += 1
for o, v in sorted(ops.items(), key=lambda x: -x[1])[:30]: print(f"{o:28s} {v/tot*100:5.1f}%")
if '--dump' in sys.argv:
    for e, a in lines: print(e, a)
