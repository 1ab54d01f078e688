#!/usr/bin/env python
"""Warp-stall samples per barrier-delimited segment of the top kernel. usage: ncu_stalls.py rep.ncu-rep"""
import csv, subprocess, sys, io
src = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
h = rows[hi]; ia = h.index('Source'); ie = h.index('Instructions Executed')
cols = [i for i, c in enumerate(h) if c.startswith('stall_') and 'Not Issued' not in c]
seg = 0; acc = {}
for r in rows[hi + 1:]:
    if len(r) <= ie or not r[ie].isdigit(): continue
    d = acc.setdefault(seg, {})
    for i in cols:
        v = int(r[i] or 0)
        if v: d[h[i]] = d.get(h[i], 0) + v
    if 'BAR.SYNC' in r[ia] or 'EXIT' in r[ia]: seg += 1
tot = sum(sum(d.values()) for d in acc.values())
for s, d in acc.items():
    t = sum(d.values())
    if t < tot * 0.005: continue
    print(f"segment {s}: {t/tot*100:5.1f}% of samples | " + ", ".join(f"{k[6:]} {v/t*100:.0f}%" for k, v in sorted(d.items(), key=lambda x: -x[1])[:7]))
