#!/usr/bin/env python
"""Executed warp instructions per CUDA source line of the profiled kernel (needs -lineinfo and --import-source on).
usage: ncu_lines.py rep.ncu-rep [top]"""
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(io.StringIO(out)))
fname, lines, ie, isamp = "?", [], None, None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif len(r) > 8 and r[0] == "Line No":
        ie, isamp = r.index("Instructions Executed"), r.index("# Samples")
    elif ie is not None and len(r) > ie and r[0].isdigit():
        try:
            lines.append((int(r[ie]), int(r[isamp] or 0), fname, int(r[0]), r[1].strip()))
        except ValueError:
            pass
tot = sum(x[0] for x in lines) or 1
stot = sum(x[1] for x in lines) or 1
print(f"total warp instructions {tot}, stall samples {stot}")
for n, sm, f, ln, src in sorted(lines, reverse=True)[:top]:
    print(f"{n / tot * 100:5.1f}% inst {sm / stot * 100:5.1f}% samp  {f}:{ln}  {src[:110]}")
