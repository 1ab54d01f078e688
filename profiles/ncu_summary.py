#!/usr/bin/env python
"""Summarise an .ncu-rep: headline metrics + instruction share per barrier-delimited segment.
usage: ncu_summary.py report.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__cycles_elapsed.max',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed_op_shared_ld.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed']
for r in rows[2:]:
    print("kernel:", r[h.index('Kernel Name')] if 'Kernel Name' in h else '?')
    for i, k in enumerate(h):
        if k in keep or ('issue_stalled' in k and k.endswith('per_issue_active.ratio')):
            print(f"  {k} [{u[i]}] = {r[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
h = rows[hi]; ia = h.index('Source'); ie = h.index('Instructions Executed'); it = h.index('Thread Instructions Executed')
body = [r for r in rows[hi + 1:] if len(r) > it and r[ie].isdigit()]
tot = sum(int(r[ie]) for r in body)
print(f"total warp instructions {tot}")
seg = acc = tacc = 0; ops = {}
for r in body:
    v = int(r[ie]); acc += v; tacc += int(r[it])
    t = r[ia].split(); op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    ops[op] = ops.get(op, 0) + v
    if 'BAR.SYNC' in r[ia] or 'EXIT' in r[ia]:
        if acc: print(f"  segment {seg}: {acc/tot*100:5.1f}% of instructions, {tacc/max(acc,1):4.1f} active threads/inst (ends at {r[ia].strip()[:36]})")
        seg += 1; acc = tacc = 0
print("  opcode mix:", ", ".join(f"{o} {v/tot*100:.1f}%" for o, v in sorted(ops.items(), key=lambda x: -x[1])[:16]))
