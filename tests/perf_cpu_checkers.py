#!/usr/bin/env python
"""Timing of the CPU checkers (oracle/) on one host core, for the secondary lines next to profiles/bench_paths.py.
Lives under tests/ because only tests/, smoke() and bench.py's CPU-baseline legs may execute oracle/.
usage: python tests/perf_cpu_checkers.py   -> one JSON line per checker"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libadsb_b200 import synth, synth978   # noqa: E402
from oracle import oracle1090b as B, oracle978 as O, oraclecpr as P   # noqa: E402

BUF = 262144


def rate(n, fn):
    t0 = time.perf_counter()
    fn()
    return n / (time.perf_counter() - t0) / 1e6


iq = synth.config_b(seconds=2.0, seed=0x1090B002)[0][:16 * BUF]
o = B.Oracle1090B()
print(json.dumps({"path": "B cpu checker (1 core)", "msamples_s": rate(iq.size / 2, lambda: [o.handle_data(iq[k:k + BUF]) for k in range(0, iq.size, BUF)])}))
iq = synth978.config5(seconds=2.0)[0][:16 * BUF]
u = O.Oracle978()
print(json.dumps({"path": "C cpu checker (1 core)", "msamples_s": rate(iq.size / 2, lambda: [u.handle_data(iq[k:k + BUF]) for k in range(0, iq.size, BUF)])}))
rng = np.random.default_rng(3)
raw = rng.integers(0, 256, (1 << 20, 14), dtype=np.uint8)
raw[::2, 0] = (17 << 3) | 5
print(json.dumps({"path": "B fields cpu checker (1 core)", "mrecords_s": rate(raw.shape[0], lambda: o.fields_bulk(raw, raw[:, 0] >> 3))}))
p5 = rng.integers(0, 131072, (1 << 20, 5)).astype(np.int32)
print(json.dumps({"path": "B cpr cpu checker (1 core)", "mpairs_s": rate(p5.shape[0], lambda: o.cpr_bulk(p5))}))
recs = P.random_records(1 << 20, 5)
print(json.dumps({"path": "A cpr cpu checker (1 core)", "mrecords_s": rate(recs.size, lambda: P.decode(recs))}))
