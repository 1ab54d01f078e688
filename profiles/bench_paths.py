#!/usr/bin/env python
"""Secondary timing lines (not the headline metric): Path B (libadsb's own 2.0 MS/s handler) and Path C
(UAT 978) through their C-ABI entry points with HOST buffers.  The CPU checkers are timed by
tests/perf_cpu_checkers.py (only tests/ may execute oracle/).
usage: python profiles/bench_paths.py [seconds]   -> one JSON line per path"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libadsb_b200 import adsbcu, synth, synth978   # noqa: E402

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
BUF = 262144   # RTLSDR.hpp:55: bytes per HandleData call in live mode


def timed(fn, reps=3):
    fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        n = fn()
        t.append(time.perf_counter() - t0)
    return min(t), n


# ---- Path B: one HandleData call per 262144-byte buffer, and one call per 8 MB (the largest the reference's
# fixed magnitude vector allows, ADSB1090.cpp:109)
iq = synth.config_b(seconds=seconds, seed=0x1090B002)[0]
d = adsbcu.Demod1090()
for name, step in (("262144-byte buffers", BUF), ("8 MiB buffers", 8 << 20)):
    out_b = np.zeros(1 << 17, dtype=adsbcu.MSG_B_DTYPE)      # reused: a fresh 4 MB array per call would be most of a live-size call

    def run_b():
        d.reset_b()
        return sum(len(d.handle_data_b(iq[k:k + step], out=out_b)) for k in range(0, iq.size, step))
    dt, n = timed(run_b)
    line = {"path": "B (ADSB1090Handler, 2.0 MS/s)", "feed": name, "msamples_s": iq.size / 2 / dt / 1e6, "messages": n,
            "realtime_factor": iq.size / 2 / dt / 2.0e6}
    print(json.dumps(line))

# ---- Path C
iq = synth978.config5(seconds=seconds)[0]
for name, step in (("262144-byte buffers", BUF), ("whole capture", iq.size)):
    out_c = np.zeros(1 << 14, dtype=adsbcu.UAT_DTYPE)

    def run_c():
        d.reset_uat()
        return sum(len(d.handle_data_uat(iq[k:k + step], out=out_c)) for k in range(0, iq.size, step))
    dt, n = timed(run_c)
    print(json.dumps({"path": "C (UAT978Handler, 2.083334 MS/s)", "feed": name, "msamples_s": iq.size / 2 / dt / 1e6,
                      "frames": n, "realtime_factor": iq.size / 2 / dt / 2083334.0}))

# ---- Path B next layer: bulk field decode and CPR (host buffers in, host buffers out) next to the CPU checker
rng = np.random.default_rng(3)
N = 4 << 20
rec = np.zeros(N, dtype=adsbcu.MSG_B_DTYPE)
rec["msg"] = rng.integers(0, 256, (N, 14), dtype=np.uint8)
rec["msg"][::2, 0] = (17 << 3) | 5
rec["msgtype"] = rec["msg"][:, 0] >> 3
rec["msgbits"] = 112
pairs = np.zeros(N, dtype=adsbcu.CPR_PAIR_DTYPE)
for f in ("even_lat", "even_lon", "odd_lat", "odd_lon"):
    pairs[f] = rng.integers(0, 131072, N)
dt, _ = timed(lambda: d.decode_fields_b(rec).size)
print(json.dumps({"path": "B fields (adsbcu_1090b_decode_fields)", "records": N, "mrecords_s": N / dt / 1e6,
                  "pcie_bytes_per_record": 112}))
dt, _ = timed(lambda: d.cpr_decode(pairs).size)
print(json.dumps({"path": "B cpr (adsbcu_1090b_cpr_decode)", "pairs": N, "mpairs_s": N / dt / 1e6, "pcie_bytes_per_pair": 32}))

# ---- Path A next layer: dump1090's CPR decoders, all four modes mixed
rng5 = np.random.default_rng(5)
recs = np.zeros(4 << 20, dtype=adsbcu.CPR_IN_DTYPE)
for f in ("even_lat", "even_lon", "odd_lat", "odd_lon"):
    recs[f] = rng5.integers(0, 131072, recs.size)
recs["mode"] = rng5.integers(0, 4, recs.size)
recs["fflag"] = rng5.integers(0, 2, recs.size)
recs["reflat"] = rng5.uniform(-90, 90, recs.size)
recs["reflon"] = rng5.uniform(-180, 180, recs.size)
dt, _ = timed(lambda: d.cpr_decode_a(recs).size)
print(json.dumps({"path": "A cpr (adsbcu_1090_cpr_decode)", "records": int(recs.size), "mrecords_s": recs.size / dt / 1e6,
                  "pcie_bytes_per_record": 64}))
