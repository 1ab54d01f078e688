"""TEST INFRASTRUCTURE ONLY.  Random-case search on the CPU for Path B: scanb_kernel's source on the SIMT emulator against the
restatement of ADSB1090.cpp (oracle/oracle1090b.cpp), HandleData call by call, over traffic, noise, bit errors, Address/Parity
shares and ragged call sizes.  usage: python tests/emu/fuzz_pathb.py [seconds_to_run] [first_seed]"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import emu_load  # noqa: E402
from libadsb_b200 import synth  # noqa: E402
from oracle import oracle1090b as B  # noqa: E402


def one_case(A, seed):
    rng = np.random.default_rng(seed)
    seconds = float(rng.choice([0.05, 0.15, 0.3]))
    rate = float(rng.choice([500.0, 2500.0, 7000.0]))
    sigma = float(rng.choice([0.5, 1.5, 5.0]))
    pool = int(rng.choice([4, 60, 300]))
    flip, ap = float(rng.choice([0.0, 0.3, 0.6])), float(rng.choice([0.0, 0.25, 0.6]))
    iq = synth.config_b(seconds=seconds, seed=seed, msgs_per_s=rate, sigma=sigma, pool=pool, flip_frac=flip, ap_frac=ap)[0]
    if rng.random() < 0.2:
        a = int(rng.integers(0, iq.size // 2))
        b = min(iq.size // 2, a + 30000)
        iq[2 * a:2 * b] = rng.integers(0, 256, 2 * (b - a), dtype=np.uint8)
    sizes = [int(x) for x in rng.choice([100, 239, 240, 241, 480, 1000, 5001, 40000, 131072, 400000], int(rng.integers(2, 30)))]
    params = dict(seed=seed, seconds=seconds, rate=rate, sigma=sigma, pool=pool, flip=flip, ap=ap, sizes=sizes)
    o, d = B.Oracle1090B(), A.Demod1090()
    k = i = total = 0
    while k < iq.size:
        nb = 2 * sizes[i % len(sizes)]
        want, got = o.handle_data(iq[k:k + nb]), d.handle_data_b(iq[k:k + nb])
        if len(want) != len(got) or want.tobytes() != got.astype(want.dtype).tobytes():
            return params, total, f"call {i} at byte {k}: {len(want)} vs {len(got)} messages"
        total += len(want)
        k += nb
        i += 1
    return params, total, ""


if __name__ == "__main__":
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    A = emu_load.load_emulated()
    t0 = time.time()
    ncase = 0
    while time.time() - t0 < budget:
        params, n, bad = one_case(A, seed)
        ncase += 1
        print(f"seed {seed}: {n} messages, {'ok' if not bad else 'MISMATCH ' + bad}  {params if bad else ''}", flush=True)
        if bad:
            sys.exit(1)
        seed += 1
    print(f"{ncase} cases, all equal to the oracle")
