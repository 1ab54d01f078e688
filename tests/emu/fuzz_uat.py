"""TEST INFRASTRUCTURE ONLY.  Random-case search on the CPU for Path C: the UAT kernel sources on the SIMT emulator against
the restatement (oracle/oracle978.c) call by call, over traffic, noise, corrupted bytes, and ragged HandleData sizes (the
reference's staging loop makes the frame set depend on the call sizes).
usage: python tests/emu/fuzz_uat.py [seconds_to_run] [first_seed]"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import emu_load  # noqa: E402
from libadsb_b200 import synth978 as S  # noqa: E402
from oracle import oracle978 as O  # noqa: E402


def one_case(A, seed):
    rng = np.random.default_rng(seed)
    seconds = float(rng.choice([0.1, 0.25, 0.5]))
    sigma = float(rng.choice([1.0, 2.0, 5.0]))
    adsb = float(rng.choice([100.0, 800.0, 3000.0]))
    up = float(rng.choice([0.0, 4.0, 20.0]))
    iq = S.config5(seconds=seconds, seed=seed, sigma=sigma, adsb_per_s=adsb, uplink_per_s=up)[0]
    sizes = [int(x) for x in rng.choice([10, 70, 72, 74, 999, 5000, 20000, 65536, 70001, 131072, 300000], int(rng.integers(3, 40)))]
    params = dict(seed=seed, seconds=seconds, sigma=sigma, adsb=adsb, up=up, sizes=sizes)
    d, o = A.Demod1090(), O.Oracle978()
    k = i = total = 0
    while k < iq.size:
        nb = 2 * sizes[i % len(sizes)]
        want, got = o.handle_data(iq[k:k + nb]), d.handle_data_uat(iq[k:k + nb])
        if len(want) != len(got) or want.tobytes() != got.tobytes():
            return params, total, f"call {i} at byte {k}: {len(want)} vs {len(got)} frames"
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
        print(f"seed {seed}: {n} frames, {'ok' if not bad else 'MISMATCH ' + bad}  {params if bad else ''}", flush=True)
        if bad:
            sys.exit(1)
        seed += 1
    print(f"{ncase} cases, all equal to the oracle")
