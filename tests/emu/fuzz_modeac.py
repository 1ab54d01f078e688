"""TEST INFRASTRUCTURE ONLY.  Random-case search on the CPU for Mode A/C (adsbcu_1090_modeac_block beside the Mode S feed, the
dump1090.c:895-905 loop): kernel sources on the SIMT emulator against the oracle with --modeac, over reply density, noise,
repair settings and block sizes.  usage: python tests/emu/fuzz_modeac.py [seconds_to_run] [first_seed]"""
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
from oracle import oracle1090 as O  # noqa: E402


def one_case(A, seed):
    rng = np.random.default_rng(seed)
    seconds = float(rng.choice([0.03, 0.08, 0.15]))
    sigma = float(rng.choice([1.0, 2.0, 6.0]))
    ac = float(rng.choice([500.0, 3000.0, 20000.0]))
    ms = float(rng.choice([0.0, 1500.0, 6000.0]))
    nfix = int(rng.integers(0, 3))
    block = int(rng.choice([131072, 131072, 40000, 16384, 4096]))
    params = dict(seed=seed, seconds=seconds, sigma=sigma, ac=ac, ms=ms, nfix=nfix, block=block)
    iq = synth.config_ac(seconds=seconds, seed=seed, msgs_per_s=ms, ac_per_s=ac, sigma=sigma)[0]
    o = O.Oracle1090(nfix, 1, block)
    O.lib().oracle1090_set_mode_ac(o.h, 1)
    want = o.feed(iq, cap=1 << 18)
    d = A.Demod1090(nfix, 1, block_samples=block)
    out, n = [], iq.size // 2
    for s in range(0, n, block):
        e = min(n, s + block)
        out.append(d.feed(iq[2 * s:2 * e]))
        out.append(d.modeac_block(iq[2 * s:2 * e], iq[2 * max(0, s - 326):2 * s], 5 * s))
    got = np.concatenate(out)
    return params, len(want), O.same_messages(want, got)


if __name__ == "__main__":
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    A = emu_load.load_emulated()
    t0, ncase = time.time(), 0
    while time.time() - t0 < budget:
        params, n, diffs = one_case(A, seed)
        ncase += 1
        print(f"seed {seed}: {n} messages, {'ok' if not diffs else 'MISMATCH'}  {params if diffs else ''}", flush=True)
        if diffs:
            print(diffs[:10])
            sys.exit(1)
        seed += 1
    print(f"{ncase} cases, all equal to the oracle")
