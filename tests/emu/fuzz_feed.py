"""TEST INFRASTRUCTURE ONLY.  Random-case search on the CPU: the Path A kernel sources on the SIMT emulator against the
oracle, over traffic densities, address-pool sizes, noise levels, repair settings, ragged feed splits, mag_buf sizes and the
resolve kernel's batch knobs.  usage: python tests/emu/fuzz_feed.py [seconds_to_run] [first_seed]
Prints one line per case; a mismatch prints the parameters that reproduce it and exits 1."""
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
    seconds = float(rng.choice([0.02, 0.05, 0.1]))
    rate = float(rng.choice([800.0, 6000.0, 30000.0, 90000.0]))
    pool = int(rng.choice([3, 40, 1500]))
    sigma = float(rng.choice([1.0, 2.0, 6.0]))
    nfix, fdf = int(rng.integers(0, 3)), int(rng.integers(0, 2))
    block = int(rng.choice([0, 0, 4096, 16384]))
    batch = rng.choice(["", "", "40", "300"])
    seq = rng.random() < 0.1
    iq = (synth.config3(seconds=seconds, seed=seed, msgs_per_s=rate, pool=pool, sigma=sigma)[0] if rng.random() < 0.8 else
          synth.config2(seconds=seconds, seed=seed, msgs_per_s=min(rate, 7000.0), pool=pool, sigma=sigma)[0])
    if rng.random() < 0.15:                                   # full-scale noise stretch in the middle
        a, b = sorted(rng.integers(0, iq.size // 2, 2))
        b = min(b, a + 20000)
        iq[2 * a:2 * b] = rng.integers(0, 256, 2 * (b - a), dtype=np.uint8)
    cuts = [0, iq.size]
    if rng.random() < 0.6:
        cuts = sorted(set([0, iq.size] + [2 * int(x) for x in rng.integers(0, iq.size // 2, int(rng.integers(1, 6)))]))
    params = dict(seed=seed, seconds=seconds, rate=rate, pool=pool, sigma=sigma, nfix=nfix, fdf=fdf, block=block, batch=str(batch), seq=seq,
                  cuts=cuts)
    for k, v in (("ADSBCU_RESOLVE_BATCH", str(batch)), ("ADSBCU_RESOLVE_SEQ", "1" if seq else "")):
        if v:
            os.environ[k] = v
        else:
            os.environ.pop(k, None)
    o = O.Oracle1090(nfix, fdf, block)
    d = A.Demod1090(nfix, fdf, block_samples=block) if block else A.Demod1090(nfix, fdf)
    mode = rng.choice(["uc8", "uc8", "uc8", "expire", "sc16", "sc16q11"])
    params["mode"] = str(mode)
    prefill = int(rng.choice([0, 0, 0, 0, 3900, 4090, 4096, 5000]))     # a filter that is (nearly) full: icao_filter.c:71-84 stops adding
    params["prefill"] = prefill
    for a in rng.integers(1, 1 << 24, prefill):
        o.filter_add(int(a))
        d.filter_add(int(a))
    if mode == "expire":
        # live mode: the filter's tables flip between some of the feeds (icao_filter.c:112-127), on both sides at the same places
        want, got = [], []
        for i in range(len(cuts) - 1):
            part = iq[cuts[i]:cuts[i + 1]]
            want.append(o.feed(part, cap=1 << 16))
            got.append(d.feed(part, cap=1 << 16))
            if rng.random() < 0.5:
                o.filter_expire()
                d.filter_expire()
        want, got = np.concatenate(want), np.concatenate(got)
    elif mode in ("sc16", "sc16q11"):
        q11 = mode == "sc16q11"
        s16 = synth.to_sc16(iq, q11)
        o.set_input_format(2 if q11 else 1)
        want = o.feed(s16, cap=1 << 16)
        got = np.concatenate([d.feed_sc16(s16[cuts[i]:cuts[i + 1]], q11, cap=1 << 16) for i in range(len(cuts) - 1)])
    else:
        want = o.feed(iq, cap=1 << 16)
        got = np.concatenate([d.feed(iq[cuts[i]:cuts[i + 1]], cap=1 << 16) for i in range(len(cuts) - 1)])
    diffs = O.same_messages(want, got)
    return params, len(want), diffs


if __name__ == "__main__":
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    A = emu_load.load_emulated()
    t0 = time.time()
    ncase = 0
    while time.time() - t0 < budget:
        params, n, diffs = one_case(A, seed)
        ncase += 1
        print(f"seed {seed}: {n} messages, {'ok' if not diffs else 'MISMATCH'}  {params if diffs else ''}", flush=True)
        if diffs:
            print(diffs[:10])
            sys.exit(1)
        seed += 1
    print(f"{ncase} cases, all equal to the oracle")
