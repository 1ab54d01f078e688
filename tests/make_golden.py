"""Generate tests/golden/*.npz: small seeded synthetic captures plus the message sets the
REFERENCE ITSELF (oracle/_ref/dump1090_ref, built from /root/reference/dump1090 by
oracle/Makefile) decodes from them, for every parity-relevant flag combination.

Run in the build container only (needs /root/reference):  python tests/make_golden.py
The fixtures travel to the GPU box; /root/reference does not."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth          # noqa: E402
from oracle import oracle1090 as O      # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
CONFIGS = [(0, 1), (0, 0), (1, 1), (1, 0), (2, 1), (2, 0)]   # (nfix_crc, fix_df)


def boundary_capture(seed=0x1090B0B0):
    """Frames placed so that accepted messages straddle the 131072-sample mag_buf boundaries
    (exercises demod_2400.c:432-443) and the very start of the stream (zero overlap region)."""
    rng = np.random.default_rng(seed)
    n = 3 * 131072 + 5000
    frames, starts = [], []
    icaos = rng.integers(1, 1 << 24, size=8)
    for b in (1, 2, 3):
        edge = b * 131072 - 326      # stream sample whose start position is the first of mag_buf b
        for off in (-2000, -1400, -800, -150, 420, 1000, 1600):
            frames.append(synth.make_frame(17 if rng.random() < 0.6 else 11, int(icaos[rng.integers(0, 8)]), rng))
            starts.append(edge + off + rng.uniform(-40, 40))
    for s0 in (0.0, 703.3, 1500.6):      # messages at the very beginning of the stream
        frames.append(synth.make_frame(17, int(icaos[0]), rng)); starts.append(s0)
    order = np.argsort(starts)
    frames = [frames[i] for i in order]; starts = np.array(starts, dtype=float)[order]
    keep = [0]
    for i in range(1, len(starts)):      # drop overlapping frames
        if starts[i] - starts[keep[-1]] > 150:
            keep.append(i)
    frames = [frames[i] for i in keep]; starts = starts[keep]
    amps = rng.uniform(30, 80, size=len(frames)); ph = rng.uniform(0, 6.28, size=len(frames))
    return synth.synth_capture(n, frames, starts, amps, ph, 2.0, rng)


def main():
    if not O.have_ref():
        raise SystemExit("oracle/_ref/dump1090_ref missing: run `make -C oracle ref` where /root/reference exists")
    os.makedirs(GOLD, exist_ok=True)
    caps = {
        "c2_sparse": synth.config2(seconds=0.15, seed=0x10900002, msgs_per_s=4000.0)[0],
        "c3_dense": synth.config3(seconds=0.15, seed=0x10900003, msgs_per_s=8000.0, pool=300)[0],
        "boundary": boundary_capture(),
    }
    for name, iq in caps.items():
        arrays = {"iq": iq}
        for nfix, fdf in CONFIGS:
            ref = O.run_ref(iq, nfix, fdf)
            arrays[f"ref_fix{nfix}_df{fdf}"] = ref
            print(name, nfix, fdf, len(ref), "messages")
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **arrays)
    # the reference's magnitude table, dumped through the reference binary's own LUT is not reachable
    # from outside; the table fixture is the oracle's, which test_oracle pins against the golden messages.
    np.save(os.path.join(GOLD, "mag_table_crc32.npy"),
            np.array([__import__("zlib").crc32(O.mag_table().tobytes())], dtype=np.uint32))


if __name__ == "__main__":
    main()
