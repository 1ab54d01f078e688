"""Writes tests/golden/sc16.npz: one capture in the reference's two int16 input formats and what the REFERENCE
binary (oracle/_ref/dump1090_ref --iformat SC16 / SC16Q11 --fix-2bit) hands to useModesMessage."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth          # noqa: E402
from oracle import oracle1090 as O      # noqa: E402

if __name__ == "__main__":
    iq8 = synth.config3(seconds=0.15, seed=0x10905C16, msgs_per_s=6000.0, pool=300)[0]
    out = {}
    for q11, name in ((False, "SC16"), (True, "SC16Q11")):
        iq = synth.to_sc16(iq8, q11)
        ref = O.run_ref(iq, 2, 1, extra=("--iformat", name))
        assert len(ref) > 100
        out["iq_" + name.lower()] = iq
        out["ref_" + name.lower()] = ref
        print(name, len(ref), "messages")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sc16.npz"), **out)
