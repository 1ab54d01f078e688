"""Writes tests/golden/ac_mixed.npz: Mode S + Mode A/C capture and what the REFERENCE binary
(oracle/_ref/dump1090_ref --modeac) hands to useModesMessage for it.  Run in the build container."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth          # noqa: E402
from oracle import oracle1090 as O      # noqa: E402

if __name__ == "__main__":
    iq, _, _ = synth.config_ac(seconds=0.3, seed=0x1090AC01)
    ref = O.run_ref(iq, 1, 1, extra=("--modeac",))
    assert (ref["msgtype"] == 32).sum() > 50
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ac_mixed.npz"), iq=iq, ref_fix1_modeac=ref)
    print("ac_mixed:", len(ref), "messages,", int((ref["msgtype"] == 32).sum()), "Mode A/C")
