"""tests/golden/fields_a.npz: accepted messages of a dense mixed-DF capture and the surveillance fields (mode_s.c:626-779) the
reference's own decodeModesMessage filled in for them (oracle/_ref/dump1090_ref, second tap of oracle/ref_harness.c).
Run in the build container."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth  # noqa: E402
from oracle import oracle1090 as O  # noqa: E402

if __name__ == "__main__":
    iq = synth.config3(seconds=1.5, seed=779, msgs_per_s=3000.0, pool=300)[0]
    msgs, fields = O.run_ref_fields(iq, 1, 1)
    assert len(msgs) == len(fields) and len(msgs) > 1500
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fields_a.npz"), msgs=msgs, fields=fields)
    print(len(msgs), "messages; DFs:", {int(d): int((msgs["msgtype"] == d).sum()) for d in np.unique(msgs["msgtype"])})
