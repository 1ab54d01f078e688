"""Writes tests/golden/b_mixed.npz: a synthetic 2.0 MS/s capture and what the REFERENCE's own
ADSB1090Handler (oracle/_ref/libref1090b.so = /root/reference/ADSB1090.cpp compiled in place) does with
it, buffer by buffer: IListener::OnChanged callbacks, statistics counters, cached ICAO addresses.  Also the
oracle's message records for the same buffers (checked against the reference by tests/test_oracle_b.py).
Run in the build container (needs /root/reference): python tests/make_golden_b.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth          # noqa: E402
from oracle import oracle1090b as B     # noqa: E402

BUF = 262144                             # RTLSDR.hpp:55 buffer size in bytes

if __name__ == "__main__":
    iq, _, _ = synth.config_b(seconds=1.2, seed=0x1090B001, msgs_per_s=1500.0)
    ref, ora = B.Ref1090B(), B.Oracle1090B()
    cbs, recs, nper, stats = [], [], [], []
    for k in range(0, iq.size, BUF):
        c = ref.handle_data(iq[k:k + BUF])
        r = ora.handle_data(iq[k:k + BUF])
        assert len(c) == len(r) and np.array_equal(c["addr"], r["addr"])
        cbs.append(c); recs.append(r); nper.append(len(c)); stats.append(list(ref.stats().values()))
    assert ref.stats() == ora.stats() and ref.cache() == ora.cache()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "b_mixed.npz"), iq=iq, ref_callbacks=np.concatenate(cbs),
                        oracle_msgs=np.concatenate(recs), per_buffer=np.array(nper), ref_stats=np.array(stats, dtype=np.uint64),
                        ref_cache=np.array(sorted(ref.cache()), dtype=np.uint32))
    print("b_mixed:", sum(nper), "messages in", len(nper), "buffers")
