"""tests/golden/uat_ref.npz: what the reference's own UAT978.cpp (oracle/_ref/libref978.so, compiled where it lies)
produces for the two Path C stages whose source is in the tree: the 65536-entry phase table of InitATan2Table and the
sequence of process_buffer(len, offset) calls HandleData makes for a capture fed in 262144-byte buffers (with
oracle978's process_buffer plugged in for dump978's).  Run in the build container."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libadsb_b200 import synth978  # noqa: E402
from oracle import oracle978 as O, ref978 as R  # noqa: E402

SEED, SECONDS, BUF = 31, 0.6, 262144


def run_reference(iq, feed):
    o = O.Oracle978()
    frames = []

    def process(phi, offset):
        done, f = o.process_buffer(phi, offset)
        frames.append(f)
        return done
    r = R.Ref978(process)
    for k in range(0, iq.size, feed):
        r.handle_data(iq[k:k + feed])
    return r, np.array(r.calls, np.int64), np.concatenate(frames) if frames else np.zeros(0, O.FRAME_DTYPE)


if __name__ == "__main__":
    iq = synth978.config5(seconds=SECONDS, seed=SEED)[0]
    r, calls, frames = run_reference(iq, BUF)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "uat_ref.npz"), phase_table=r.phase_table(), calls=calls,
                        frame_offsets=frames["offset"], frame_rs=frames["rs_errors"], frame_len=frames["len"],
                        seed=SEED, seconds=SECONDS, buf=BUF)
    print(len(calls), "calls", len(frames), "frames")
