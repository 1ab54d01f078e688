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

def crc24(frames, nbytes):
    rem = np.zeros(len(frames), np.uint32)
    for byte in range(nbytes):
        rem ^= frames[:, byte].astype(np.uint32) << 16
        for _ in range(8):
            top = (rem & 0x800000) != 0
            rem = (rem << 1) & 0xFFFFFF
            rem[top] ^= 0xFFF409
    return rem


def pin_restatement(n=1500000, seed=779):
    """oracle1090_decode_fields against the reference's own decodeModesMessage (third tap of ref_harness.c: frames from a file)
    on n random frames of the nine downlink formats the demodulator accepts, every header field; the derived fields for all
    formats but DF17 / DF18, where the ES decoder has the last word (tests/make_golden_es.py covers those)."""
    rng = np.random.default_rng(seed)
    f = rng.integers(0, 256, (n, 14), dtype=np.uint8)
    df = rng.choice([0, 4, 5, 11, 16, 17, 18, 20, 21], n).astype(np.uint8)
    f[:, 0] = (df << 3) | (f[:, 0] & 7)
    f[:, 1] &= 0x03
    f[:, 2] = 0x5A                                            # announced addresses from a pool the reference's filter can hold
    short = df < 16
    addr = (0x5A0000 | rng.integers(0, 1024, n)).astype(np.uint32)
    pi = np.isin(df, [11, 17, 18])                            # Parity/Interrogator formats: plain CRC; the others carry CRC ^ address
    p = np.where(short, crc24(f, 4), crc24(f, 11)) ^ np.where(pi, 0, addr).astype(np.uint32)
    pos = np.where(short, 4, 11)
    for k in range(3):
        f[np.arange(n), pos + k] = (p >> (8 * (2 - k))) & 255
    f[short, 7:] = 0
    ref, rc = O.run_ref_decode(f, header=True)
    assert (rc == 0).all()
    m = np.zeros(n, O.MSG_DTYPE)
    m["msg"], m["msgtype"] = f, df
    got = O.decode_fields(m)
    es = (df == 17) | (df == 18)
    for k in ("AC", "ID", "CA", "CC", "CF", "DR", "FS", "KE", "ND", "RI", "SL", "UM", "VS", "es_pending"):
        assert np.array_equal(got[k], ref[k]), k
    for k in ("airground", "alert_valid", "alert", "spi_valid", "spi", "altitude_baro_valid", "altitude_baro_unit", "squawk_valid", "squawk",
              "altitude_baro"):
        assert np.array_equal(got[k][~es], ref[k][~es]), k
    print("restatement == reference on", n, "random frames")


if __name__ == "__main__":
    pin_restatement()
    iq = synth.config3(seconds=1.5, seed=779, msgs_per_s=3000.0, pool=300)[0]
    msgs, fields = O.run_ref_fields(iq, 1, 1)
    assert len(msgs) == len(fields) and len(msgs) > 1500
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fields_a.npz"), msgs=msgs, fields=fields)
    print(len(msgs), "messages; DFs:", {int(d): int((msgs["msgtype"] == d).sum()) for d in np.unique(msgs["msgtype"])})
