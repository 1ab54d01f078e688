"""tests/golden/fields_es.npz: DF17 / DF18 frames and what the reference's own decodeModesMessage (mode_s.c:520-796 with
decodeExtendedSquitter, :798-1551) left in struct modesMessage for them (oracle/_ref/dump1090_ref, third tap of
oracle/ref_harness.c).  Before writing the fixture the script pins the restatement (oracle1090_decode_es) on a much
larger random set - every frame, every field, bit for bit (the float fields as their bit patterns).
Run in the build container."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle1090 as O  # noqa: E402


def crc24(frames):
    """Mode S parity of the first 11 bytes of every frame (generator 0xFFF409), vectorised."""
    rem = np.zeros(len(frames), np.uint32)
    for byte in range(11):
        rem ^= frames[:, byte].astype(np.uint32) << 16
        for _ in range(8):
            top = (rem & 0x800000) != 0
            rem = (rem << 1) & 0xFFFFFF
            rem[top] ^= 0xFFF409
    return rem


def es_frames(rng, n, metype=None):
    """n valid DF17 / DF18 frames: random CA / CF, address and ME (metype fixed if given), correct parity."""
    f = rng.integers(0, 256, (n, 14), dtype=np.uint8)
    df = np.where(rng.random(n) < 0.5, 17, 18).astype(np.uint8)
    f[:, 0] = (df << 3) | rng.integers(0, 8, n, dtype=np.uint8)
    f[:, 1] &= 0x03                     # a pool of 1024 addresses per run: the reference's filter holds 4096
    f[:, 2] = 0x5A
    if metype is not None:
        f[:, 4] = (np.uint8(metype) << 3) | (f[:, 4] & 7)
    # some all-zero position / altitude corners (mode_s.c:1056) and zero speed fields
    z = rng.random(n) < 0.03
    f[z, 5:11] = 0
    p = crc24(f)
    f[:, 11], f[:, 12], f[:, 13] = p >> 16, (p >> 8) & 255, p & 255
    return f


def with_repairs(rng, frames, share=0.1):
    """`as received`: a copy where some DF17 frames are followed by a same-address frame with one ME bit flipped (the reference
    repairs it against its filter: correctedbits = 1, reliable = 0).  Returns (received, corrected frames, corrected bits)."""
    recv, fixed, bits = [], [], []
    for k in range(len(frames)):
        recv.append(frames[k]); fixed.append(frames[k]); bits.append(0)
        if frames[k, 0] >> 3 == 17 and rng.random() < share:
            g = es_frames(rng, 1)[0]
            g[0] = frames[k, 0]
            g[1:4] = frames[k, 1:4]
            p = int(crc24(g[None, :])[0])
            g[11:14] = (p >> 16, (p >> 8) & 255, p & 255)
            bad = g.copy()
            bit = int(rng.integers(32, 88))
            bad[bit >> 3] ^= 0x80 >> (bit & 7)
            recv.append(bad); fixed.append(g); bits.append(1)
    return np.array(recv, np.uint8), np.array(fixed, np.uint8), np.array(bits, np.uint8)


def differences(a, b):
    out = []
    for name in a.dtype.names:
        x, y = a[name], b[name]
        if x.dtype.kind == "f":
            x, y = x.view(np.uint32), y.view(np.uint32)
        bad = np.nonzero(x != y)[0]
        if len(bad):
            out.append((name, len(bad), int(bad[0])))
    return out


if __name__ == "__main__":
    rng = np.random.default_rng(1551)
    # 1. pin the restatement: random frames, then every metype by itself
    big = np.concatenate([es_frames(rng, 400000)] + [es_frames(rng, 40000, m) for m in range(32)])
    ref, rc = O.run_ref_decode(big)
    assert (rc == 0).all(), np.unique(rc, return_counts=True)
    d = differences(O.decode_es(big), ref)
    assert not d, d
    # every (E/W, N/S) velocity pair of metype 19 subtypes 1 and 2 (sqrtf / atan2, mode_s.c:894-902), both signs
    ew, ns = np.meshgrid(np.arange(1024, dtype=np.uint64), np.arange(1024, dtype=np.uint64), indexing="ij")
    ew, ns = ew.ravel(), ns.ravel()
    for sub in (1, 2):
        for signs in range(4):
            me = (np.uint64(19) << np.uint64(51)) | (np.uint64(sub) << np.uint64(48)) | (np.uint64(signs >> 1) << np.uint64(42)) | \
                (ew << np.uint64(32)) | (np.uint64(signs & 1) << np.uint64(31)) | (ns << np.uint64(21))
            f = np.zeros((len(me), 14), np.uint8)
            f[:, 0] = 17 << 3 | 5
            f[:, 1:4] = (0x4B, 0x17, 0x2A)
            for b in range(7):
                f[:, 4 + b] = (me >> np.uint64(8 * (6 - b))) & np.uint64(255)
            p = crc24(f)
            f[:, 11], f[:, 12], f[:, 13] = p >> 16, (p >> 8) & 255, p & 255
            ref_v, rc_v = O.run_ref_decode(f)
            assert (rc_v == 0).all()
            d = differences(O.decode_es(f), ref_v)
            assert not d, (sub, signs, d)
    print("restatement == reference on", len(big) + 8 * len(ew), "frames")
    # 2. the committed fixture: 200 frames per metype + random ones, some of them repaired by the reference
    small = np.concatenate([es_frames(rng, 200, m) for m in range(32)] + [es_frames(rng, 1600)])
    recv, fixed, bits = with_repairs(rng, small)
    ref, rc = O.run_ref_decode(recv)
    assert (rc == 0).all() and int(bits.sum()) > 100
    assert (ref["reliable"][bits == 1] == 0).all()
    d = differences(O.decode_es(fixed, bits), ref)
    assert not d, d
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fields_es.npz"), received=recv, frames=fixed, corrected=bits, fields=ref)
    print(len(recv), "frames in the fixture,", int(bits.sum()), "repaired; metypes:", np.bincount(ref["metype"], minlength=32).tolist())
