"""tests/golden/fields_mb.npz: DF20 / DF21 frames and what the reference's own decodeModesMessage -> decodeCommB (comm_b.c:50-985)
left in struct modesMessage for them (oracle/_ref/dump1090_ref, third tap of oracle/ref_harness.c).  The frames are built per
register (empty, BDS 1,0 2,0 3,0 1,7 4,0 5,0 6,0 4,4 0,5) with mostly plausible field values and some out-of-range ones, plus
plain random MB fields.  Before writing the fixture the script pins the restatement (oracle1090_decode_commb) on a much
larger set - every frame, every field, floats as bit patterns.  Run in the build container."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle1090 as O  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
from make_golden_es import crc24, differences  # noqa: E402

U = np.uint64


def pack(n, fields):
    """fields: list of (values[n] or scalar, width); MSB first, 56 bits in total."""
    assert sum(w for _, w in fields) == 56
    v = np.zeros(n, U)
    for val, w in fields:
        v = (v << U(w)) | (np.broadcast_to(np.asarray(val, U), (n,)) & U((1 << w) - 1))
    return v


def pick(rng, n, lo, hi, width, p_wild=0.12):
    """Mostly values in [lo, hi], sometimes anything the field can hold, sometimes zero."""
    v = rng.integers(lo, hi + 1, n).astype(U)
    wild = rng.random(n) < p_wild
    v[wild] = rng.integers(0, 1 << width, int(wild.sum())).astype(U)
    v[rng.random(n) < 0.03] = 0
    return v


def flag(rng, n, p=0.85):
    return (rng.random(n) < p).astype(U)


def registers(rng, n):
    """A dict name -> 56-bit MB values."""
    r = {}
    r["random"] = rng.integers(0, 1 << 56, n, dtype=U)
    first = rng.choice([0x00, 0x40, 0x50, 0x60, 0x10, 0x20, 0x30, 0x70], n).astype(U)
    rest = np.where(rng.random(n) < 0.8, 0, rng.integers(0, 1 << 48, n, dtype=U))
    r["empty"] = (first << U(48)) | rest
    r["bds10"] = pack(n, [(0x10, 8), (rng.integers(0, 2, n), 1), (np.where(rng.random(n) < 0.8, 0, rng.integers(0, 32, n)), 5), (rng.integers(0, 1 << 42, n, dtype=U), 42)])
    caps = rng.integers(0, 1 << 24, n, dtype=U)
    es = rng.choice([0x3F, 0x3E, 0x00, 0x0E, 0x0F, 0x15], n).astype(U)
    caps = (caps & U(0x03FFFF)) | (es << U(18))
    caps &= ~np.where(rng.random(n) < 0.7, U(0b000000000111110000011100), U(0))       # the "unlikely" registers mostly off
    r["bds17"] = pack(n, [(caps, 24), (np.where(rng.random(n) < 0.9, 0, rng.integers(0, 1 << 32, n, dtype=U)), 32)])
    good = np.array([1 + k for k in range(26)] + [32] + [48 + k for k in range(10)], U)
    chars = good[rng.integers(0, len(good), (n, 8))]
    odd = rng.random((n, 8))
    chars[odd < 0.04] = 0                                                                 # '@' padding
    chars[odd > 0.99] = 43                                                                # an invalid character
    r["bds20"] = pack(n, [(0x20, 8)] + [(chars[:, k], 6) for k in range(8)])
    r["bds30"] = pack(n, [(0x30, 8), (rng.integers(0, 1 << 48, n, dtype=U), 48)])
    same = rng.random(n) < 0.5
    mcp = pick(rng, n, 63, 3125, 12)
    fms = np.where(same, mcp, pick(rng, n, 63, 3125, 12))
    level = ((rng.integers(2, 100, n) * 500 + 8) // 16).astype(U)                          # a multiple of 500 ft, as selected altitudes are
    mcp = np.where(rng.random(n) < 0.5, level, mcp)
    fms = np.where(same, mcp, fms)
    r["bds40"] = pack(n, [(flag(rng, n), 1), (mcp, 12), (flag(rng, n), 1), (fms, 12), (flag(rng, n), 1), (pick(rng, n, 1000, 3000, 12), 12),
                          (np.where(rng.random(n) < 0.93, 0, rng.integers(0, 256, n)), 8), (flag(rng, n, 0.6), 1), (rng.integers(0, 8, n), 3),
                          (np.where(rng.random(n) < 0.95, 0, rng.integers(0, 4, n)), 2), (flag(rng, n, 0.6), 1), (rng.integers(0, 4, n), 2)])
    sign = rng.integers(0, 2, n).astype(U)
    roll = np.where(sign == 1, pick(rng, n, 285, 511, 9), pick(rng, n, 0, 227, 9))
    r["bds50"] = pack(n, [(flag(rng, n, 0.95), 1), (sign, 1), (roll, 9), (flag(rng, n, 0.95), 1), (rng.integers(0, 2, n), 1), (rng.integers(0, 1024, n), 10),
                          (flag(rng, n, 0.95), 1), (pick(rng, n, 25, 350, 10), 10), (flag(rng, n), 1), (rng.integers(0, 2, n), 1),
                          (pick(rng, n, 0, 511, 9), 9), (flag(rng, n, 0.95), 1), (pick(rng, n, 25, 350, 10), 10)])
    bsign = rng.integers(0, 2, n).astype(U)
    brate = np.where(bsign == 1, pick(rng, n, 325, 511, 9), pick(rng, n, 0, 187, 9))
    isign = np.where(rng.random(n) < 0.8, bsign, rng.integers(0, 2, n).astype(U))
    irate = np.where(rng.random(n) < 0.6, brate, np.where(isign == 1, pick(rng, n, 325, 511, 9), pick(rng, n, 0, 187, 9)))
    r["bds60"] = pack(n, [(flag(rng, n, 0.95), 1), (rng.integers(0, 2, n), 1), (rng.integers(0, 1024, n), 10), (flag(rng, n, 0.95), 1),
                          (pick(rng, n, 50, 700, 10), 10), (flag(rng, n, 0.95), 1), (pick(rng, n, 25, 225, 10), 10), (flag(rng, n), 1), (bsign, 1),
                          (brate, 9), (flag(rng, n), 1), (isign, 1), (irate, 9)])
    ssign = rng.integers(0, 2, n).astype(U)
    sat = np.where(ssign == 1, pick(rng, n, 192, 511, 9), pick(rng, n, 0, 240, 9))
    asp_v, tur_v, hum_v = flag(rng, n, 0.5), flag(rng, n, 0.5), flag(rng, n, 0.5)
    keep = rng.random(n) < 0.1                                                            # field left non-zero although not valid
    r["bds44"] = pack(n, [(rng.choice([0, 1, 2, 3, 4, 5, 9], n, p=[.05, .25, .25, .2, .15, .05, .05]), 4), (flag(rng, n, 0.95), 1), (pick(rng, n, 0, 250, 9), 9),
                          (rng.integers(0, 512, n), 9), (flag(rng, n, 0.95), 1), (ssign, 1), (sat, 9), (asp_v, 1),
                          (np.where((asp_v == 1) | keep, pick(rng, n, 25, 1100, 11), 0), 11), (tur_v, 1), (np.where((tur_v == 1) | keep, rng.integers(0, 4, n), 0), 2),
                          (hum_v, 1), (np.where((hum_v == 1) | keep, rng.integers(0, 64, n), 0), 6)])
    r["bds05"] = pack(n, [(pick(rng, n, 9, 18, 5, 0.1), 5), (rng.integers(0, 8, n), 3), (rng.integers(0, 4096, n), 12), (flag(rng, n, 0.1), 1),
                          (rng.integers(0, 2, n), 1), (pick(rng, n, 1, (1 << 17) - 1, 17, 0.02), 17), (pick(rng, n, 1, (1 << 17) - 1, 17, 0.02), 17)])
    return r


def commb_frames(rng, mb, bds05=False):
    """DF20 / DF21 replies around the MB values: FS random, DR mostly 0, UM random, AC/ID random (for bds05: the AC that matches the
    MB's altitude, mostly), Address/Parity for a pool of 1024 addresses."""
    n = len(mb)
    f = np.zeros((n, 14), np.uint8)
    df = np.where(rng.random(n) < (0.9 if bds05 else 0.55), 20, 21).astype(np.uint32)
    dr = np.where(rng.random(n) < 0.9, 0, rng.integers(0, 32, n)).astype(np.uint32)
    um = np.where(rng.random(n) < 0.5, 0, rng.integers(0, 64, n)).astype(np.uint32)
    ac = rng.integers(0, 1 << 13, n).astype(np.uint32)
    if bds05:
        ac12 = ((mb >> U(36)) & U(0xFFF)).astype(np.uint32)
        match = ((ac12 & 0x0FC0) << 1) | (ac12 & 0x003F)
        ac = np.where(rng.random(n) < 0.85, match, ac)
    head = (df << 27) | (rng.integers(0, 8, n).astype(np.uint32) << 24) | (dr << 19) | (um << 13) | ac
    for b in range(4):
        f[:, b] = (head >> (8 * (3 - b))) & 255
    for b in range(7):
        f[:, 4 + b] = (mb >> U(8 * (6 - b))) & U(255)
    addr = (0x5A0000 | rng.integers(0, 1024, n)).astype(np.uint32)
    p = crc24(f) ^ addr
    f[:, 11], f[:, 12], f[:, 13] = p >> 16, (p >> 8) & 255, p & 255
    return f


def build(rng, n_each):
    regs = registers(rng, n_each)
    return np.concatenate([commb_frames(rng, regs[k], k == "bds05") for k in sorted(regs)])


if __name__ == "__main__":
    rng = np.random.default_rng(985)
    big = build(rng, 250000)
    ref, rc = O.run_ref_decode(big, commb=True)
    assert (rc == 0).all(), np.unique(rc, return_counts=True)
    print("formats seen:", np.bincount(ref["commb_format"], minlength=13).tolist())
    d = differences(O.decode_commb(big), ref)
    assert not d, d
    print("restatement == reference on", len(big), "frames")
    ambiguous = big[ref["commb_format"] == 1][:80]                                       # ties are rare: carry some over from the big set
    small = np.concatenate([build(rng, 700), ambiguous])
    ref, rc = O.run_ref_decode(small, commb=True)
    assert (rc == 0).all()
    assert (np.bincount(ref["commb_format"], minlength=13) > 20).all(), np.bincount(ref["commb_format"], minlength=13)
    for bit in range(20):
        assert (ref["valid"] & (1 << bit)).any(), bit
    assert not differences(O.decode_commb(small), ref)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fields_mb.npz"), frames=small, fields=ref)
    print(len(small), "frames in the fixture; formats:", np.bincount(ref["commb_format"], minlength=13).tolist())
