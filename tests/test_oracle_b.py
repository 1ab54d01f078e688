"""CPU suite for Path B (libadsb's own 2.0 MS/s handler, ADSB1090.cpp): the oracle restatement against
the reference-generated golden fixture, and -- where the reference could be compiled (this container) --
against the reference's own functions directly."""
import os

import numpy as np
import pytest

from conftest import GOLD

BUF = 262144


@pytest.fixture(scope="module")
def B():
    from oracle import oracle1090b as B
    return B


@pytest.fixture(scope="module")
def gold_b():
    with np.load(os.path.join(GOLD, "b_mixed.npz")) as z:
        return {k: z[k] for k in z.files}


def test_known_answers(B):
    o = B.Oracle1090B()
    # ADSB1090.cpp:266-275: first and last non-zero entries of ModesChecksumTable
    assert o.checksum(bytes([0x80] + [0] * 13), 112) == 0x3935EA
    assert o.checksum(bytes([0] * 10 + [0x01] + [0] * 3), 112) == 0xFFF409
    # dump1090/net_io.c:1375 example frame: a valid DF17
    f = bytes.fromhex("8D4B969699155600E87406F5B69F")
    assert o.checksum(f, 112) == int.from_bytes(f[11:], "big")
    d = o.decode(f)
    assert (d["msgtype"], d["msgbits"], d["crcok"], d["errorbit"], d["addr"]) == (17, 112, 1, -1, 0x4B9696)
    lut = o.lut().reshape(129, 129)
    assert lut[0, 0] == 0 and lut[128, 128] == round((2 * 128 * 128) ** 0.5 * 360) and lut[3, 4] == 1800


def test_oracle_vs_reference_golden(B, gold_b):
    """Buffer by buffer (RTLSDR.hpp:55 buffers): as many accepted messages as the reference made
    IListener::OnChanged callbacks, for the same aircraft, and the same counters / ICAO cache at the end."""
    o = B.Oracle1090B()
    iq, k0 = gold_b["iq"], 0
    for b, nref in enumerate(gold_b["per_buffer"]):
        got = o.handle_data(iq[b * BUF:(b + 1) * BUF])
        ref = gold_b["ref_callbacks"][k0:k0 + nref]
        assert len(got) == nref
        assert np.array_equal(got["addr"], ref["addr"])
        assert list(o.stats().values()) == [int(x) for x in gold_b["ref_stats"][b]]
        want = gold_b["oracle_msgs"][k0:k0 + nref]
        assert got.tobytes() == want.tobytes()
        k0 += nref
    assert sorted(o.cache()) == [int(x) for x in gold_b["ref_cache"]]
    m = gold_b["oracle_msgs"]
    assert len(m) > 1000 and (m["errorbit"] >= 0).sum() > 50 and m["corrected"].sum() > 5
    assert np.isin(m["msgtype"], (0, 4, 5, 16, 20, 21)).sum() > 100      # Address/Parity replies recovered via the cache


def test_oracle_vs_reference_functions(B):
    if not B.have_ref():
        pytest.skip("oracle/_ref/libref1090b.so not built (needs /root/reference)")
    from libadsb_b200 import synth
    o, r = B.Oracle1090B(), B.Ref1090B()
    assert np.array_equal(o.lut(), r.lut())
    rng = np.random.default_rng(11)
    for _ in range(300):
        df = int(rng.choice([0, 4, 5, 11, 16, 17, 18, 19, 20, 21, 24, 3]))
        f = synth.make_frame(df, int(rng.integers(1, 1 << 24)), rng).ljust(14, b"\0")
        u = rng.random()
        if u < 0.4:
            f = synth.flip_bits(f, [int(rng.integers(0, 112 if df & 16 else 56))])
        elif u < 0.5:
            f = synth.flip_bits(f, [int(x) for x in rng.choice(56, size=2, replace=False)])
        for bits in (56, 112):
            assert o.checksum(f, bits) == r.checksum(f, bits)
        assert o.fix_single(f, 112 if df & 16 else 56) == r.fix_single(f, 112 if df & 16 else 56)
        assert o.decode(f) == r.decode(f)        # same cache state on both sides: clean DF11/17 are added as we go
    assert o.cache() == r.cache()


def test_oracle_vs_reference_windows(B):
    """Fresh capture, ragged windows, including ones shorter than a frame."""
    if not B.have_ref():
        pytest.skip("oracle/_ref/libref1090b.so not built (needs /root/reference)")
    from libadsb_b200 import synth
    iq = synth.config_b(seconds=0.6, seed=77, msgs_per_s=2500.0, flip_frac=0.3)[0]
    o, r = B.Oracle1090B(), B.Ref1090B()
    rng = np.random.default_rng(2)
    k = 0
    while k < iq.size:
        nb = 2 * int(rng.choice([480, 1000, 5000, 40000, 131072]))
        a, c = o.handle_data(iq[k:k + nb]), r.handle_data(iq[k:k + nb])
        assert len(a) == len(c) and np.array_equal(a["addr"], c["addr"])
        assert o.stats() == r.stats()
        k += nb
    assert o.cache() == r.cache() and o.stats()["goodcrc"] > 300
