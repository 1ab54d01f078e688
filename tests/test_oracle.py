"""CPU suite: pins the oracle (oracle/oracle1090.c) against the reference's own outputs
(tests/golden, produced by oracle/_ref/dump1090_ref) and against the known-answer material
the reference tree holds for this path."""
import zlib

import numpy as np
import pytest

from conftest import CONFIGS, GOLDEN_NAMES


@pytest.mark.parametrize("name", GOLDEN_NAMES)
@pytest.mark.parametrize("nfix,fdf", CONFIGS)
def test_oracle_matches_reference_golden(oracle_mod, golden, name, nfix, fdf):
    O = oracle_mod
    g = golden[name]
    got = O.Oracle1090(nfix, fdf).feed(g["iq"])
    ref = g[f"ref_fix{nfix}_df{fdf}"]
    assert len(ref) > 0
    assert O.same_messages(ref, got) == []


def test_oracle_matches_reference_binary_live(oracle_mod):
    """When the reference binary is present (build container, or shipped to the GPU box) run
    it on a fresh seed and compare message for message."""
    O = oracle_mod
    if not O.have_ref():
        pytest.skip("oracle/_ref/dump1090_ref not built")
    from libadsb_b200 import synth
    iq, _, _ = synth.config3(seconds=0.4, seed=1234, msgs_per_s=5000.0, pool=500)
    for nfix, fdf in CONFIGS:
        ref = O.run_ref(iq, nfix, fdf)
        got = O.Oracle1090(nfix, fdf).feed(iq)
        assert len(ref) > 100
        assert O.same_messages(ref, got) == [], (nfix, fdf)


def test_crc_known_answers(oracle_mod):
    O = oracle_mod
    # example frame in a comment of the reference, dump1090/net_io.c:1375: a valid DF17
    frame = bytes.fromhex("8D4B969699155600E87406F5B69F")
    assert O.crc(frame, 112) == 0
    # ModesChecksumTable[0] of the libadsb handler (ADSB1090.cpp:266-275) == CRC of a lone bit 0
    one = bytes([0x80] + [0] * 13)
    assert O.crc(one, 112) == 0x3935EA
    # generator polynomial (crc.c:28): a lone bit 24 places before the parity field
    assert O.crc(bytes([0] * 10 + [1, 0, 0, 0]), 112) == 0xFFF409
    # parity field is xored in verbatim
    assert O.crc(bytes([0] * 11 + [0x12, 0x34, 0x56]), 112) == 0x123456
    assert O.crc(bytes([0] * 4 + [0x12, 0x34, 0x56]), 56) == 0x123456


def test_syndrome_table_sizes(oracle_mod):
    """crc.c built with -DCRCDEBUG reports 56/112 entries for --fix and 1596/3694 for --fix-2bit
    (SURVEY.md section 8a row A8)."""
    O = oracle_mod
    o1, o2 = O.Oracle1090(1, 1), O.Oracle1090(2, 1)
    assert len(o1.syndrome_table(56)) == 56 and len(o1.syndrome_table(112)) == 112
    t56, t112 = o2.syndrome_table(56), o2.syndrome_table(112)
    assert len(t56) == 1596 and len(t112) == 3694
    assert np.all(np.diff(t112[:, 0].astype(np.int64)) > 0)          # sorted, unique syndromes
    # every entry reproduces its syndrome
    for row in t112[::97]:
        msg = bytearray(14)
        for b in row[2:2 + row[1]]:
            msg[int(b) >> 3] ^= 0x80 >> (int(b) & 7)
        assert O.crc(bytes(msg), 112) == row[0]


def test_df_prefilter_equals_correctable_sets():
    """valid_df_*_bitset (demod_2400.c:80-93) vs df_correctable_* (mode_s.c:260-265): any DF that gets a
    112-bit CRC must have been sliced to 14 bytes."""
    def dmg(df, d):
        s = 1 << df
        if d:
            for b in range(5):
                s |= dmg(df ^ (1 << b), d - 1)
        return s
    fix_short = [0x00000800, 0x08008e08, 0x08008e08]
    fix_long = [0x00060000, 0x066f0006, 0x6fff066f]
    base_long = sum(1 << d for d in (16, 17, 18, 20, 21))
    for nfix in (1, 2):
        valid_long = base_long | dmg(17, nfix) | dmg(18, nfix)
        assert fix_long[nfix] & ~valid_long == 0
        assert dmg(17, nfix) | dmg(18, nfix) == fix_long[nfix]
    assert dmg(11, 1) == fix_short[1]
    assert fix_long[0] & ~base_long == 0


def test_mag_table_numpy_restatement(oracle_mod):
    """tables.c:24-35 restated in numpy float32 must give the oracle's table, and the register
    formula used on the GPU (2^34 scaling, floor(2v)+1>>1) must give it too."""
    O = oracle_mod
    t = O.mag_table()
    assert zlib.crc32(t.tobytes()) == int(np.load(__import__("os").path.join(
        __import__("conftest").GOLD, "mag_table_crc32.npy"))[0])
    v = np.arange(256, dtype=np.float64)
    f = ((v - 127.4) / 128).astype(np.float32)
    sq = f * f
    i, q = np.meshgrid(np.arange(256), np.arange(256), indexing="xy")      # index = I | Q << 8
    magsq = (sq[i] + sq[q]).astype(np.float32)
    mag = np.sqrt(magsq).astype(np.float32) * np.float32(65536.0)
    r = np.floor(mag.astype(np.float64) + 0.5)                              # round(): half away from zero
    r = np.minimum(r, 65535).astype(np.uint16).reshape(-1)
    assert np.array_equal(r, t)
    sq34 = sq * np.float32(2.0 ** 34)
    s = np.sqrt((sq34[i] + sq34[q]).astype(np.float32)).astype(np.float32)
    w = np.floor(s.astype(np.float64)).astype(np.int64)
    r2 = np.minimum((w + 1) >> 1, 65535).astype(np.uint16).reshape(-1)
    assert np.array_equal(r2, t)
    assert t[127 | (127 << 8)] == 290 and t[255 | (255 << 8)] == 65535


def test_icao_filter_semantics(oracle_mod):
    O = oracle_mod
    o = O.Oracle1090()
    assert not o.filter_test(0x4B9696)
    o.filter_add(0x4B9696)
    assert o.filter_test(0x4B9696)
    assert not o.filter_test(0x4B9696 | (1 << 25))      # DF18 keys live beside, not on top of, plain ones
    o.filter_add(0x4B9696 | (1 << 25))
    assert o.filter_test(0x4B9696 | (1 << 25))
    # table-full behaviour (icao_filter.c:71-84): 4096 slots, further adds are dropped silently
    for a in range(1, 5000):
        o.filter_add(a)
    assert o.filter_test(1) and o.filter_test(4000)
    assert not o.filter_test(4999)


def test_scores(oracle_mod):
    O = oracle_mod
    o = O.Oracle1090(2, 1)
    frame = bytes.fromhex("8D4B969699155600E87406F5B69F")
    assert o.score(frame) == 25                      # DF17_UNKNOWN, as the reference prints for a first-seen DF17
    o.filter_add(0x4B9696)
    assert o.score(frame) == 27                      # DF17_KNOWN
    one = bytearray(frame); one[5] ^= 0x10
    assert o.score(bytes(one)) == 22                 # DF17_1ERROR_KNOWN
    two = bytearray(one); two[9] ^= 0x01
    assert o.score(bytes(two)) in (14, 3)            # 2-bit repair only if the syndrome survived alias pruning
    assert o.score(bytes(14)) == 1                   # ALL_ZEROS
    assert O.Oracle1090(0, 1).score(bytes(one)) == 3  # no repair tables -> UNCORRECTABLE


def test_empty_and_tiny_feeds(oracle_mod):
    O = oracle_mod
    o = O.Oracle1090()
    assert len(o.feed(np.zeros(0, dtype=np.uint8))) == 0
    assert len(o.feed(np.full(2 * 100, 127, dtype=np.uint8))) == 0
