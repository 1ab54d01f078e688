"""bulk field decode and CPR for libadsb's own handler (ADSB1090.cpp:525-672, :986-1121).
CPU part: the oracle restatement against the reference's own DecodeModesMessage / DecodeCpr (oracle/_ref).
GPU part: adsbcu_1090b_decode_fields / adsbcu_1090b_cpr_decode against the oracle, bit-exact."""
import os

import numpy as np
import pytest

from oracle import oracle1090b as B

HAVE_REF = os.path.exists(B.REF_PATH)


def crc24(msg: bytes) -> int:
    return B.Oracle1090B().checksum(msg, len(msg) * 8)


def with_parity(body: bytes) -> bytes:
    """body = the frame without its last three bytes -> frame whose CRC field equals the checksum (DF11/17 style)."""
    c = crc24(body + b"\0\0\0")
    return body + bytes([(c >> 16) & 255, (c >> 8) & 255, c & 255])


def frames(n, seed):
    """Random 14-byte frames biased to the branches of the field decoder: half DF17, a quarter of them velocity."""
    rng = np.random.default_rng(seed)
    m = rng.integers(0, 256, (n, 14), dtype=np.uint8)
    m[1::2, 0] = (17 << 3) | (m[1::2, 0] & 7)
    m[3::4, 4] = (19 << 3) | rng.integers(1, 5, m[3::4].shape[0])
    m[5::8, 4] = (rng.integers(1, 5, m[5::8].shape[0]) << 3) | (m[5::8, 4] & 7)
    m[7::16, 0] = (np.array([0, 4, 16, 20])[rng.integers(0, 4, m[7::16].shape[0])] << 3) | (m[7::16, 0] & 7)
    return m


def pairs(n, seed):
    """Half uniformly random 17-bit values (mostly rejected by the zone check), half near-consistent pairs: the odd
    frame's values are what the same aircraft would send, derived from the even ones through the decoder's own
    zone arithmetic with a small perturbation."""
    rng = np.random.default_rng(seed)
    p = np.zeros(n, dtype=[("even_lat", "<i4"), ("even_lon", "<i4"), ("odd_lat", "<i4"), ("odd_lon", "<i4"), ("even_is_newer", "<i4")])
    for f in ("even_lat", "even_lon", "odd_lat", "odd_lon"):
        p[f] = rng.integers(0, 131072, n)
    p["even_is_newer"] = rng.integers(0, 2, n)
    # consistent pairs: pick a latitude/longitude, encode both parities
    k = n // 2
    lat = rng.uniform(-89.5, 89.5, k)
    lon = rng.uniform(-180, 180, k)
    for odd, (fl, fo) in enumerate((("even_lat", "even_lon"), ("odd_lat", "odd_lon"))):
        dlat = 360.0 / (60 - odd)
        yz = np.floor(131072 * (np.mod(lat, dlat) / dlat) + 0.5)
        rlat = dlat * (yz / 131072 + np.floor(lat / dlat))
        nl = np.array([nl_of(v) for v in rlat])
        dlon = 360.0 / np.maximum(nl - odd, 1)
        xz = np.floor(131072 * (np.mod(lon, dlon) / dlon) + 0.5)
        p[fl][:k] = yz.astype(np.int64) & 0x1FFFF
        p[fo][:k] = xz.astype(np.int64) & 0x1FFFF
    return p


def nl_of(lat):
    lat = abs(lat)
    if lat >= 87:
        return 1
    if lat == 0:
        return 59
    a = 1 - np.cos(np.pi / 30)
    return int(np.floor(2 * np.pi / np.arccos(1 - a / np.cos(np.pi / 180 * lat) ** 2)))


# ---- CPU: oracle pinned against the reference ----------------------------------------------------------------

@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_fields_match_reference():
    o, r = B.Oracle1090B(), B.Ref1090B()
    for m in frames(20000, 11):
        assert o.fields(m.tobytes()).tobytes() == r.fields(m.tobytes()).tobytes(), m.tobytes().hex()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_fields_repaired_df_bit():
    """A repair inside the DF field leaves mm.msgtype as sliced (ADSB1090.cpp:501,513-519): the fields follow it."""
    o, r = B.Oracle1090B(), B.Ref1090B()
    body = bytes([16 << 3 | 5]) + bytes(range(1, 11))           # a DF16-looking frame with plain parity
    good = bytearray(with_parity(body))
    good[0] ^= 0x08                                               # bit 4: now sliced as DF17 with a single-bit error
    d = o.decode(bytes(good))
    assert d["msgtype"] == 17 and d["errorbit"] == 4 and d["msg"][0] >> 3 == 16
    assert o.fields(bytes(good)).tobytes() == r.fields(bytes(good)).tobytes()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_velocity_heading_all_octants():
    o, r = B.Oracle1090B(), B.Ref1090B()
    for ewd in (0, 1):
        for nsd in (0, 1):
            for ewv, nsv in ((0, 0), (0, 5), (5, 0), (5, 5), (1023, 1023), (1, 1023), (1023, 1), (300, 400), (17, 999)):
                m = bytearray(14)
                m[0] = 17 << 3 | 5
                m[4] = 19 << 3 | 1
                m[5] = ewd << 2 | ewv >> 8
                m[6] = ewv & 255
                m[7] = nsd << 7 | nsv >> 3
                m[8] = (nsv & 7) << 5
                assert o.fields(bytes(m)).tobytes() == r.fields(bytes(m)).tobytes(), (ewd, nsd, ewv, nsv)


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_cpr_matches_reference():
    o, r = B.Oracle1090B(), B.Ref1090B()
    p = pairs(40000, 5)
    ok = 0
    for q in p:
        a, b = o.cpr(*[int(x) for x in q]), r.cpr(*[int(x) for x in q])
        assert a == b, q
        ok += a[0]
    assert ok > p.size // 3          # the consistent half decodes


def test_oracle_cpr_known_position():
    """The worked example every CPR write-up uses (even 93000/51372, odd 74158/50194): 52.2572 N, 3.9194 E."""
    o = B.Oracle1090B()
    ok, lat, lon = o.cpr(93000, 51372, 74158, 50194, 1)
    assert ok == 1 and abs(lat / 1e7 - 52.2572021484375) < 1e-6 and abs(lon / 1e7 - 3.91937255859375) < 1e-6
    ok, lat, lon = o.cpr(93000, 51372, 74158, 50194, 0)
    assert ok == 1 and abs(lat / 1e7 - 52.26578017412606) < 1e-6 and abs(lon / 1e7 - 3.938912527901786) < 1e-6


# ---- GPU: the C ABI against the oracle -----------------------------------------------------------------------

def to_records(o, raw):
    """What adsbcu_1090b_feed would hand out for these frames: msg after the CRC stage, msgtype as sliced."""
    from libadsb_b200 import adsbcu
    rec = np.zeros(raw.shape[0], dtype=adsbcu.MSG_B_DTYPE)
    for k, m in enumerate(raw):
        d = o.decode(m.tobytes())
        rec["msg"][k] = np.frombuffer(d["msg"], dtype=np.uint8)
        rec["msgbits"][k] = d["msgbits"]
        rec["msgtype"][k] = d["msgtype"]
        rec["errorbit"][k] = d["errorbit"]
    return rec


@pytest.mark.gpu
def test_gpu_fields_match_oracle():
    from libadsb_b200 import adsbcu
    o = B.Oracle1090B()
    raw = frames(30000, 21)
    rec = to_records(o, raw)
    want = np.concatenate([o.fields(m.tobytes()).reshape(1) for m in raw])
    got = adsbcu.Demod1090().decode_fields_b(rec)
    bad = np.nonzero([a.tobytes() != b.tobytes() for a, b in zip(got, want)])[0]
    assert bad.size == 0, (raw[bad[0]].tobytes().hex(), got[bad[0]], want[bad[0]])


@pytest.mark.gpu
def test_gpu_heading_exhaustive():
    """All 2^22 velocity vectors: heading and speed equal the oracle's (glibc atan2 / sqrt)."""
    from libadsb_b200 import adsbcu
    i = np.arange(1 << 22, dtype=np.uint32)
    ewd, nsd, ewv, nsv = i >> 21, (i >> 20) & 1, (i >> 10) & 1023, i & 1023
    rec = np.zeros(i.size, dtype=adsbcu.MSG_B_DTYPE)
    rec["msgtype"] = 17
    rec["msgbits"] = 112
    m = rec["msg"]
    m[:, 0] = 17 << 3 | 5
    m[:, 4] = 19 << 3 | 1
    m[:, 5] = ewd << 2 | ewv >> 8
    m[:, 6] = ewv & 255
    m[:, 7] = nsd << 7 | nsv >> 3
    m[:, 8] = (nsv & 7) << 5
    got = adsbcu.Demod1090().decode_fields_b(rec)
    sx = np.where(ewd == 1, -ewv.astype(np.int64), ewv.astype(np.int64))
    sy = np.where(nsd == 1, -nsv.astype(np.int64), nsv.astype(np.int64))
    vel = np.sqrt((sx * sx + sy * sy).astype(np.float64)).astype(np.int64)
    assert np.array_equal(got["velocity"], vel)
    assert np.array_equal(got["ew_velocity"], ewv) and np.array_equal(got["ns_velocity"], nsv)
    # a strided sample through the oracle itself (the full 4 M would take minutes in ctypes calls)
    o = B.Oracle1090B()
    for k in np.concatenate([np.arange(0, 1 << 22, 4099), np.arange(0, 2048)]):
        assert got[k].tobytes() == o.fields(rec["msg"][k].tobytes()).tobytes(), k
    # and everything against numpy's arctan2, which is the same libm call, where the velocity is non-zero
    h = (np.arctan2(sx.astype(np.float64), sy.astype(np.float64)) * 360 / (np.pi * 2)).astype(np.int64)
    h = np.where(h < 0, h + 360, h)
    h = np.where(vel == 0, 0, h)
    assert np.array_equal(got["heading"], h)


@pytest.mark.gpu
def test_gpu_cpr_matches_oracle():
    from libadsb_b200 import adsbcu
    o = B.Oracle1090B()
    p = pairs(60000, 9)
    got = adsbcu.Demod1090().cpr_decode(p)
    for q, g in zip(p, got):
        ok, lat, lon = o.cpr(*[int(x) for x in q])
        assert int(g["valid"]) == ok, q
        if ok:
            assert (int(g["lat1e7"]), int(g["lon1e7"])) == (lat, lon), q
    assert int(got["valid"].sum()) > p.size // 3


@pytest.mark.gpu
def test_gpu_fields_empty_and_after_feed():
    """n = 0 is a no-op; records straight out of adsbcu_1090b_feed decode to the oracle's fields."""
    from libadsb_b200 import adsbcu, synth
    d = adsbcu.Demod1090()
    assert d.decode_fields_b(np.zeros(0, dtype=adsbcu.MSG_B_DTYPE)).size == 0
    assert d.cpr_decode(np.zeros(0, dtype=adsbcu.CPR_PAIR_DTYPE)).size == 0
    iq = synth.config_b(seconds=0.5, seed=77)[0]
    msgs = d.handle_data_b(iq)
    assert msgs.size > 10
    got = d.decode_fields_b(msgs)
    o = B.Oracle1090B()
    for mrec, g in zip(msgs, got):
        # the frame is already repaired: the oracle's CRC stage leaves it as it is, msgtype is msg[0] >> 3 here
        if mrec["msgtype"] == mrec["msg"][0] >> 3:
            assert g.tobytes() == o.fields(mrec["msg"].tobytes()).tobytes()
