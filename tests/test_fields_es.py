"""dump1090's extended-squitter payload decode (decodeExtendedSquitter and the decodeES* functions, mode_s.c:798-1551;
SURVEY 8f rank 1).  tests/golden/fields_es.npz holds 8419 DF17 / DF18 frames (every ME type, every CF, 419 of them repaired
by the reference against its own filter) and what the reference binary's struct modesMessage carried after
decodeModesMessage (tests/make_golden_es.py, which also pins the restatement on ten million frames when it is run).
Integer fields must match exactly; the float fields are compared as bit patterns too."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "fields_es.npz")


def differences(a, b):
    out = []
    assert a.dtype.names == b.dtype.names
    for name in a.dtype.names:
        x, y = a[name], b[name]
        if x.dtype.kind == "f":
            x, y = x.view(np.uint32), y.view(np.uint32)
        bad = np.nonzero(x != y)[0]
        if len(bad):
            out.append((name, len(bad), int(bad[0])))
    return out


def crc24(frames):
    rem = np.zeros(len(frames), np.uint32)
    for byte in range(11):
        rem ^= frames[:, byte].astype(np.uint32) << 16
        for _ in range(8):
            top = (rem & 0x800000) != 0
            rem = (rem << 1) & 0xFFFFFF
            rem[top] ^= 0xFFF409
    return rem


def as_msgs(dtype, frames, corrected=None):
    m = np.zeros(len(frames), dtype)
    m["msg"] = frames
    m["verbatim"] = frames
    m["msgbits"] = 112
    m["msgtype"] = frames[:, 0] >> 3
    if corrected is not None:
        m["correctedbits"] = corrected
    return m


def velocity_frames(sub, signs):
    """Every (E/W, N/S) raw pair of an airborne velocity message (ME type 19) of one subtype and one pair of direction bits."""
    ew, ns = np.meshgrid(np.arange(1024, dtype=np.uint64), np.arange(1024, dtype=np.uint64), indexing="ij")
    ew, ns = ew.ravel(), ns.ravel()
    me = (np.uint64(19) << np.uint64(51)) | (np.uint64(sub) << np.uint64(48)) | (np.uint64(signs >> 1) << np.uint64(42)) | \
        (ew << np.uint64(32)) | (np.uint64(signs & 1) << np.uint64(31)) | (ns << np.uint64(21))
    f = np.zeros((len(me), 14), np.uint8)
    f[:, 0] = 17 << 3 | 5
    f[:, 1:4] = (0x4B, 0x17, 0x2A)
    for b in range(7):
        f[:, 4 + b] = (me >> np.uint64(8 * (6 - b))) & np.uint64(255)
    p = crc24(f)
    f[:, 11], f[:, 12], f[:, 13] = p >> 16, (p >> 8) & 255, p & 255
    return f


def test_oracle_es_equals_reference_fixture(oracle_mod):
    with np.load(GOLD) as z:
        frames, corrected, ref = z["frames"], z["corrected"], z["fields"]
    assert set(range(32)) == set(int(t) for t in np.unique(ref["metype"])) and int(corrected.sum()) > 100
    for bit in range(31):                                       # every *_valid flag occurs
        assert (ref["valid"] & (1 << bit)).any(), bit
    assert not differences(oracle_mod.decode_es(frames, corrected), ref.view(oracle_mod.FIELDS_ES_DTYPE))


def test_oracle_es_other_formats_are_empty(oracle_mod):
    rng = np.random.default_rng(5)
    f = rng.integers(0, 256, (2000, 14), dtype=np.uint8)
    f[:, 0] = (rng.choice([0, 4, 5, 11, 16, 20, 21, 24], 2000).astype(np.uint8) << 3) | (f[:, 0] & 7)
    assert not oracle_mod.decode_es(f).tobytes().strip(b"\0")


@pytest.mark.gpu
def test_gpu_es_equals_reference_fixture(adsbcu_mod):
    with np.load(GOLD) as z:
        frames, corrected, ref = z["frames"], z["corrected"], z["fields"]
    d = adsbcu_mod.Demod1090()
    got = d.decode_es(as_msgs(adsbcu_mod.MSG_DTYPE, frames, corrected))
    assert not differences(got, ref.view(adsbcu_mod.FIELDS_ES_DTYPE))
    assert len(d.decode_es(as_msgs(adsbcu_mod.MSG_DTYPE, frames[:0]))) == 0


@pytest.mark.gpu
def test_gpu_es_equals_oracle_on_random_payloads(adsbcu_mod, oracle_mod):
    rng = np.random.default_rng(1447)
    n = 300000 if os.environ.get("ADSBCU_TEST_EMU") != "1" else 60000
    f = rng.integers(0, 256, (n, 14), dtype=np.uint8)
    df = rng.choice([17, 18, 17, 18, 11, 4, 20], n).astype(np.uint8)
    f[:, 0] = (df << 3) | (f[:, 0] & 7)
    f[: n // 2, 4] = (rng.choice([19, 29, 31, 5, 11], n // 2).astype(np.uint8) << 3) | (f[: n // 2, 4] & 7)     # the long decoders
    cor = rng.integers(0, 3, n).astype(np.uint8)
    d = adsbcu_mod.Demod1090()
    got = d.decode_es(as_msgs(adsbcu_mod.MSG_DTYPE, f, cor))
    want = oracle_mod.decode_es(f, cor)
    assert not differences(got, want.view(adsbcu_mod.FIELDS_ES_DTYPE))
    assert got.tobytes() == want.tobytes()                      # padding and unused fields too
    assert (got["es"] == ((df == 17) | (df == 18))).all()


@pytest.mark.gpu
def test_gpu_es_velocity_every_component_pair(adsbcu_mod, oracle_mod):
    """sqrtf and atan2 of the velocity components (mode_s.c:894-902) for every raw (E/W, N/S) pair, both subtypes, all four
    direction combinations: the device's double atan2 against glibc's, narrowed to the reference's float - bit for bit."""
    d = adsbcu_mod.Demod1090()
    combos = [(1, 0), (2, 3)] if os.environ.get("ADSBCU_TEST_EMU") == "1" else [(s, q) for s in (1, 2) for q in range(4)]
    for sub, signs in combos:
        f = velocity_frames(sub, signs)
        got = d.decode_es(as_msgs(adsbcu_mod.MSG_DTYPE, f))
        want = oracle_mod.decode_es(f).view(adsbcu_mod.FIELDS_ES_DTYPE)
        assert (want["valid"] & adsbcu_mod.ES_VALID["heading"]).sum() > 1000000
        assert not differences(got, want), (sub, signs)


def check_capture_es(es, ref, msgs, valid_bits):
    """fields_a.npz carries what the reference's struct modesMessage held for the accepted messages of a real (synthetic-air)
    capture AFTER its ES decoder ran: for the DF17 / DF18 messages the ES record must give exactly those final values."""
    sel = (msgs["msgtype"] == 17) | (msgs["msgtype"] == 18)
    assert sel.sum() > 300 and (es["es"] == sel).all()
    e, r = es[sel], ref[sel]
    assert np.array_equal(e["airground"], r["airground"])
    for name, bit in (("alert", "alert"), ("spi", "spi"), ("altitude_baro", "altitude_baro"), ("squawk", "squawk")):
        have = (e["valid"] & valid_bits[bit]) != 0
        assert np.array_equal(have, r[name + "_valid"] == 1), name
        assert np.array_equal(e[name][have], r[name][have]), name
    assert np.array_equal(e["altitude_baro_unit"], r["altitude_baro_unit"])
    assert np.array_equal(e["CA_CF"], np.where(msgs["msgtype"][sel] == 17, r["CA"], r["CF"]))
    assert np.array_equal(e["addr"], msgs["addr"][sel])                         # incl. the non-ICAO flag of the demodulator's record


def test_oracle_es_gives_the_final_values_of_the_reference_capture(oracle_mod):
    with np.load(os.path.join(os.path.dirname(GOLD), "fields_a.npz")) as z:
        msgs, ref = z["msgs"], z["fields"]
    bits = {"alert": 1 << 14, "spi": 1 << 13, "altitude_baro": 1 << 0, "squawk": 1 << 8}
    check_capture_es(oracle_mod.decode_es(msgs["msg"], msgs["correctedbits"]), ref, msgs, bits)


@pytest.mark.gpu
def test_gpu_es_gives_the_final_values_of_the_reference_capture(adsbcu_mod):
    with np.load(os.path.join(os.path.dirname(GOLD), "fields_a.npz")) as z:
        msgs, ref = z["msgs"], z["fields"]
    m = np.zeros(len(msgs), adsbcu_mod.MSG_DTYPE)
    for f in adsbcu_mod.MSG_DTYPE.names:
        m[f] = msgs[f]
    check_capture_es(adsbcu_mod.Demod1090().decode_es(m), ref, msgs, adsbcu_mod.ES_VALID)
