"""dump1090's surveillance field decode (decodeModesMessage outside ME / MB, mode_s.c:626-779).
tests/golden/fields_a.npz holds what the reference binary's own struct modesMessage carried for 2079 accepted messages of a
dense mixed-DF capture (tests/make_golden_fields.py).  For DF17 / DF18 the reference's extended-squitter decoder refines the
derived fields afterwards (not built): there only the header fields are compared."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "fields_a.npz")
RAW = ("AC", "ID", "CA", "CC", "CF", "DR", "FS", "KE", "ND", "RI", "SL", "UM", "VS", "es_pending")
DERIVED = ("airground", "alert_valid", "alert", "spi_valid", "spi", "altitude_baro_valid", "altitude_baro_unit", "squawk_valid", "squawk",
           "altitude_baro")


def check(got, ref, msgs):
    es = (msgs["msgtype"] == 17) | (msgs["msgtype"] == 18)
    for f in RAW:
        assert np.array_equal(got[f], ref[f]), f
    for f in DERIVED:
        assert np.array_equal(got[f][~es], ref[f][~es]), f


def random_messages(dtype, n, seed):
    rng = np.random.default_rng(seed)
    m = np.zeros(n, dtype)
    m["msg"] = rng.integers(0, 256, (n, 14), dtype=np.uint8)
    m["msgtype"] = rng.choice([0, 4, 5, 11, 16, 17, 18, 20, 21, 24], n)
    return m


def test_oracle_fields_equal_reference_fixture(oracle_mod):
    with np.load(GOLD) as z:
        msgs, ref = z["msgs"], z["fields"]
    assert {0, 4, 5, 11, 16, 17, 18, 20, 21} <= set(int(d) for d in np.unique(msgs["msgtype"]))
    assert (ref["altitude_baro_valid"] == 1).sum() > 100 and (ref["squawk_valid"] == 1).sum() > 100
    check(oracle_mod.decode_fields(msgs), ref, msgs)


@pytest.mark.gpu
def test_gpu_fields_equal_reference_fixture_and_oracle(adsbcu_mod, oracle_mod):
    with np.load(GOLD) as z:
        msgs, ref = z["msgs"], z["fields"]
    d = adsbcu_mod.Demod1090()
    m = np.zeros(len(msgs), adsbcu_mod.MSG_DTYPE)
    for f in adsbcu_mod.MSG_DTYPE.names:
        m[f] = msgs[f]
    check(d.decode_fields_a(m), ref, msgs)
    # every value of the 13-bit altitude / identity fields (Gillham and 25 ft encodings, the M bit), all formats
    r = random_messages(adsbcu_mod.MSG_DTYPE, 4 * 8192, 13)
    r["msg"][:, 2] = (r["msg"][:, 2] & 0xE0) | ((np.arange(len(r)) >> 8) & 0x1F)
    r["msg"][:, 3] = np.arange(len(r)) & 0xFF
    r["msgtype"][:8192], r["msgtype"][8192:16384], r["msgtype"][16384:24576], r["msgtype"][24576:] = 4, 5, 20, 21
    got = d.decode_fields_a(r)
    want = oracle_mod.decode_fields(r)
    assert got.tobytes() == want.tobytes()
    r2 = random_messages(adsbcu_mod.MSG_DTYPE, 20000, 14)
    assert d.decode_fields_a(r2).tobytes() == oracle_mod.decode_fields(r2).tobytes()
    assert len(d.decode_fields_a(r2[:0])) == 0
