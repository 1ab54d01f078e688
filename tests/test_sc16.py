"""SC16 / SC16Q11 input converters (convert.c:46-92, magnitude_sc16*.c)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sc16.npz")
FORMATS = [("sc16", 1, False, "SC16"), ("sc16q11", 2, True, "SC16Q11")]


@pytest.mark.parametrize("key,fmt,q11,name", FORMATS)
def test_oracle_sc16_matches_reference_fixture(oracle_mod, key, fmt, q11, name):
    with np.load(GOLD) as z:
        iq, ref = z["iq_" + key], z["ref_" + key]
    o = oracle_mod.Oracle1090(2, 1)
    o.set_input_format(fmt)
    assert oracle_mod.same_messages(ref, o.feed(iq, cap=1 << 16)) == []


@pytest.mark.gpu
@pytest.mark.parametrize("key,fmt,q11,name", FORMATS)
def test_gpu_sc16_matches_reference_fixture(adsbcu_mod, oracle_mod, key, fmt, q11, name):
    with np.load(GOLD) as z:
        iq, ref = z["iq_" + key], z["ref_" + key]
    assert oracle_mod.same_messages(ref, adsbcu_mod.Demod1090(2, 1).feed_sc16(iq, q11)) == []
    # HandleData-sized and ragged feeds continue one stream; internal takes are invisible
    d = adsbcu_mod.Demod1090(2, 1, max_feed_samples=70000)
    parts, k, rng = [], 0, np.random.default_rng(3)
    while k < iq.size // 2:
        n = int(rng.choice([1, 325, 327, 5000, 131072]))
        parts.append(d.feed_sc16(iq[2 * k:2 * (k + n)], q11))
        k += n
    assert oracle_mod.same_messages(ref, np.concatenate(parts)) == []


@pytest.mark.gpu
@pytest.mark.parametrize("q11", [False, True])
def test_gpu_sc16_full_range_vs_oracle(adsbcu_mod, oracle_mod, q11):
    """Random full-range int16 samples (incl. -32768 and clamped magnitudes) around injected messages."""
    from libadsb_b200 import synth
    iq8 = synth.config2(seconds=0.2, seed=77)[0]
    iq = synth.to_sc16(iq8, q11).copy()
    rng = np.random.default_rng(9)
    idx = rng.integers(0, iq.size, size=2000)
    iq[idx] = rng.choice(np.array([-32768, 32767, -1, 0, 1, 12345, -23456], dtype=np.int16), size=idx.size)
    o = oracle_mod.Oracle1090(1, 1)
    o.set_input_format(2 if q11 else 1)
    want = o.feed(iq, cap=1 << 16)
    assert len(want) > 50
    assert oracle_mod.same_messages(want, adsbcu_mod.Demod1090(1, 1).feed_sc16(iq, q11)) == []
