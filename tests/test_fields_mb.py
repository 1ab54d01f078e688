"""dump1090's Comm-B payload decode (decodeCommB and its ten register decoders, comm_b.c:50-985).
tests/golden/fields_mb.npz holds 7780 DF20 / DF21 replies built register by register (plausible and implausible field values,
ties included) and what the reference binary's struct modesMessage carried after decodeModesMessage
(tests/make_golden_mb.py, which also pins the restatement on 2.75 million replies when it is run).  Every field must match
exactly, the floats as bit patterns."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
from make_golden_mb import build  # noqa: E402
from test_fields_es import as_msgs, differences  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden", "fields_mb.npz")


def test_oracle_commb_equals_reference_fixture(oracle_mod):
    with np.load(GOLD) as z:
        frames, ref = z["frames"], z["fields"]
    assert (np.bincount(ref["commb_format"], minlength=13) > 20).all()          # every commb_format_t value occurs
    for bit in range(20):
        assert (ref["valid"] & (1 << bit)).any(), bit
    assert not differences(oracle_mod.decode_commb(frames), ref.view(oracle_mod.FIELDS_MB_DTYPE))


def test_oracle_commb_repaired_and_other_formats(oracle_mod):
    with np.load(GOLD) as z:
        frames = z["frames"][:500]
    got = oracle_mod.decode_commb(frames, np.ones(len(frames), np.uint8))        # comm_b.c:55: nothing is decoded from a repaired reply
    assert (got["commb_format"] == 2).all() and (got["valid"] == 0).all()
    other = frames.copy()
    other[:, 0] = (17 << 3) | (other[:, 0] & 7)
    assert not oracle_mod.decode_commb(other).tobytes().strip(b"\0")


def test_bds50_turn_rate_test_has_margin():
    """comm_b.c:584-590 compares |68625 tan(roll) / (20 pi tas) - track_rate| with 2.0 in double.  The device's tan is not
    glibc's bit for bit (a few ulp); over every (roll, tas, track rate) the register accepts, the quantity never comes
    closer to 2.0 than 1e-9 - except at roll == 0, where tan is exactly 0 on both sides - so they cannot disagree on the penalty."""
    raw = np.arange(512, dtype=np.float64)
    roll = np.concatenate([raw * 45.0 / 256.0, raw * 45.0 / 256.0 - 90.0]).astype(np.float32).astype(np.float64)
    roll = roll[(roll >= -40) & (roll < 40) & (roll != 0)]
    tas = np.arange(25, 351, dtype=np.float64) * 2
    rate = np.concatenate([raw * 8.0 / 256.0, raw * 8.0 / 256.0 - 16.0])
    rate = np.unique(rate[(rate >= -10) & (rate <= 10)])
    implied = (68625 * np.tan(roll * np.pi / 180.0))[:, None] / (tas * 20 * np.pi)[None, :]
    margin = np.inf
    for r in rate:
        margin = min(margin, float(np.abs(np.abs(implied - r) - 2.0).min()))
    assert margin > 1e-9, margin


@pytest.mark.gpu
def test_gpu_commb_equals_reference_fixture(adsbcu_mod):
    with np.load(GOLD) as z:
        frames, ref = z["frames"], z["fields"]
    d = adsbcu_mod.Demod1090()
    got = d.decode_commb(as_msgs(adsbcu_mod.MSG_DTYPE, frames))
    assert not differences(got, ref.view(adsbcu_mod.FIELDS_MB_DTYPE))
    assert len(d.decode_commb(as_msgs(adsbcu_mod.MSG_DTYPE, frames[:0]))) == 0


@pytest.mark.gpu
def test_gpu_commb_equals_oracle_on_generated_replies(adsbcu_mod, oracle_mod):
    emu = os.environ.get("ADSBCU_TEST_EMU") == "1"
    rng = np.random.default_rng(50)
    frames = build(rng, 3000 if emu else 60000)
    plain = rng.integers(0, 256, (20000, 14), dtype=np.uint8)                     # any downlink format, any payload
    frames = np.concatenate([frames, plain])
    cor = (rng.random(len(frames)) < 0.05).astype(np.uint8)
    d = adsbcu_mod.Demod1090()
    got = d.decode_commb(as_msgs(adsbcu_mod.MSG_DTYPE, frames, cor))
    want = oracle_mod.decode_commb(frames, cor)
    assert not differences(got, want.view(adsbcu_mod.FIELDS_MB_DTYPE))
    assert got.tobytes() == want.tobytes()
    assert (np.bincount(got["commb_format"], minlength=13) > 10).all()
