"""GPU suite for Path B (adsbcu_1090b_feed): bit-exact against the oracle restatement of ADSB1090.cpp and,
through the golden fixture, against what the reference's own handler accepted."""
import os

import numpy as np
import pytest

from conftest import GOLD

pytestmark = pytest.mark.gpu
BUF = 262144


@pytest.fixture(scope="module")
def B():
    from oracle import oracle1090b as B
    return B


def same(a, b):
    return len(a) == len(b) and a.tobytes() == b.astype(a.dtype).tobytes()


def test_golden_b(adsbcu_mod, B):
    with np.load(os.path.join(GOLD, "b_mixed.npz")) as z:
        g = {k: z[k] for k in z.files}
    d = adsbcu_mod.Demod1090()
    iq, k0 = g["iq"], 0
    for b, nref in enumerate(g["per_buffer"]):
        got = d.handle_data_b(iq[b * BUF:(b + 1) * BUF])
        assert np.array_equal(got["addr"], g["ref_callbacks"]["addr"][k0:k0 + nref])     # the reference itself
        assert same(g["oracle_msgs"][k0:k0 + nref], got)                                   # every field
        k0 += nref
    assert k0 > 1000


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_seeded_b_vs_oracle(adsbcu_mod, B, seed):
    from libadsb_b200 import synth
    iq = synth.config_b(seconds=0.8, seed=seed, msgs_per_s=2500.0, flip_frac=0.3)[0]
    o, d = B.Oracle1090B(), adsbcu_mod.Demod1090()
    rng = np.random.default_rng(seed)
    k, total = 0, 0
    while k < iq.size:                                   # ragged HandleData calls, some shorter than one frame
        nb = 2 * int(rng.choice([100, 480, 1000, 5001, 40000, 131072, 400000]))
        want, got = o.handle_data(iq[k:k + nb]), d.handle_data_b(iq[k:k + nb])
        assert same(want, got)
        total += len(got)
        k += nb
    assert total > 400
    # whole capture in one call, as tests/test_1090.cpp:55-72 does; a reset handle forgets the ICAO cache
    d.reset_b()
    assert same(B.Oracle1090B().handle_data(iq), d.handle_data_b(iq))


def test_edge_inputs_b(adsbcu_mod, B):
    d, o = adsbcu_mod.Demod1090(), B.Oracle1090B()
    assert len(d.handle_data_b(np.zeros(0, dtype=np.uint8))) == 0
    for fill in (0, 127, 255):
        iq = np.full(2 * 100000, fill, dtype=np.uint8)
        assert len(d.handle_data_b(iq)) == len(o.handle_data(iq)) == 0
    rng = np.random.default_rng(9)                        # full-scale noise: many preambles and random frames
    iq = rng.integers(0, 256, size=2 * 600000, dtype=np.uint8)
    assert same(o.handle_data(iq), d.handle_data_b(iq))
