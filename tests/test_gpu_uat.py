"""GPU suite for Path C (adsbcu_978_feed): frame-for-frame, byte-for-byte equal to the CPU restatement
(oracle/oracle978.c; parity with dump978 itself is unpinned, see its header) plus round-trip properties."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def O():
    from oracle import oracle978 as O
    O.lib()
    return O


def same(a, b):
    return len(a) == len(b) and a.tobytes() == b.tobytes()


def test_phase_exhaustive(adsbcu_mod, O):
    idx = np.arange(65536)
    iq = np.zeros(2 * 65536, dtype=np.uint8)
    iq[0::2], iq[1::2] = idx & 255, idx >> 8
    assert np.array_equal(adsbcu_mod.Demod1090().uat_phase(iq), O.Oracle978().phase_table())


@pytest.mark.parametrize("seed,sigma", [(1, 1.0), (2, 2.0), (3, 4.0)])
def test_seeded_uat_vs_oracle(adsbcu_mod, O, seed, sigma):
    from libadsb_b200 import synth978 as S
    iq, sent = S.config5(seconds=2.0, seed=seed, sigma=sigma, adsb_per_s=400.0, uplink_per_s=4.0)
    want = O.Oracle978().handle_data(iq)
    d = adsbcu_mod.Demod1090()
    assert same(want, d.handle_data_uat(iq))                       # whole capture in one call
    assert len(want) > 0.5 * len(sent)
    d.reset_uat()                                                  # HandleData-sized calls (262144 bytes, RTLSDR.hpp:55)
    o = O.Oracle978()
    for k in range(0, iq.size, 262144):
        assert same(o.handle_data(iq[k:k + 262144]), d.handle_data_uat(iq[k:k + 262144]))
    d.reset_uat()                                                  # ragged calls, some far shorter than one frame
    o, rng, k = O.Oracle978(), np.random.default_rng(seed), 0
    while k < iq.size:
        nb = 2 * int(rng.choice([10, 999, 5000, 20000, 65536, 70001, 300000]))
        assert same(o.handle_data(iq[k:k + nb]), d.handle_data_uat(iq[k:k + nb]))
        k += nb


def test_uat_edges(adsbcu_mod, O):
    d, o = adsbcu_mod.Demod1090(), O.Oracle978()
    assert len(d.handle_data_uat(np.zeros(0, dtype=np.uint8))) == 0
    rng = np.random.default_rng(4)
    iq = rng.integers(0, 256, size=2 * 700000, dtype=np.uint8)     # full-scale noise: random sync hits, no frames
    assert same(o.handle_data(iq), d.handle_data_uat(iq))
    from libadsb_b200 import synth978 as S
    # back-to-back frames: the scan's skip over a decoded frame and its stale sync registers come into play
    bits = [S.frame_bits(S.ADSB_SYNC, S.adsb_frame(bytes([8 + i] + [i] * 33))) for i in range(40)]
    starts = np.cumsum([50000.0] + [2 * len(b) + g for b, g in zip(bits[:-1], [0, 1, 2, 3, 7, 20, 71, 72, 73, 100] * 4)])[:40]
    iq = S.synth_capture(200000, bits, starts, np.full(40, 50.0), 1.0, rng)
    want, got = O.Oracle978().handle_data(iq), adsbcu_mod.Demod1090().handle_data_uat(iq)
    assert len(want) >= 30 and same(want, got)


def test_gpu_uat_equals_reference_fixture(adsbcu_mod):
    """tests/golden/uat_ref.npz was produced by the reference's own UAT978.cpp compiled in place (phase table, staging loop with
    its short memmove, UAT978.cpp:43-60,76-100; tests/make_golden_978.py): same table, same frames at the same offsets."""
    import os
    from libadsb_b200 import synth978 as S
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "uat_ref.npz")) as z:
        table, offs, rs, ln = z["phase_table"], z["frame_offsets"], z["frame_rs"], z["frame_len"]
        iq = S.config5(seconds=float(z["seconds"]), seed=int(z["seed"]))[0]
        buf = int(z["buf"])
    d = adsbcu_mod.Demod1090()
    idx = np.arange(65536)
    probe = np.zeros(2 * 65536, dtype=np.uint8)
    probe[0::2], probe[1::2] = idx & 255, idx >> 8
    assert np.array_equal(d.uat_phase(probe), table)
    frames = np.concatenate([d.handle_data_uat(iq[k:k + buf]) for k in range(0, iq.size, buf)])
    assert np.array_equal(frames["offset"], offs) and np.array_equal(frames["rs_errors"], rs) and np.array_equal(frames["len"], ln)


def test_gpu_uat_device_feed_equals_host_feed(adsbcu_mod):
    import os
    from libadsb_b200 import synth978 as S
    iq = S.config5(seconds=1.0, seed=12)[0]
    d = adsbcu_mod.Demod1090()
    want = d.handle_data_uat(iq)
    d.reset_uat()
    if os.environ.get("ADSBCU_TEST_EMU") == "1":
        keep, ptr = iq, iq.ctypes.data
    else:
        import torch
        keep = torch.from_numpy(iq).cuda()
        ptr = keep.data_ptr()
    got = d.handle_data_uat_device(ptr, iq.size, np.zeros(1 << 14, adsbcu_mod.UAT_DTYPE))
    assert len(want) > 100 and same(want, got)
    assert d.uat_stats()["calls"] >= 30
