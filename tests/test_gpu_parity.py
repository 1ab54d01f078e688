"""GPU suite (-m gpu): the CUDA path, called through the C ABI, against the reference-generated
golden fixtures, against the oracle on seeded inputs, and through size-independent properties."""
import numpy as np
import pytest

from conftest import CONFIGS, GOLDEN_NAMES

pytestmark = pytest.mark.gpu


def test_magnitude_exhaustive(adsbcu_mod, oracle_mod):
    """All 65536 (I,Q) inputs: register-computed magnitude == the reference LUT (tables.c:9-41)."""
    d = adsbcu_mod.Demod1090()
    idx = np.arange(65536)
    iq = np.zeros(2 * 65536, dtype=np.uint8)
    iq[0::2], iq[1::2] = idx & 255, idx >> 8
    assert np.array_equal(d.magnitude(iq), oracle_mod.mag_table())
    # ragged length (tail handled by the scalar path)
    rng = np.random.default_rng(5)
    iq = rng.integers(0, 256, size=2 * 100003, dtype=np.uint8)
    lut = oracle_mod.mag_table()
    assert np.array_equal(d.magnitude(iq), lut[iq[0::2].astype(np.int64) | (iq[1::2].astype(np.int64) << 8)])


@pytest.mark.parametrize("name", GOLDEN_NAMES)
@pytest.mark.parametrize("nfix,fdf", CONFIGS)
def test_golden(adsbcu_mod, oracle_mod, golden, name, nfix, fdf):
    g = golden[name]
    got = adsbcu_mod.Demod1090(nfix, fdf).feed(g["iq"])
    assert oracle_mod.same_messages(g[f"ref_fix{nfix}_df{fdf}"], got) == []


@pytest.mark.parametrize("nfix,fdf", CONFIGS)
def test_seeded_vs_oracle(adsbcu_mod, oracle_mod, nfix, fdf):
    from libadsb_b200 import synth
    for iq in (synth.config2(seconds=2.0, seed=77 + nfix)[0],
               synth.config3(seconds=1.0, seed=99 + fdf, msgs_per_s=6000.0, pool=1500)[0]):
        want = oracle_mod.Oracle1090(nfix, fdf).feed(iq)
        got = adsbcu_mod.Demod1090(nfix, fdf).feed(iq)
        assert len(want) > 500
        assert oracle_mod.same_messages(want, got) == []


@pytest.mark.parametrize("block", [4096, 16384])
def test_small_mag_buf_boundaries(adsbcu_mod, oracle_mod, block):
    """Shrinking the reference's mag_buf makes the boundary rule (demod_2400.c:432-443) fire often."""
    from libadsb_b200 import synth
    iq = synth.config3(seconds=0.5, seed=4242, msgs_per_s=9000.0, pool=200)[0]
    want = oracle_mod.Oracle1090(2, 1, block).feed(iq)
    got = adsbcu_mod.Demod1090(2, 1, block_samples=block).feed(iq)
    assert len(want) > 500
    assert oracle_mod.same_messages(want, got) == []


def test_streaming_feeds_equal_one_shot(adsbcu_mod, oracle_mod, golden):
    """HandleData-sized feeds (131072 samples = 262144 bytes, RTLSDR.hpp:55) continue one stream."""
    iq = golden["c3_dense"]["iq"]
    whole = adsbcu_mod.Demod1090(2, 1).feed(iq)
    d = adsbcu_mod.Demod1090(2, 1)
    parts = [d.feed(iq[k:k + 262144]) for k in range(0, iq.size, 262144)]
    assert oracle_mod.same_messages(whole, np.concatenate(parts)) == []
    # internal chunking (max_feed_samples smaller than the feed) is invisible too
    d2 = adsbcu_mod.Demod1090(2, 1, max_feed_samples=50000)
    assert oracle_mod.same_messages(whole, d2.feed(iq)) == []
    # tiny ragged feeds, including ones shorter than the 326-sample overlap
    d3 = adsbcu_mod.Demod1090(2, 1)
    rng = np.random.default_rng(1)
    parts, k = [], 0
    while k < iq.size // 2:
        n = int(rng.choice([1, 7, 100, 325, 326, 327, 5000, 40000]))
        parts.append(d3.feed(iq[2 * k:2 * (k + n)]))
        k += n
    assert oracle_mod.same_messages(whole, np.concatenate(parts)) == []


def test_device_feed_equals_host_feed(adsbcu_mod, oracle_mod, golden):
    import torch
    iq = golden["c2_sparse"]["iq"]
    want = adsbcu_mod.Demod1090(1, 1).feed(iq)
    t = torch.from_numpy(iq).cuda()
    out = np.zeros(4096, dtype=adsbcu_mod.MSG_DTYPE)
    got = adsbcu_mod.Demod1090(1, 1).feed_device(t.data_ptr(), iq.size // 2, out)
    assert oracle_mod.same_messages(want, got) == []
    with pytest.raises(adsbcu_mod.AdsbcuError):                      # misaligned device pointer
        adsbcu_mod.Demod1090().feed_device(t.data_ptr() + 2, 1000, out)


def test_reset_and_filter_state(adsbcu_mod, oracle_mod, golden):
    iq = golden["c3_dense"]["iq"]
    d = adsbcu_mod.Demod1090(0, 1)
    first = d.feed(iq).copy()
    d.reset(clear_filter=True)
    again = d.feed(iq)
    assert oracle_mod.same_messages(first, again) == []
    # without clearing the filter, Address/Parity replies of now-known aircraft are accepted from the start
    d.reset(clear_filter=False)
    warm = d.feed(iq)
    assert len(warm) >= len(first)
    o = oracle_mod.Oracle1090(0, 1)
    o.feed(iq)
    o2 = oracle_mod.Oracle1090(0, 1)
    for a in np.unique(first["addr"][np.isin(first["msgtype"], (11, 17))]):
        o2.filter_add(int(a))
    d4 = adsbcu_mod.Demod1090(0, 1)
    for a in np.unique(first["addr"][np.isin(first["msgtype"], (11, 17))]):
        d4.filter_add(int(a))
        assert d4.filter_test(int(a))
    assert oracle_mod.same_messages(o2.feed(iq), d4.feed(iq)) == []


def test_edge_inputs(adsbcu_mod, oracle_mod):
    d = adsbcu_mod.Demod1090(2, 1)
    assert len(d.feed(np.zeros(0, dtype=np.uint8))) == 0
    for fill in (0, 127, 255):                                     # silence, DC, saturation
        iq = np.full(2 * 300000, fill, dtype=np.uint8)
        assert len(d.feed(iq)) == len(oracle_mod.Oracle1090(2, 1).feed(iq)) == 0
    rng = np.random.default_rng(3)                                  # full-scale uniform noise: many preambles
    iq = rng.integers(0, 256, size=2 * 400000, dtype=np.uint8)
    want = oracle_mod.Oracle1090(2, 1).feed(iq)
    got = adsbcu_mod.Demod1090(2, 1).feed(iq)
    assert oracle_mod.same_messages(want, got) == []
    out = np.zeros(3, dtype=adsbcu_mod.MSG_DTYPE)                   # out[] too small -> overflow status
    from libadsb_b200 import synth
    with pytest.raises(adsbcu_mod.AdsbcuError, match="-4"):
        adsbcu_mod.Demod1090().feed(synth.config2(seconds=0.2, msgs_per_s=2000)[0], out=out)


def test_full_size_properties(adsbcu_mod, oracle_mod):
    """BASELINE config 2 at full size (60 s, 144 M samples): properties that need no oracle run,
    plus exact oracle parity on the first 5 s of the same stream."""
    from libadsb_b200 import synth
    iq, frames, starts = synth.config2(seconds=60.0)
    d = adsbcu_mod.Demod1090(1, 1, max_feed_samples=1 << 26)
    msgs = d.feed(iq, cap=200000)
    st = d.stats()
    assert st["samples"] == 144_000_000
    assert 40000 < len(msgs) <= 70000
    assert np.all(np.diff(msgs["timestamp"].astype(np.int64)) > 0)             # stream order
    # every accepted DF17/DF11 frame is CRC-consistent after correction
    for m in msgs[::257]:
        nb = int(m["msgbits"]) // 8
        syn = oracle_mod.crc(bytes(m["msg"][:nb]), nb * 8)
        if m["msgtype"] == 17:
            assert syn == 0
        elif m["msgtype"] == 11:
            assert syn & 0xFFFF80 == 0
    # injected frames come back verbatim (clean ones) at the right time
    sent = {bytes(f) for f in frames}
    clean = msgs[(msgs["correctedbits"] == 0) & np.isin(msgs["msgtype"], (11, 17))]
    hit = sum(bytes(m["msg"][:int(m["msgbits"]) // 8]) in sent for m in clean[::50])
    assert hit >= 0.98 * len(clean[::50])
    # the stream prefix decodes identically (demodulation is causal up to the 326-sample lookahead)
    n5 = 5 * 2_400_000
    want = oracle_mod.Oracle1090(1, 1).feed(iq[:2 * n5])
    t_end = 5 * (n5 - 326)
    head = msgs[msgs["timestamp"] < t_end]
    assert oracle_mod.same_messages(want[want["timestamp"] < t_end], head) == []


@pytest.mark.gpu
def test_filter_saturation(adsbcu_mod, oracle_mod):
    """More aircraft than the reference's filter has slots (4096, icao_filter.c:26): insertions are dropped once the table is full
    and membership then depends on the insertion order.  The engine has to follow the reference through the transition (its
    ordered pass falls back to the sequential replay for the batch that fills the table) and afterwards (static table)."""
    from libadsb_b200 import synth
    iq = synth.config3(seconds=7.0, seed=4096, msgs_per_s=3500.0, pool=20000)[0]
    want = oracle_mod.Oracle1090(1, 1).feed(iq, cap=1 << 18)
    d = adsbcu_mod.Demod1090(1, 1)
    got = d.feed(iq)
    assert oracle_mod.same_messages(want, got) == []
    _, addrs = d.state_export()
    assert len(addrs) == 4096                              # the table is full
    got2 = d.feed(iq)                                      # and stays as it is: a second pass over the same samples
    want2 = oracle_mod.Oracle1090(1, 1)
    want2.feed(iq, cap=1 << 18)
    assert oracle_mod.same_messages(want2.feed(iq, cap=1 << 18), got2) == []


@pytest.mark.gpu
def test_live_mode_filter_expiry(adsbcu_mod, oracle_mod):
    """Live SDR input: the reference flips its two filter tables every 60 s (icaoFilterExpire, icao_filter.c:112-127,
    dump1090.c:469-472), so an address not heard for two flips is forgotten.  adsbcu_1090_filter_expire is that flip; the
    flips are injected between feeds at the same stream positions on both sides.  Traffic: every aircraft announces itself
    (DF17/DF11) only in the first part of the capture and keeps sending Address/Parity replies afterwards, so the accepted
    set depends on the flips."""
    from libadsb_b200 import synth
    rng = np.random.default_rng(60)
    icaos = [int(x) for x in rng.integers(1, 1 << 24, size=40)]

    def part(dfs, m=300):
        n = int(0.12 * synth.FS_1090)
        frames = [synth.make_frame(int(rng.choice(dfs)), icaos[int(rng.integers(0, len(icaos)))], rng) for _ in range(m)]
        starts = (np.arange(m) + rng.random(m) * 0.2) * (n - 400) / m              # no overlaps
        return synth.synth_capture(n, frames, starts, np.full(m, 40.0), rng.uniform(0, 2 * np.pi, m), 2.0, rng)
    parts = [part([17, 11])] + [part([4, 5, 20, 21]) for _ in range(5)]
    o = oracle_mod.Oracle1090(1, 1)
    d = adsbcu_mod.Demod1090(1, 1)
    plain = adsbcu_mod.Demod1090(1, 1)
    n_diff = 0
    for k, p in enumerate(parts):
        if k in (1, 2, 4, 5):                       # the first flip comes at once in the reference (next_flip starts at 0)
            o.filter_expire()
            d.filter_expire()
        want = o.feed(p, cap=1 << 16)
        got = d.feed(p)
        assert oracle_mod.same_messages(want, got) == []
        n_diff += len(plain.feed(p)) != len(got)
    assert n_diff > 0                                # the flips changed what is accepted: the test can tell
    del rng


@pytest.mark.gpu
def test_overflow_recovery_full_scale_noise(adsbcu_mod, oracle_mod):
    """2^24 samples of full-scale uniform noise with --fix-2bit: several times more preambles and records per position than the
    lists of a chunk are sized for.  The engine has to notice the overflow, double its lists, run the chunk again and still
    return exactly the reference's set; the stream state must be usable afterwards."""
    rng = np.random.default_rng(2 ** 24)
    n = 1 << 24
    iq = rng.integers(0, 256, size=2 * n, dtype=np.uint8)
    d = adsbcu_mod.Demod1090(2, 1, max_feed_samples=n)
    o = oracle_mod.Oracle1090(2, 1)
    want = o.feed(iq, cap=1 << 20)
    got = d.feed(iq, cap=1 << 20)
    assert oracle_mod.same_messages(want, got) == []
    tail = iq[:2 * 400_000]
    assert oracle_mod.same_messages(o.feed(tail, cap=1 << 18), d.feed(tail, cap=1 << 18)) == []
