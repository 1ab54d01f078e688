"""GPU: a capture split over N shards (each with its own handle, as N GPUs would have) gives exactly the
message set of the one-shot stream and of the reference, for every N and window length."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def demod_emulated(adsbcu, shard, iq, world, window, nfix=2, fdf=1):
    """shard.demod_sharded with the N ranks run one after the other on one GPU."""
    n = iq.size // 2
    hs = [adsbcu.Demod1090(nfix, fdf, max_feed_samples=max(1 << 16, min(n, window))) for _ in range(world)]
    res = adsbcu.Demod1090(nfix, fdf, device=-1)
    msgs, base, revived = [], 0, 0
    while base < n:
        w = iq[2 * base:2 * min(n, base + window)]
        nw = w.size // 2
        tail = iq[2 * max(0, base - shard.HALO):2 * base]
        cands, known = [], np.zeros(adsbcu.KNOWN_WORDS, np.uint32)
        for r, d in enumerate(hs):
            lo, hi = shard.shard_range(nw, r, world)
            halo = w[2 * (lo - shard.HALO):2 * lo] if lo >= shard.HALO else np.concatenate([tail, w[:2 * lo]])[-2 * shard.HALO:]
            d.seek(base + lo, halo)
            cands.append(d.scan_candidates(w[2 * lo:2 * hi]))
            known |= d.known_export()
        for r, d in enumerate(hs):
            d.known_import(known)
            rv = d.revive()
            revived += len(rv)
            cands[r] = shard.merge_sorted(cands[r], rv)
        msgs.append(res.resolve(np.concatenate(cands)))
        base += window
    return np.concatenate(msgs), revived


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_equals_reference(adsbcu_mod, oracle_mod, golden, world):
    from libadsb_b200 import shard
    g = golden["c3_dense"]
    iq = g["iq"]
    for window in (iq.size // 2, 100_000):
        got, _ = demod_emulated(adsbcu_mod, shard, iq, world, window)
        assert oracle_mod.same_messages(g["ref_fix2_df1"], got) == []


def test_sharded_revives_address_parity_replies(adsbcu_mod, oracle_mod):
    """Dense traffic with many DF4/5/20/21 replies: their addresses are often announced in ANOTHER shard, so the
    revive step must fire, and the result must still equal the one-shot stream and the oracle."""
    from libadsb_b200 import shard, synth
    iq = synth.config3(seconds=1.0, seed=321, msgs_per_s=8000.0, pool=400)[0]
    want = oracle_mod.Oracle1090(1, 1).feed(iq)
    one = adsbcu_mod.Demod1090(1, 1).feed(iq)
    assert oracle_mod.same_messages(want, one) == []
    got, revived = demod_emulated(adsbcu_mod, shard, iq, 4, iq.size // 2, nfix=1)
    assert revived > 0
    assert oracle_mod.same_messages(want, got) == []
    got, _ = demod_emulated(adsbcu_mod, shard, iq, 8, 300_000, nfix=1)
    assert oracle_mod.same_messages(want, got) == []


def test_demod_sharded_local_group(adsbcu_mod, oracle_mod, golden):
    from libadsb_b200 import shard
    g = golden["c2_sparse"]
    d = adsbcu_mod.Demod1090(0, 1)
    got = shard.demod_sharded(d, shard.LocalGroup(), g["iq"], window_samples=120_000)
    assert oracle_mod.same_messages(g["ref_fix0_df1"], got) == []
