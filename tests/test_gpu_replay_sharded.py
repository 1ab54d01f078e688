"""GPU: adsbcu_1090_replay_sharded -- every shard demodulates AND resolves its own range on its own handle; the start
state of a range (skip state + ICAO filter contents) is found by the fixed-point exchange.  The ranks are threads
here (one handle each, as N GPUs would have); the union of their message lists must equal the reference's message
set for every N, with uneven ranges and ranges made of several pieces."""
import os
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EMU = os.environ.get("ADSBCU_TEST_EMU") == "1"


class ThreadGroup:
    """all-gather between the rank threads of one process."""

    def __init__(self, world):
        self.world, self.slots, self.bar = world, [None] * world, threading.Barrier(world)

    def allgather(self, rank, words):
        self.slots[rank] = words
        self.bar.wait()
        parts = list(self.slots)
        self.bar.wait()
        return parts


def device_copy(iq):
    if EMU:
        return iq, iq.ctypes.data
    import torch
    t = torch.from_numpy(iq).cuda()
    return t, t.data_ptr()


def replay(adsbcu, iq, cuts, nfix, fdf, pieces_per_rank=1, tail=False):
    """cuts: sample indices [0, c1, ..., n] -> len(cuts)-1 ranks; tail: the last range is rank 0's second one (len(cuts)-2 ranks)."""
    world = len(cuts) - 1 - (1 if tail else 0)
    keep, base = device_copy(iq)
    group = ThreadGroup(world)
    results, errors = [None] * world, []

    def body(rank):
        try:
            lo, hi = cuts[rank], cuts[rank + 1]
            d = adsbcu.Demod1090(nfix, fdf, max_feed_samples=1 << 20)
            edges = [lo + (hi - lo) * k // pieces_per_rank // 8 * 8 for k in range(pieces_per_rank)] + [hi]
            pieces = [(base + 2 * a, b - a) for a, b in zip(edges[:-1], edges[1:])]
            out = np.zeros(1 << 16, adsbcu.MSG_DTYPE)
            t = None
            if tail and rank == 0:
                tlo, thi = cuts[-2], cuts[-1]
                t = ([(base + 2 * tlo, thi - tlo)], tlo, iq[2 * max(0, tlo - 326):2 * tlo])
            msgs, rep = d.replay_sharded(rank, world, pieces, lo, iq[2 * max(0, lo - 326):2 * lo],
                                         lambda w: group.allgather(rank, w), out, has_tail=tail, tail=t)
            results[rank] = (msgs.copy(), rep.passes, rep.rounds, rep.start_addrs, d.n_first)
        except Exception as e:      # noqa: BLE001
            errors.append(e)
            group.bar.abort()

    threads = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    if errors:
        raise errors[0]
    del keep
    return results


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_replay_sharded_equals_reference(adsbcu_mod, oracle_mod, golden, world):
    g = golden["c3_dense"]
    iq = g["iq"]
    n = iq.size // 2
    cuts = [0] + [(n * (k + 1) // (world + 1)) // 8 * 8 for k in range(1, world)] + [n]     # rank 0 twice as long
    res = replay(adsbcu_mod, iq, cuts, 2, 1, pieces_per_rank=2 if world == 3 else 1)
    got = np.concatenate([r[0] for r in res])
    assert oracle_mod.same_messages(g["ref_fix2_df1"], got) == []
    assert res[0][1] == 1                                  # rank 0 never runs twice
    assert all(r[2] == res[0][2] for r in res)             # same number of rounds everywhere
    if world > 1:
        assert max(r[1] for r in res) >= 2 and res[-1][3] > 0      # later ranks re-ran with a filled filter


def test_replay_sharded_cut_inside_messages(adsbcu_mod, oracle_mod):
    """Dense traffic, cuts every 60 000 samples: many accepted messages straddle a cut, so the skip state has to travel."""
    from libadsb_b200 import synth
    iq = synth.config3(seconds=0.3, seed=77, msgs_per_s=6000.0, pool=100)[0]
    n = iq.size // 2
    want = oracle_mod.Oracle1090(1, 1).feed(iq)
    cuts = list(range(0, n, 60000)) + [n]
    res = replay(adsbcu_mod, iq, cuts, 1, 1)
    got = np.concatenate([r[0] for r in res])
    assert len(want) > 200
    assert oracle_mod.same_messages(want, got) == []


@pytest.mark.parametrize("world", [2, 4])
def test_replay_sharded_with_tail_on_rank0(adsbcu_mod, oracle_mod, golden, world):
    """world + 1 ranges: rank 0 demodulates the first range in round 0 and the last one in round 1 (nobody waits for the
    last range's end state), the others make their two passes.  Same message set."""
    g = golden["c3_dense"]
    iq = g["iq"]
    n = iq.size // 2
    cuts = [0] + [(n * k // (world + 1)) // 8 * 8 for k in range(1, world + 1)] + [n]
    res = replay(adsbcu_mod, iq, cuts, 2, 1, tail=True)
    first = res[0][4]
    got = np.concatenate([res[0][0][:first]] + [r[0] for r in res[1:]] + [res[0][0][first:]])
    assert oracle_mod.same_messages(g["ref_fix2_df1"], got) == []
    assert res[0][1] == 2 and first > 0 and len(res[0][0]) > first     # one pass over each of its two ranges
    assert all(r[2] == res[0][2] for r in res)


def test_replay_sharded_refuses_a_saturating_filter(adsbcu_mod):
    """Once the reference's 4096-slot filter is (nearly) full its contents depend on the insertion order, so a sharded replay could
    not be exact: the orchestrator must say so instead of returning a different message set."""
    from libadsb_b200 import synth
    iq = synth.config3(seconds=7.0, seed=4096, msgs_per_s=3500.0, pool=20000)[0]
    n = iq.size // 2
    with pytest.raises(adsbcu_mod.AdsbcuError, match="ADSBCU_SHARD_MAX_ADDRS"):
        replay(adsbcu_mod, iq, [0, n // 2 // 8 * 8, n], 1, 1)
