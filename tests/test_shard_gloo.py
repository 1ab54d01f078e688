"""Host side of the sharded-capture path on CPU: world_size 2 over gloo.

No GPU here, so the candidates are synthesised from the reference-generated golden records (one
candidate per accepted message); what is exercised is everything around the kernels: range
arithmetic, the known-address exchange, the candidate gather in rank order, and the ordered pass
(adsbcu_1090_resolve on a merge-only handle), whose scores depend on the ICAO filter state."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NT, NONE = 1 << 25, 0xFFFFFFFF


def cands_from_records(recs):
    from libadsb_b200 import adsbcu
    c = np.zeros(len(recs), dtype=adsbcu.CAND_DTYPE)
    ts = recs["timestamp"].astype(np.int64)
    c["pos"] = (ts - 772) // 5
    c["phase"] = ts - 768 - 5 * c["pos"].astype(np.int64)
    c["corrections"], c["msgtype"] = recs["correctedbits"], recs["msgtype"]
    c["verbatim"], c["msg"], c["crc"], c["addr"] = recs["verbatim"], recs["msg"], recs["crc"], recs["addr"]
    c["signal_level"] = recs["signal_level"]
    for i, r in enumerate(recs):
        df, nc, addr, crc = int(r["msgtype"]), int(r["correctedbits"]), int(r["addr"]), int(r["crc"])
        kind, test, add = 0, crc, NONE                       # K_AP
        if df == 17:
            kind, test = 5 + nc, addr
            add = addr if nc == 0 else NONE
        elif df == 18:
            kind, test = 8 + nc, (addr & 0xFFFFFF) | NT
            add = (addr | NT) if nc == 0 else NONE
        elif df == 11:
            acq = (crc & 0x7F) == 0
            kind, test = (1 if acq else 2) + 2 * nc, addr
            add = addr if (acq and nc == 0) else NONE
        c["kind"][i], c["test_key"][i], c["add_key"][i] = kind, test, add
    return c


def same(a, b):
    for f in ("timestamp", "msg", "verbatim", "msgbits", "correctedbits", "msgtype", "score", "crc", "addr"):
        assert np.array_equal(a[f], b[f]), f
    assert np.allclose(a["signal_level"], b["signal_level"], rtol=1e-12)


def test_shard_ranges_partition():
    from libadsb_b200.shard import shard_range
    for n in (0, 1, 7, 8, 1000, 144_000_000, 207_360_000_000):
        for world in (1, 2, 3, 4, 8):
            cuts = [shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[r][1] == cuts[r + 1][0] for r in range(world - 1))
            assert all(lo % 8 == 0 and lo <= hi for lo, hi in cuts)


def test_resolve_replays_reference_scores(golden):
    """One candidate per reference message, in order: the ordered pass must accept every one of them with the
    reference's score (which depends on the filter state built by the earlier messages) and timestamp."""
    from libadsb_b200 import adsbcu
    for name in ("c2_sparse", "c3_dense"):
        ref = golden[name]["ref_fix2_df1"]
        got = adsbcu.Demod1090(2, 1, device=-1).resolve(cands_from_records(ref))
        assert len(got) == len(ref)
        same(ref, got)


def test_merge_only_handle_refuses_gpu_work():
    from libadsb_b200 import adsbcu
    d = adsbcu.Demod1090(device=-1)
    with pytest.raises(adsbcu.AdsbcuError, match="merge-only"):
        d.feed(np.zeros(64, np.uint8))
    with pytest.raises(adsbcu.AdsbcuError, match="merge-only"):
        d.scan_candidates(np.zeros(64, np.uint8))


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from libadsb_b200 import adsbcu, shard
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        with np.load(os.path.join(ROOT, "tests", "golden", "c3_dense.npz")) as z:
            ref, n = z["ref_fix2_df1"], z["iq"].size // 2
        g = shard.TorchGroup()
        allc = cands_from_records(ref)
        lo, hi = shard.shard_range(n, rank, world)
        mine = allc[(allc["pos"] >= lo) & (allc["pos"] < hi)]
        # the known-address exchange: every rank contributes disjoint words, all see the OR
        w = np.zeros(adsbcu.KNOWN_WORDS, np.uint32)
        w[rank::world] = 1 << rank
        merged = g.or_words(w)
        assert all(np.all(merged[r::world] == (1 << r)) for r in range(world))
        parts = g.gather_to_root(mine)
        if rank == 0:
            assert [len(p) for p in parts] == [int(((allc["pos"] >= a) & (allc["pos"] < b)).sum())
                                               for a, b in (shard.shard_range(n, r, world) for r in range(world))]
            got = adsbcu.Demod1090(2, 1, device=-1).resolve(np.concatenate(parts))
            same(ref, got)
            ret.put(("ok", len(got)))
        else:
            assert parts is None
    except Exception as e:      # surface the failure in the parent
        ret.put(("fail", f"rank {rank}: {e!r}"))
        raise
    finally:
        dist.destroy_process_group()


def test_two_rank_gather_and_resolve_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    status, n = ret.get(timeout=5)
    assert status == "ok" and n > 100
