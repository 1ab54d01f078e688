"""One capture, several GPUs (BASELINE config 4): contiguous sample ranges, no data-path
collective.  Rank r of N scans start positions [r*n/N, (r+1)*n/N) of every window state-free on its own
GPU (reading a 326-sample left halo for the overlap magnitudes, fifo.c:187-196); the only exchanges are
host-side and tiny: the OR of the "addresses that may enter the ICAO filter" sets (4 MiB per rank) and
the gather of candidate records to rank 0, which replays the sequential rules of demodulate2400
(demod_2400.c:336,349-355,410-443; mode_s.c:781-788) once.  The message set is independent of N.

Plumbing only: all work happens behind the C ABI (adsbcu_1090_seek / _scan_candidates / _known_export /
_known_import / _revive / _resolve)."""
from __future__ import annotations

import os

import numpy as np

from . import adsbcu

HALO = 326   # dump1090.c:159 trailing samples == left halo of a shard


def shard_range(n: int, rank: int, world: int, align: int = 8) -> tuple[int, int]:
    """Start positions owned by `rank`: contiguous, disjoint, covering [0, n); interior cuts are multiples
    of `align` samples (16-byte alignment of device pointers)."""
    cut = lambda k: n if k >= world else (n * k // world) // align * align   # noqa: E731
    return cut(rank), cut(rank + 1)


def merge_sorted(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Two candidate arrays, each ascending in (pos, phase) -> one."""
    if len(b) == 0:
        return a
    if len(a) == 0:
        return b
    c = np.concatenate([a, b])
    return c[np.lexsort((c["phase"], c["pos"]))]


class LocalGroup:
    """world_size 1 stand-in for torch.distributed (also used to emulate N shards on one GPU)."""
    rank, world = 0, 1

    def or_words(self, w: np.ndarray) -> np.ndarray:
        return w

    def gather_to_root(self, c: np.ndarray) -> list[np.ndarray] | None:
        return [c]


class TorchGroup:
    """torch.distributed plumbing (gloo or nccl): all_gather of the known-address words, gather of candidates."""

    def __init__(self, device=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.device = device if device is not None else torch.device("cpu")

    def or_words(self, w: np.ndarray) -> np.ndarray:
        t = self.torch.from_numpy(w.view(np.int32)).to(self.device)
        parts = [self.torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(parts, t)
        acc = parts[0]
        for p in parts[1:]:
            acc = acc | p
        return acc.cpu().numpy().view(np.uint32)

    def gather_to_root(self, c: np.ndarray) -> list[np.ndarray] | None:
        t = self.torch
        n = t.tensor([len(c)], dtype=t.int64, device=self.device)
        sizes = [t.zeros_like(n) for _ in range(self.world)]
        self.dist.all_gather(sizes, n)
        sizes = [int(s) for s in sizes]
        m = max(max(sizes), 1)
        buf = np.zeros(m, dtype=adsbcu.CAND_DTYPE)
        buf[:len(c)] = c
        mine = t.from_numpy(buf.view(np.uint8).reshape(-1)).to(self.device)
        parts = [t.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(parts, mine)      # all_gather: nccl has no gather-to-host; candidates are a few MB
        if self.rank != 0:
            return None
        return [p.cpu().numpy().view(adsbcu.CAND_DTYPE)[:k].copy() for p, k in zip(parts, sizes)]


def scan_window(d: adsbcu.Demod1090, group, iq_window: np.ndarray, window_base: int, prev_tail: np.ndarray | None):
    """This rank's part of one window.  iq_window: the whole window's interleaved uint8 I/Q (host; a rank only
    touches its own range plus halo).  window_base: capture index of the window's first sample.  prev_tail: the
    (up to) 326 samples before the window (None at capture start).  Returns this rank's candidates."""
    n = iq_window.size // 2
    lo, hi = shard_range(n, group.rank, group.world)
    if lo >= HALO:
        halo = iq_window[2 * (lo - HALO):2 * lo]
    else:
        before = prev_tail if prev_tail is not None else np.zeros(0, np.uint8)
        halo = np.concatenate([before, iq_window[:2 * lo]])[-2 * HALO:]
    d.seek(window_base + lo, halo)
    cands = d.scan_candidates(iq_window[2 * lo:2 * hi])
    d.known_import(group.or_words(d.known_export()))
    return merge_sorted(cands, d.revive())


def demod_sharded(d: adsbcu.Demod1090, group, iq: np.ndarray, resolver: adsbcu.Demod1090 | None = None,
                  window_samples: int | None = None) -> np.ndarray | None:
    """Demodulate one capture with group.world shards per window.  `d`: this rank's GPU handle; `resolver`:
    the handle holding the sequential state (rank 0 only; may be merge-only, device=-1).  Returns the
    messages on rank 0, None elsewhere."""
    iq = np.ascontiguousarray(iq, dtype=np.uint8)
    n = iq.size // 2
    window = window_samples or n
    msgs, base = [], 0
    while base < n or (n == 0 and base == 0):
        w = iq[2 * base:2 * min(n, base + window)]
        tail = iq[2 * max(0, base - HALO):2 * base] if base else None
        mine = scan_window(d, group, w, base, tail)
        if os.environ.get("ADSBCU_SHARD_DEBUG"):
            print(f"rank {group.rank}: window @{base}: {len(mine)} candidates, pos {mine['pos'].min() if len(mine) else '-'}..{mine['pos'].max() if len(mine) else '-'}", flush=True)
        parts = group.gather_to_root(mine)
        if os.environ.get("ADSBCU_SHARD_DEBUG") and parts is not None:
            print("root got", [(len(p), int(p['pos'].min()) if len(p) else -1) for p in parts], flush=True)
        if parts is not None:
            allc = np.concatenate(parts) if parts else mine     # rank order == position order
            msgs.append((resolver or d).resolve(allc))
        base += max(window, 1)
        if n == 0:
            break
    if group.rank != 0:
        return None
    return np.concatenate(msgs) if msgs else np.zeros(0, adsbcu.MSG_DTYPE)


def _selftest():
    """torchrun --nproc-per-node N -m libadsb_b200.shard : one capture over N GPUs == the reference's message set
    (tests/golden/c3_dense.npz was produced by the reference binary) and == the single-GPU stream."""
    import os
    import torch
    import torch.distributed as dist
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    group = TorchGroup(device=torch.device("cuda", local))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with np.load(os.path.join(root, "tests", "golden", "c3_dense.npz")) as z:
        iq, ref = z["iq"], z["ref_fix2_df1"]
    d = adsbcu.Demod1090(2, 1, device=local)
    resolver = adsbcu.Demod1090(2, 1, device=-1) if rank == 0 else None
    ok = True
    for window in (iq.size // 2, 100_000):
        if resolver is not None:
            resolver.reset(True)
        d.reset(True)
        got = demod_sharded(d, group, iq, resolver=resolver, window_samples=window)
        if rank == 0:
            same = len(got) == len(ref) and all(np.array_equal(got[f], ref[f]) for f in
                                                ("timestamp", "msgbits", "correctedbits", "score", "crc", "addr"))
            if same:    # bytes past the frame length are scratch in the reference (demod_2400.c:103,345)
                live = np.arange(14)[None, :] < (ref["msgbits"] // 8)[:, None]
                same = bool(np.all((got["msg"] == ref["msg"]) | ~live) and np.all((got["verbatim"] == ref["verbatim"]) | ~live))
            print(f"shard selftest: world={group.world} window={window}: {len(got)} messages, "
                  f"{'bit-identical to the reference' if same else 'MISMATCH'}")
            ok = ok and same
    dist.destroy_process_group()
    if not ok:
        raise SystemExit(1)


if __name__ == "__main__":
    _selftest()
