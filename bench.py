#!/usr/bin/env python
"""bench.py -- Msamples/s through the 1090 MHz demod+CRC hot path (BASELINE.json metric).

A step is one pass of the whole path (magnitude -> slicer decisions -> preamble -> CRC/repair -> order -> the ordered
pass of demodulate2400 -> signal level -> messages; every stage on the GPU) over one synthetic capture.

  --workload config2 (default: 60 s, 2.4 MS/s, configs[1]) | config3 (1 h dense, --fix-2bit) | config4 (24 h, strong
             scaling) | config5 (UAT 978)
  N > 1      ONE capture cut into contiguous sample ranges (adsbcu_1090_replay_sharded_tail), no data-path collective
  value      whole-job Msamples/s, I/Q already resident in HBM
  e2e        the same metric with HOST buffers (H2D inside the timed region, messages written into pinned out[])
  roofline   per kernel: 2 algorithmic bytes per complex sample / its CUDA-event duration, against MEASURED_PEAKS hbm_gbs
  cpu_baseline / --impl reference   the reference's own dump1090 (oracle/_ref) on the host cores
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Msamples/s demod+CRC"
UNIT = "Msamples/s"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi sampled during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ref_run(iq: np.ndarray, nproc: int, flags: list[str]) -> float:
    """Wall seconds for `nproc` reference dump1090 processes, each demodulating one equal
    contiguous shard of `iq` from tmpfs (process start + table init + file read included)."""
    from oracle import oracle1090 as O
    n = iq.size // 2
    shard = (n + nproc - 1) // nproc
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as d:
        files = []
        for k in range(nproc):
            p = os.path.join(d, f"s{k}.uc8")
            iq[2 * k * shard: 2 * min(n, (k + 1) * shard)].tofile(p)
            files.append(p)
        t0 = time.perf_counter()
        procs = [subprocess.Popen([O.REF_BIN, "--ifile", f, "--quiet", *flags], stdout=subprocess.DEVNULL,
                                  stderr=subprocess.DEVNULL) for f in files]
        for p in procs:
            p.wait()
        return time.perf_counter() - t0


def oracle_run(iq: np.ndarray, nfix: int) -> float:
    from oracle import oracle1090 as O
    o = O.Oracle1090(nfix, 1)
    t0 = time.perf_counter()
    o.feed(iq, cap=1 << 20)
    return time.perf_counter() - t0


def cpu_baseline(iq: np.ndarray, nfix: int, cores: int, sample_desc: str):
    from oracle import oracle1090 as O
    flags = {0: [], 1: ["--fix"], 2: ["--fix-2bit"]}[nfix]
    n = iq.size // 2
    if O.have_ref():
        dt = ref_run(iq, cores, flags)
        kind = "reference"
    else:
        dt = oracle_run(iq, nfix)
        kind, cores = "port", 1
    return {"value": n / dt / 1e6, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_desc}, dt


def message_set_hash(msgs: np.ndarray) -> tuple[int, int]:
    """Order- and partition-independent 64-bit digest of a message list (sum of per-message mixes mod 2^64)."""
    if len(msgs) == 0:
        return 0, 0
    raw = np.ascontiguousarray(msgs).view(np.uint8).reshape(len(msgs), -1)
    pad = (-raw.shape[1]) % 8
    if pad:
        raw = np.concatenate([raw, np.zeros((len(msgs), pad), np.uint8)], axis=1)
    w = raw.view(np.uint64)
    acc = np.full(len(msgs), 0x9E3779B97F4A7C15, np.uint64)
    with np.errstate(over="ignore"):
        for k in range(w.shape[1]):
            acc = (acc ^ w[:, k]) * np.uint64(0xFF51AFD7ED558CCD)
            acc ^= acc >> np.uint64(33)
        return len(msgs), int(acc.sum(dtype=np.uint64))


def pieces_of(base_ptr: int, tile_samples: int, lo: int, hi: int):
    """[lo, hi) of the capture `tile repeated` as (address, nsamples) pieces, cut at tile boundaries."""
    out, pos = [], lo
    while pos < hi:
        off = pos % tile_samples
        take = min(tile_samples - off, hi - pos)
        out.append((base_ptr + 2 * off, take))
        pos += take
    return out


def main_978(args):
    """config5: UAT 978 MHz, 2.083334 MS/s, through UAT978Handler::HandleData's replacement (adsbcu_978_feed).
    A step = one 60 s capture.  value: samples resident in HBM (adsbcu_978_feed_device, one call); e2e: host buffers fed in the
    reference's 262144-byte HandleData buffers (RTLSDR.hpp:55).  cpu_baseline: the oracle port (dump978 itself is not in the
    reference tree, so there is no reference binary for this path) on a bounded sample."""
    import torch
    from libadsb_b200 import adsbcu, build, synth978
    build.build()
    metric, unit = "Msamples/s UAT demod+RS", "Msamples/s"
    steps = args.steps if args.steps is not None else 5
    iq = synth978.config5(seconds=args.seconds)[0]
    n = iq.size // 2
    config = {"workload": f"config5: synthetic 2.083334 MS/s uint8 I/Q, {args.seconds:g} s, 200 ADS-B frames/s (80 % long) + 1 uplink frame/s, "
                          "0..t corrupted bytes per RS block, sigma=2 LSB", "samples_per_gpu": n,
              "l2_policy": f"input ({2 * n / 1e6:.0f} MB) larger than L2"}
    if args.impl == "reference":
        from oracle import oracle978 as O
        sample = iq[:2 * min(n, int(10 * synth978.FS_978))]
        times = []
        for _ in range(max(1, args.warmup) + steps):
            o = O.Oracle978()
            t0 = time.perf_counter()
            o.handle_data(sample, cap=1 << 16)
            times.append(time.perf_counter() - t0)
        dt = float(np.mean(times[max(1, args.warmup):]))
        v = sample.size / 2 / dt / 1e6
        cb = {"value": v, "unit": unit, "cores": 1, "kind": "port",
              "sample": "first 10 s of the capture, one call (oracle/oracle978.c; dump978 is an absent submodule of the reference)"}
        print(json.dumps({"impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": args.gpus, "steps": steps,
                          "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "u16/i16/gf256", "data": "synthetic", "config": config, "cpu_baseline": cb,
                          "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    torch.cuda.set_device(0)
    host = torch.from_numpy(iq).pin_memory()
    dev = host.cuda()
    hview = host.numpy()
    d = adsbcu.Demod1090()
    out = np.zeros(1 << 16, adsbcu.UAT_DTYPE)
    BUF = 262144

    def step_resident():
        d.reset_uat()
        f = d.handle_data_uat_device(dev.data_ptr(), iq.size, out)
        return f, d.uat_stats()

    def step_e2e():
        d.reset_uat()
        return sum(len(d.handle_data_uat(hview[k:k + BUF], out=out)) for k in range(0, iq.size, BUF))

    sampler = ClockSampler(0)
    sampler.start()
    for _ in range(max(3, args.warmup)):
        frames, st = step_resident()
    nframes = len(frames)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dev_ms, launches, calls = [], 0, 0
    for _ in range(steps):
        _, st = step_resident()
        dev_ms.append(st["ms_device"]); launches += st["launches"]; calls = st["calls"]
    torch.cuda.synchronize()
    t_res = time.perf_counter() - t0
    step_e2e()
    t0 = time.perf_counter()
    for _ in range(steps):
        ne = step_e2e()
    t_e2e = time.perf_counter() - t0
    clocks = sampler.stop()
    peak, peak_src = load_peaks()
    k_ms = float(np.mean(dev_ms))
    achieved = 2.0 * n / (k_ms * 1e-3) / 1e9
    line = {"metric": metric, "value": n * steps / t_res / 1e6, "unit": unit, "n_gpus": 1, "steps": steps, "warmup": max(3, args.warmup),
            "ms_per_step": t_res / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16/i16/gf256",
            "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": n * steps / t_e2e / 1e6, "unit": unit, "h2d_bytes_per_step": 2 * n, "d2h_bytes_per_step": int(ne) * adsbcu.UAT_DTYPE.itemsize,
                    "note": "262144-byte HandleData buffers from pageable host memory, as the reference's RTL-SDR callback delivers them"},
            "roofline": {"bound": "hbm", "kernel": "all kernels of the feed (phase table gather + per process_buffer() call: sync search, frame decode, staging shift)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                         "launches_per_step": launches // steps, "algorithmic_bytes_per_launch": 2.0 * n / max(launches // steps, 1),
                         "avg_launch_ms": k_ms / max(launches // steps, 1), "traffic": None, "kernel_ms_per_step": k_ms,
                         "note": "device time includes the host round trip between process_buffer() calls: the reference's staging is a serial "
                                 "state machine (what a call presents depends on what the previous one consumed), so the calls run one after another"},
            "process_buffer_calls_per_step": calls, "frames_per_step": nframes,
            "realtime_factor": n * steps / t_res / synth978.FS_978}
    if not args.no_cpu_baseline:
        from oracle import oracle978 as O
        sample = iq[:2 * min(n, int(10 * synth978.FS_978))]
        o = O.Oracle978()
        t0 = time.perf_counter()
        o.handle_data(sample, cap=1 << 16)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": sample.size / 2 / dt / 1e6, "unit": unit, "cores": 1, "kind": "port",
                                "sample": "first 10 s of the capture, one call (oracle/oracle978.c; dump978 is an absent submodule of the reference)"}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seconds", type=float, default=60.0, help="length of the synthetic tile (60 s)")
    ap.add_argument("--hours", type=float, default=None, help="capture length: the tile repeated (config3: 1, config4: 24)")
    ap.add_argument("--nfix", type=int, default=None, help="0 none, 1 --fix, 2 --fix-2bit (default: 1; config3: 2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="config2", choices=["config2", "config3", "config4", "config5"],
                    help="config2 (default, the metric's configuration), config3 (1 h dense traffic, --fix-2bit), "
                         "config4 (24 h sharded over the GPUs, strong scaling), config5 (UAT 978)")
    args = ap.parse_args()
    if args.workload == "config5":
        return main_978(args)
    if args.steps is None:
        args.steps = 100 if args.workload == "config2" else 3
    if args.nfix is None:
        args.nfix = 2 if args.workload == "config3" else 1
    if args.hours is None:
        args.hours = {"config2": 0.0, "config3": 1.0, "config4": 24.0}[args.workload]
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from libadsb_b200 import synth
    tile_n = int(round(args.seconds * synth.FS_1090)) // 8 * 8
    ntiles = max(1, int(round(args.hours * 3600.0 / args.seconds)))      # tiles per capture (strong) or per GPU (weak)
    strong = args.workload == "config4"
    if args.workload == "config3":
        traffic = "Poisson 2000 msg/s with overlaps, DF17/11/4/5/20/21/18 mix, 1-2 bit errors"
        make = lambda seed: synth.config3(seconds=args.seconds, seed=seed)[0]                   # noqa: E731
    else:
        traffic = "1000 msg/s DF17/DF11, 6-30 dB SNR"
        make = lambda seed: synth.config2(seconds=args.seconds, seed=seed)[0]                   # noqa: E731
    total_tiles = ntiles if strong else ntiles * world
    total = tile_n * total_tiles
    sharded = world > 1 or ntiles > 1
    workload = (f"{args.workload}: synthetic 2.4 MS/s uint8 I/Q, {traffic}, sigma=2 LSB, nfix_crc={args.nfix}; "
                + (f"ONE capture of {total_tiles} x {args.seconds:g} s ({total / synth.FS_1090 / 3600:.3g} h; the {args.seconds:g} s tile repeated), "
                   f"contiguous ranges over {world} GPU(s)" if sharded else f"{args.seconds:g} s capture on one GPU"))
    config = {"workload": workload, "samples_per_gpu": total // world, "capture_samples": total,
              "l2_policy": f"input tile ({2 * tile_n / 1e6:.0f} MB) larger than L2",
              "sharding": ("none" if world == 1 else
                           "adsbcu_1090_replay_sharded_tail: world + 1 equal contiguous sample ranges of one capture, range r on rank r, the last one on "
                           "rank 0 too (one pass over each of its two, the other ranks two passes over theirs); start states (skip state + ICAO "
                           "filter) by fixed-point exchange of <= 16 KiB per rank and round; no data-path collective")}
    scaling = "strong" if strong else "weak"

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        iq = make(0x10900002)
        cores = os.cpu_count() or 1
        for _ in range(args.warmup):
            cpu_baseline(iq, args.nfix, cores, "")
        times = []
        desc = (f"one {args.seconds:g} s tile of the workload per step, split into {cores} contiguous shards, one dump1090 process each")
        cb = None
        for _ in range(args.steps):
            cb, dt = cpu_baseline(iq, args.nfix, cores, desc)
            times.append(dt)
        dt = float(np.mean(times))
        v = (iq.size // 2) / dt / 1e6
        cb["value"] = v
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                          "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u16/i32",
                          "data": "synthetic", "config": config, "cpu_baseline": cb,
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    # ------------------------------------------------------------------ our arm
    import torch
    import torch.distributed as dist
    from libadsb_b200 import adsbcu, build
    if rank == 0 or world == 1:
        build.build()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    W1 = adsbcu.SHARD_WORDS + 1
    if world > 1:                           # exchange buffers: pinned staging + device tensors (NCCL all-gather: ~50 us per round)
        ag_send_h = torch.zeros(W1, dtype=torch.int32).pin_memory()
        ag_recv_h = torch.zeros(world * W1, dtype=torch.int32).pin_memory()
        ag_send_d = torch.zeros(W1, dtype=torch.int32, device="cuda")
        ag_recv_d = torch.zeros(world * W1, dtype=torch.int32, device="cuda")

    def allgather(words: np.ndarray):
        ag_send_h[0] = int(words.size)
        ag_send_h[1:1 + words.size] = torch.from_numpy(words.view(np.int32))
        ag_send_d.copy_(ag_send_h, non_blocking=True)
        dist.all_gather_into_tensor(ag_recv_d, ag_send_d)
        ag_recv_h.copy_(ag_recv_d, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        flat = ag_recv_h.numpy().view(np.uint32)
        return [flat[q * W1 + 1:q * W1 + 1 + int(flat[q * W1])] for q in range(world)]

    iq = make(0x10900002)                   # the tile: the same on every rank (one capture)
    n = tile_n
    host = torch.from_numpy(iq[:2 * n]).pin_memory()
    dev = host.cuda(non_blocking=False)
    hview = host.numpy()
    # this rank's range(s) of the capture: world + 1 equal ranges, range r on rank r and the last one (the "tail") on rank 0
    # as well -- the first range never runs twice and nobody waits for the last one's end state (adsbcu_1090_replay_sharded_tail)
    def halo_of(pos):
        o = pos % n
        return None if pos == 0 else (hview[2 * (o - 326):2 * o] if o >= 326 else np.concatenate([hview[2 * (n - (326 - o)):], hview[:2 * o]]))
    has_tail = world > 1
    if world == 1:
        lo, hi = 0, total
        tlo = thi = total
    else:
        cut = lambda k: total if k >= world + 1 else (total * k // (world + 1)) // 8 * 8      # noqa: E731
        lo, hi = cut(rank), cut(rank + 1)
        tlo, thi = (cut(world), total) if rank == 0 else (total, total)
    my_n = (hi - lo) + (thi - tlo)
    halo = halo_of(lo)
    dev_pieces = pieces_of(dev.data_ptr(), n, lo, hi)
    host_pieces = pieces_of(hview.ctypes.data, n, lo, hi)
    dev_tail = (pieces_of(dev.data_ptr(), n, tlo, thi), tlo, halo_of(tlo)) if thi > tlo else None
    # a range that crosses tile boundaries would be fed as several pieces (one feed, with its pipeline fill and drain, per piece):
    # short ranges are laid out contiguously in HBM once, before anything is timed, as a real capture would be
    keep_alive = []

    def contiguous(pieces):
        if len(pieces) <= 1 or sum(c for _, c in pieces) > (1 << 30):
            return pieces
        buf = torch.empty(2 * sum(c for _, c in pieces), dtype=torch.uint8, device="cuda")
        pos = 0
        for p, c in pieces:
            o = p - dev.data_ptr()
            buf[pos:pos + 2 * c].copy_(dev[o:o + 2 * c])
            pos += 2 * c
        keep_alive.append(buf)
        return [(buf.data_ptr(), pos // 2)]
    if sharded:
        dev_pieces = contiguous(dev_pieces)
        if dev_tail:
            dev_tail = (contiguous(dev_tail[0]), dev_tail[1], dev_tail[2])
    host_tail = (pieces_of(hview.ctypes.data, n, tlo, thi), tlo, halo_of(tlo)) if thi > tlo else None
    d = adsbcu.Demod1090(nfix_crc=args.nfix, fix_df=1, max_feed_samples=max(n, 1 << 20), device=local_rank)
    # out[] in pinned memory: the engine's level kernel writes the accepted messages there itself (no host copy)
    probe = adsbcu.Demod1090(nfix_crc=args.nfix, fix_df=1, max_feed_samples=1 << 24, device=local_rank)
    per_sample = (len(probe.feed_device(dev.data_ptr(), min(n, 1 << 24), np.zeros(1 << 18, adsbcu.MSG_DTYPE))) + 64) / min(n, 1 << 24)
    probe.close()
    out_cap = max(4096, int(max(my_n, total // world + 8) * per_sample * 1.5))       # sized from the message density of the tile's first 7 s
    out_pinned = torch.zeros(out_cap * adsbcu.MSG_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    out = out_pinned.numpy().view(adsbcu.MSG_DTYPE)
    KEYS = ("front", "messages", "order", "resolve")

    def step_resident():
        """-> messages, {ms per kernel}, launches, chunks, passes, rounds"""
        if not sharded:
            d.reset(True)
            m = d.feed_device(dev.data_ptr(), n, out)
            st = d.stats()
            return m, {k: st["ms_" + k] for k in KEYS}, st["launches"] + 1, st["chunks"], 1, 0
        m, rep = d.replay_sharded(rank, world, dev_pieces, lo, halo, allgather, out, has_tail=has_tail, tail=dev_tail)
        return m, {k: getattr(rep, "ms_" + k) for k in KEYS}, rep.launches, rep.chunks, rep.passes, rep.rounds

    # sharded e2e: the range goes host -> device once (pinned pieces, inside the timed region) and both passes read it from
    # HBM; a range that does not fit (24 h on one GPU) is uploaded chunk by chunk on every pass instead
    # With host buffers the upload is the bottleneck, and the tail would make rank 0 upload twice what the others do: the
    # end-to-end arm cuts the capture into `world` equal ranges instead (adsbcu_1090_replay_sharded without a tail).
    if world == 1:
        elo, ehi = 0, total
    else:
        ecut = lambda k: total if k >= world else (total * k // world) // 8 * 8      # noqa: E731
        elo, ehi = ecut(rank), ecut(rank + 1)
    e_n = ehi - elo
    e_halo = halo_of(elo)
    e_host_pieces = pieces_of(hview.ctypes.data, n, elo, ehi)
    range_fits = sharded and 2 * e_n <= 64 * (1 << 30)
    range_dev = torch.empty(2 * e_n, dtype=torch.uint8, device="cuda") if range_fits else None

    def step_e2e():
        if not sharded:
            d.reset(True)
            m = d.feed(hview, out=out)
            return m, len(m) * adsbcu.MSG_DTYPE.itemsize, 2 * n
        if range_fits:
            pos = 0
            for p, cnt in e_host_pieces:
                o = (p - hview.ctypes.data)
                range_dev[pos:pos + 2 * cnt].copy_(host[o:o + 2 * cnt], non_blocking=True)
                pos += 2 * cnt
            torch.cuda.current_stream().synchronize()
            m, rep = d.replay_sharded(rank, world, [(range_dev.data_ptr(), e_n)], elo, e_halo, allgather, out)
            return m, len(m) * adsbcu.MSG_DTYPE.itemsize, 2 * e_n
        m, rep = d.replay_sharded(rank, world, e_host_pieces, elo, e_halo, allgather, out, host_pieces=True)
        return m, len(m) * adsbcu.MSG_DTYPE.itemsize, 2 * e_n * rep.passes

    sampler = ClockSampler(local_rank)
    sampler.start()                      # nvidia-smi needs ~100 ms to deliver its first row: start before the warm-up
    for _ in range(args.warmup):
        msgs = step_resident()[0]
    nmsg, digest = message_set_hash(msgs)
    t_w = time.perf_counter()
    while world == 1 and len(sampler.rows) < 2 and time.perf_counter() - t_w < 3.0:   # keep the GPU under load until sampling runs
        step_resident()
    barrier()
    kms = {k: [] for k in KEYS}
    launches, chunks, passes, rounds = 0, 0, 0, 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, ms, nl, chunks, passes, rounds = step_resident()
        for k in KEYS:
            kms[k].append(ms[k])
        launches += nl
    barrier()
    t_res = time.perf_counter() - t0

    # host->device ceiling for the e2e arm: one plain pinned-memory copy of the tile, CUDA events
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h2d_ms = []
    for _ in range(4):
        ev0.record()
        dev.copy_(host, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        h2d_ms.append(ev0.elapsed_time(ev1))
    h2d_gbs = 2.0 * n / (min(h2d_ms[1:]) * 1e-3) / 1e9

    e2e_steps = args.steps if total <= 4 * n else max(1, min(args.steps, 2))
    for _ in range(2 if total <= 4 * n else 0):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    d2h = h2d = 0
    for _ in range(e2e_steps):
        _, b_out, b_in = step_e2e()
        d2h += b_out
        h2d += b_in
    barrier()
    t_e2e = time.perf_counter() - t0
    clocks = sampler.stop()              # sampled over both timed regions (resident and end to end)

    tt = torch.tensor([t_res, t_e2e], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([nmsg, digest & 0xFFFFFFFF, digest >> 32, launches, passes], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    t_res, t_e2e = float(tt[0]), float(tt[1])
    nmsg_all = int(cnt[0])
    digest_all = (int(cnt[1]) + (int(cnt[2]) << 32)) & 0xFFFFFFFFFFFFFFFF

    if rank == 0:
        peak, peak_src = load_peaks()
        # Per-kernel device time of one step ON RANK 0 (CUDA events on the launching streams, summed over the step's chunks).
        # The roofline object describes the kernel with the largest share; both dense kernels are listed under
        # "kernels".  Algorithmic bytes: 2 B per sample, each kernel makes one launch per chunk.
        per_kernel = {k: float(np.mean(v)) for k, v in kms.items()}
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic_tab = json.load(f)
        except Exception:
            traffic_tab = {}
        # In the timed region the message kernel of chunk c runs on its own stream beside the front kernel of chunk c+1, so
        # an event pair around one of them also measures the share of the GPU the other one took.  The per-kernel roofline
        # numbers therefore come from a second handle with that overlap switched off (same kernels, same tile, same chunk
        # plan, CUDA events on the launching stream), measured here, live, right after the timed region; the concurrent
        # figures of the timed region are kept beside them.
        os.environ["ADSBCU_MSG_OVERLAP"] = "0"
        d_iso = adsbcu.Demod1090(nfix_crc=args.nfix, fix_df=1, max_feed_samples=max(n, 1 << 20), device=local_rank)
        iso = {"front": [], "messages": []}
        iso_chunks = 0
        for k in range(3 + 5):
            d_iso.reset(True)
            d_iso.feed_device(dev.data_ptr(), n, out)
            st = d_iso.stats()
            if k >= 3:
                iso["front"].append(st["ms_front"]); iso["messages"].append(st["ms_messages"])
                iso_chunks = st["chunks"]
        d_iso.close()
        os.environ.pop("ADSBCU_MSG_OVERLAP", None)
        kernels = {}
        bytes_step = 2.0 * my_n if sharded else 2.0 * n      # what rank 0's kernels streamed in one step (one pass over each of its ranges)
        for k, name in (("front", "front_kernel"), ("messages", "message_kernel")):
            ms = float(np.mean(iso[k]))
            gbs = 2.0 * n / (ms * 1e-3) / 1e9
            ratio = traffic_tab.get(name, {}).get("traffic_over_algorithmic")
            kernels[name] = {"ms_per_step": ms, "launches_per_step": iso_chunks, "avg_launch_ms": ms / max(iso_chunks, 1),
                             "algorithmic_bytes_per_launch": 2.0 * n / max(iso_chunks, 1), "achieved": gbs, "frac": gbs / peak,
                             "traffic": None if ratio is None else ratio * 2.0 * n / max(iso_chunks, 1),
                             "measured": "one 60 s tile, message kernel not overlapped with the next front kernel",
                             "concurrent_ms_per_step": per_kernel[k], "concurrent_frac": bytes_step / (per_kernel[k] * 1e-3) / 1e9 / peak}
        dom = max(kernels, key=lambda name: kernels[name]["ms_per_step"])
        kd = kernels[dom]
        line = {
            "metric": METRIC, "value": total * args.steps / t_res / 1e6, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_res / args.steps * 1e3,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u16/i32",
            "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": int(cnt[3]),
            "e2e": {"value": total * e2e_steps / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d // e2e_steps,
                    "d2h_bytes_per_step": d2h // e2e_steps, "steps": e2e_steps},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": kd["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": kd["frac"], "peak_source": peak_src, "launches_per_step": kd["launches_per_step"],
                         "algorithmic_bytes_per_launch": kd["algorithmic_bytes_per_launch"], "avg_launch_ms": kd["avg_launch_ms"],
                         "traffic": kd["traffic"],
                         "traffic_source": "profiles/ncu_traffic.json: dram read+write of one ncu --set full launch / that launch's algorithmic bytes",
                         "kernel_ms_per_step": kd["ms_per_step"],
                         "step_frac": 2.0 * total / world / (t_res / args.steps) / 1e9 / peak},
            "kernels": kernels,
            "pcie": {"h2d_gbs": h2d_gbs, "h2d_ceiling_msamples_s": h2d_gbs * 1e3 / 2.0,
                     "note": "plain pinned cudaMemcpy of the tile on this box; e2e cannot exceed world * this"},
            "breakdown_ms": dict(per_kernel, note="rank 0, device time per kernel summed over the chunks; order / resolve run on their own streams beside the next chunk's front + message kernels"),
            "messages_per_step": nmsg_all,
            "message_set": {"count": nmsg_all, "hash": f"{digest_all:016x}", "note": "sum mod 2^64 of a 64-bit mix of every accepted message record: independent of order and of the number of GPUs"},
        }
        if sharded:
            line["sharded"] = {"passes_rank0": passes, "passes_all_ranks": int(cnt[4]), "rounds": rounds, "range_samples_rank0": my_n}
        if world == 1 and not args.no_cpu_baseline:
            cores = 1
            cb, _ = cpu_baseline(iq, args.nfix, cores,
                                 f"one {args.seconds:g} s tile of the workload, one dump1090 process (1 demod thread + 1 reader thread)")
            line["cpu_baseline"] = cb
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
