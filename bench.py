#!/usr/bin/env python
"""bench.py -- Msamples/s through the 1090 MHz demod+CRC hot path (BASELINE.json metric).

A step is one pass of the whole path (magnitude -> preamble -> 5-phase slicer -> CRC/repair ->
compaction -> ordered host pass) over one synthetic 60 s, 2.4 MS/s capture (configs[1]).

  value   whole-job Msamples/s, I/Q already resident in HBM (adsbcu_1090_feed_device)
  e2e     same metric through adsbcu_1090_feed with HOST buffers (H2D + record D2H inside)
  roofline   the scan kernel: 2 algorithmic bytes per complex sample / its CUDA-event duration
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seconds", type=float, default=60.0, help="capture length per GPU (config 2: 60 s)")
    ap.add_argument("--nfix", type=int, default=1, help="0 none, 1 --fix, 2 --fix-2bit")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="config2", choices=["config2", "config3"],
                    help="config2 (default, the metric's configuration) or config3 (dense overlapping traffic, use --nfix 2)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from libadsb_b200 import synth
    if args.workload == "config2":
        workload = (f"config2: synthetic 2.4 MS/s uint8 I/Q, {args.seconds:g} s per GPU, 1000 msg/s DF17/DF11, "
                    f"6-30 dB SNR, sigma=2 LSB, nfix_crc={args.nfix}")
        make = lambda seed: synth.config2(seconds=args.seconds, seed=seed)[0]                   # noqa: E731
    else:
        workload = (f"config3: synthetic 2.4 MS/s uint8 I/Q, {args.seconds:g} s per GPU, Poisson 2000 msg/s with overlaps, "
                    f"DF17/11/4/5/20/21/18 mix, 1-2 bit errors, sigma=2 LSB, nfix_crc={args.nfix}")
        make = lambda seed: synth.config3(seconds=args.seconds, seed=seed)[0]                   # noqa: E731
    config = {"workload": workload, "samples_per_gpu": int(round(args.seconds * synth.FS_1090)),
              "l2_policy": "input (2 B/sample, 288 MB at 60 s) larger than L2", "sharding": "contiguous ranges, weak"}

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        iq = make(0x10900002)
        cores = os.cpu_count() or 1
        for _ in range(args.warmup):
            cpu_baseline(iq, args.nfix, cores, "")
        times = []
        desc = f"the whole {args.seconds:g} s capture per step, split into {cores} contiguous shards, one dump1090 process each"
        cb = None
        for _ in range(args.steps):
            cb, dt = cpu_baseline(iq, args.nfix, cores, desc)
            times.append(dt)
        dt = float(np.mean(times))
        v = (iq.size // 2) / dt / 1e6
        cb["value"] = v
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16/i32",
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

    iq = make(0x10900002 + rank)
    n = iq.size // 2
    host = torch.from_numpy(iq).pin_memory()
    dev = host.cuda(non_blocking=False)
    hview = host.numpy()
    d = adsbcu.Demod1090(nfix_crc=args.nfix, fix_df=1, max_feed_samples=max(n, 1 << 20), device=local_rank)
    # out[] in pinned memory: the engine's level kernel writes the accepted messages there itself (no host copy)
    out_pinned = torch.zeros(max(4096, n // 500) * adsbcu.MSG_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    out = out_pinned.numpy().view(adsbcu.MSG_DTYPE)

    def step_resident():
        d.reset(True)
        return d.feed_device(dev.data_ptr(), n, out)

    def step_e2e():
        d.reset(True)
        return d.feed(hview, out=out)

    sampler = ClockSampler(local_rank)
    sampler.start()                      # nvidia-smi needs ~100 ms to deliver its first row: start before the warm-up
    for _ in range(args.warmup):
        msgs = step_resident()
    nmsg = len(msgs)
    t_w = time.perf_counter()
    while len(sampler.rows) < 2 and time.perf_counter() - t_w < 3.0:   # keep the GPU under the same load until sampling runs
        step_resident()
    barrier()
    kms = {"front": [], "messages": [], "order": [], "resolve": []}
    launches, chunks = 0, 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_resident()
        st = d.stats()
        for k in kms:
            kms[k].append(st["ms_" + k])
        launches += st["launches"] + 1               # + the one-thread state kernel of reset(); memsets and copies are not counted
        chunks = st["chunks"]
    barrier()
    t_res = time.perf_counter() - t0

    # host->device ceiling for the e2e arm: one plain pinned-memory copy of the same capture, CUDA events
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h2d_ms = []
    for _ in range(4):
        ev0.record()
        dev.copy_(host, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        h2d_ms.append(ev0.elapsed_time(ev1))
    h2d_gbs = 2.0 * n / (min(h2d_ms[1:]) * 1e-3) / 1e9

    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(args.steps):
        m = step_e2e()
        st = d.stats()
        d2h += st["resolved_records"] * 64 + 24
    barrier()
    t_e2e = time.perf_counter() - t0
    clocks = sampler.stop()              # sampled over both timed regions (resident and end to end)

    tt = torch.tensor([t_res, t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_res, t_e2e = float(tt[0]), float(tt[1])

    if rank == 0:
        peak, peak_src = load_peaks()
        # Per-kernel device time of one step (CUDA events on the launching streams, summed over the step's chunks).
        # The roofline object describes the kernel with the largest share; both dense kernels are listed under
        # "kernels".  Algorithmic bytes: 2 B per sample, each kernel makes one launch per chunk.
        per_kernel = {k: float(np.mean(v)) for k, v in kms.items()}
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic_tab = json.load(f)
        except Exception:
            traffic_tab = {}
        kernels = {}
        for k, name in (("front", "front_kernel"), ("messages", "message_kernel")):
            ms = per_kernel[k]
            gbs = 2.0 * n / (ms * 1e-3) / 1e9
            ratio = traffic_tab.get(name, {}).get("traffic_over_algorithmic")
            kernels[name] = {"ms_per_step": ms, "launches_per_step": chunks, "avg_launch_ms": ms / max(chunks, 1),
                             "algorithmic_bytes_per_launch": 2.0 * n / max(chunks, 1), "achieved": gbs, "frac": gbs / peak,
                             "traffic": None if ratio is None else ratio * 2.0 * n / max(chunks, 1)}
        dom = max(kernels, key=lambda name: kernels[name]["ms_per_step"])
        kd = kernels[dom]
        line = {
            "metric": METRIC, "value": world * n * args.steps / t_res / 1e6, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_res / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16/i32",
            "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": world * n * args.steps / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": 2 * n,
                    "d2h_bytes_per_step": d2h // args.steps},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": kd["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": kd["frac"], "peak_source": peak_src, "launches_per_step": kd["launches_per_step"],
                         "algorithmic_bytes_per_launch": kd["algorithmic_bytes_per_launch"], "avg_launch_ms": kd["avg_launch_ms"],
                         "traffic": kd["traffic"],
                         "traffic_source": "profiles/ncu_traffic.json: dram read+write of one ncu --set full launch / that launch's algorithmic bytes",
                         "kernel_ms_per_step": kd["ms_per_step"],
                         "step_frac": 2.0 * n / (t_res / args.steps) / 1e9 / peak},
            "kernels": kernels,
            "pcie": {"h2d_gbs": h2d_gbs, "h2d_ceiling_msamples_s": h2d_gbs * 1e3 / 2.0,
                     "note": "plain pinned cudaMemcpy of the capture on this box; e2e cannot exceed world * this"},
            "breakdown_ms": dict(per_kernel, note="device time per kernel summed over the chunks; order / resolve run on their own streams beside the next chunk's front + message kernels"),
            "messages_per_step": nmsg,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import oracle1090 as O
            cores = 1
            cb, _ = cpu_baseline(iq, args.nfix, cores,
                                 f"the whole {args.seconds:g} s capture, one dump1090 process (1 demod thread + 1 reader thread)")
            line["cpu_baseline"] = cb
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
