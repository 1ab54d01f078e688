"""Throughput of the three decodeModesMessage entry points through the C ABI (HOST arrays in and out, as a caller has them):
adsbcu_1090_decode_fields / _decode_es / _decode_commb, 4 M records per call, best of 5 after a warm-up call; next to them the
oracle restatement on one host core (the reference's decodeModesMessage is the same per-message C).  Writes JSON lines."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from libadsb_b200 import adsbcu  # noqa: E402
from oracle import oracle1090 as O  # noqa: E402
from make_golden_es import es_frames  # noqa: E402
from make_golden_mb import build  # noqa: E402


def msgs_of(frames):
    m = np.zeros(len(frames), adsbcu.MSG_DTYPE)
    m["msg"] = frames
    m["verbatim"] = frames
    m["msgbits"] = 112
    m["msgtype"] = frames[:, 0] >> 3
    return m


def best_of(fn, reps=5):
    fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    return min(t)


if __name__ == "__main__":
    rng = np.random.default_rng(7)
    n = 1 << 22
    es = es_frames(rng, n)
    mb = build(rng, n // 11 + 1)[:n]
    d = adsbcu.Demod1090()
    out = []
    device_only = "--device-only" in sys.argv            # for the ncu capture of fields_all_kernel
    for name, frames, gpu, cpu, rec_bytes in () if device_only else (
            ("A fields (adsbcu_1090_decode_fields)", es, d.decode_fields_a, None, 56 + 32),
            ("A extended squitter (adsbcu_1090_decode_es)", es, d.decode_es, O.decode_es, 56 + 120),
            ("A Comm-B (adsbcu_1090_decode_commb)", mb, d.decode_commb, O.decode_commb, 56 + 84)):
        m = msgs_of(frames)
        t = best_of(lambda: gpu(m))
        line = {"path": name, "records": len(m), "mrecords_s": len(m) / t / 1e6, "pcie_bytes_per_record": rec_bytes}
        if cpu is not None:
            sample = frames[: 1 << 20]
            tc = best_of(lambda: cpu(sample), 3)
            line["oracle_one_core_mrecords_s"] = len(sample) / tc / 1e6
        out.append(line)
        print(json.dumps(line), flush=True)
    # the same three decoders in one launch over device-resident records (adsbcu_1090_decode_device), timed with CUDA events on
    # the launching stream; algorithmic bytes per record = 56 in + 32 + 120 + 84 out
    import torch
    mixed = np.concatenate([es[: n // 2], mb[: n // 2]])[rng.permutation(n)]
    m = msgs_of(mixed)
    d_in = torch.from_numpy(np.frombuffer(m.tobytes(), np.uint8).copy()).cuda()
    d_a = torch.empty(n * 32, dtype=torch.uint8, device="cuda")
    d_es = torch.empty(n * 120, dtype=torch.uint8, device="cuda")
    d_mb = torch.empty(n * 84, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 0) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 0
    for which, ptrs, nbytes in (("fields + es + commb", (d_a.data_ptr(), d_es.data_ptr(), d_mb.data_ptr()), 292), ("es only", (0, d_es.data_ptr(), 0), 176)):
        for _ in range(3):
            d.decode_device(d_in.data_ptr(), n, *ptrs, stream=stream)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 20
        ev[0].record()
        for _ in range(reps):
            d.decode_device(d_in.data_ptr(), n, *ptrs, stream=stream)
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / reps
        line = {"path": "A decode_device (" + which + ")", "records": n, "ms_per_launch": ms, "grecords_s": n / ms / 1e6,
                "algorithmic_gb_s": n * nbytes / ms / 1e6, "hbm_peak_gb_s": peak, "frac": (n * nbytes / ms / 1e6 / peak) if peak else None}
        print(json.dumps(line), flush=True)
