"""adsbcu_1090_decode_device: header fields, extended squitter and Comm-B of device-resident records in one launch, against the
three host-array entry points (which are pinned to the reference in test_fields_a / _es / _mb) - byte-identical records, with
every combination of outputs, tile-edge counts and a grid-stride count."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
from make_golden_es import es_frames  # noqa: E402
from make_golden_mb import build  # noqa: E402
from test_fields_es import as_msgs  # noqa: E402

EMU = os.environ.get("ADSBCU_TEST_EMU") == "1"


class Dev:
    """Caller-owned device arrays (torch is only the allocator; under the SIMT emulator device memory is host memory)."""

    def __init__(self, nbytes):
        if EMU:
            self.buf = np.full(nbytes, 0xEE, np.uint8)
        else:
            import torch
            self.torch = torch
            self.buf = torch.full((nbytes,), 0xEE, dtype=torch.uint8, device="cuda:0")

    @property
    def ptr(self):
        return self.buf.ctypes.data if EMU else self.buf.data_ptr()

    def put(self, a):
        raw = np.frombuffer(a.tobytes(), np.uint8)
        if EMU:
            self.buf[:raw.size] = raw
        else:
            self.buf[:raw.size].copy_(self.torch.from_numpy(raw.copy()))

    def get(self, dtype, n):
        if not EMU:
            self.torch.cuda.synchronize()
        raw = self.buf[:n * dtype.itemsize] if EMU else self.buf[:n * dtype.itemsize].cpu().numpy()
        return np.frombuffer(raw.tobytes(), dtype)


@pytest.mark.gpu
def test_device_resident_decode_equals_host_entry_points(adsbcu_mod):
    rng = np.random.default_rng(626)
    nmax = 40000 if EMU else 700000                  # more than 148 * 4 tiles of 256: the grid-stride loop runs
    frames = np.concatenate([es_frames(rng, nmax // 2), build(rng, nmax // 22 + 1), rng.integers(0, 256, (nmax // 8, 14), dtype=np.uint8)])
    frames = frames[rng.permutation(len(frames))][:nmax]
    msgs = as_msgs(adsbcu_mod.MSG_DTYPE, frames, (rng.random(len(frames)) < 0.1).astype(np.uint8))
    d = adsbcu_mod.Demod1090()
    want = {"a": d.decode_fields_a(msgs), "es": d.decode_es(msgs), "mb": d.decode_commb(msgs)}
    types = {"a": adsbcu_mod.FIELDS_A_DTYPE, "es": adsbcu_mod.FIELDS_ES_DTYPE, "mb": adsbcu_mod.FIELDS_MB_DTYPE}
    d_in = Dev(len(msgs) * adsbcu_mod.MSG_DTYPE.itemsize)
    d_in.put(msgs)
    outs = {k: Dev(len(msgs) * t.itemsize + 64) for k, t in types.items()}
    for n in (len(msgs), 1, 255, 256, 257, 5000):
        for which in (("a", "es", "mb"), ("es",), ("a", "mb"), ("mb",)):
            ptrs = {k: outs[k].ptr if k in which else 0 for k in types}
            d.decode_device(d_in.ptr, n, ptrs["a"], ptrs["es"], ptrs["mb"])
            for k in which:
                got = outs[k].get(types[k], n)
                assert got.tobytes() == want[k][:n].tobytes(), (n, which, k)
            if n == len(msgs):
                for k in which:                        # nothing written behind the last record
                    tail = outs[k].buf[n * types[k].itemsize:]
                    assert bool((tail == 0xEE).all())
    # arrays that are only 4-byte aligned (the word-copy path of the tiles)
    n = 3001
    d_off = Dev(n * adsbcu_mod.MSG_DTYPE.itemsize + 64)
    d_off.put(np.concatenate([np.zeros(4, np.uint8), np.frombuffer(msgs[:n].tobytes(), np.uint8)]))
    d.decode_device(d_off.ptr + 4, n, outs["a"].ptr + 4, outs["es"].ptr + 12, outs["mb"].ptr + 8)
    for k, off in (("a", 4), ("es", 12), ("mb", 8)):
        got = outs[k].get(np.dtype("u1"), off + n * types[k].itemsize)[off:]
        assert got.tobytes() == want[k][:n].tobytes(), k
    d.decode_device(d_in.ptr, 0, outs["a"].ptr)        # nothing to do
    with pytest.raises(adsbcu_mod.AdsbcuError):
        d.decode_device(d_in.ptr, 10)                  # no output at all
    with pytest.raises(adsbcu_mod.AdsbcuError):
        d.decode_device(d_in.ptr + 2, 10, outs["a"].ptr)
