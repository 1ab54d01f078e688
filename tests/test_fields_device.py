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


@pytest.mark.gpu
def test_message_set_stays_on_the_device(adsbcu_mod, oracle_mod):
    """adsbcu_1090_feed_device with out[] in device memory, then adsbcu_1090_decode_device on that array: I/Q in, decoded
    fields out, the messages never cross PCIe.  The records equal those of the ordinary feed (which the parity suite holds
    against the reference), the decoded fields equal the oracle's on those messages."""
    import ctypes
    from libadsb_b200 import synth
    iq = synth.config3(seconds=0.25, seed=97, msgs_per_s=4000.0, pool=200)[0]
    n = iq.size // 2
    want = adsbcu_mod.Demod1090(0, 1).feed(iq)
    assert len(want) > 300
    cap = len(want) + 100
    d_iq, d_msgs = Dev(iq.size + 64), Dev(cap * adsbcu_mod.MSG_DTYPE.itemsize)
    d_iq.put(iq)
    if EMU:
        adsbcu_mod.load().adsbcu_emu_mark_device(ctypes.c_void_p(d_msgs.ptr), ctypes.c_size_t(d_msgs.buf.size))
    try:
        d = adsbcu_mod.Demod1090(0, 1)
        got_n = d.feed_device_resident(d_iq.ptr, n, d_msgs.ptr, cap)
        assert got_n == len(want)
        got = d_msgs.get(adsbcu_mod.MSG_DTYPE, got_n)
        assert got.tobytes() == want.tobytes()
        d_es, d_mb, d_a = Dev(cap * 120), Dev(cap * 84), Dev(cap * 32)
        d.decode_device(d_msgs.ptr, got_n, d_a.ptr, d_es.ptr, d_mb.ptr)
        es = d_es.get(adsbcu_mod.FIELDS_ES_DTYPE, got_n)
        mb = d_mb.get(adsbcu_mod.FIELDS_MB_DTYPE, got_n)
        assert es.tobytes() == oracle_mod.decode_es(want["msg"], want["correctedbits"]).tobytes()
        assert mb.tobytes() == oracle_mod.decode_commb(want["msg"], want["correctedbits"]).tobytes()
        assert (es["es"] == 1).sum() > 50 and ((want["msgtype"] == 20) | (want["msgtype"] == 21)).sum() > 20
        # too small a device array: the count is still reported, the call says so
        d2 = adsbcu_mod.Demod1090(0, 1)
        with pytest.raises(adsbcu_mod.AdsbcuError, match="too small"):
            d2.feed_device_resident(d_iq.ptr, n, d_msgs.ptr, 10)
        with pytest.raises(adsbcu_mod.AdsbcuError, match="8-byte"):
            adsbcu_mod.Demod1090(0, 1).feed_device_resident(d_iq.ptr, n, d_msgs.ptr + 4, cap - 1)
    finally:
        if EMU:
            adsbcu_mod.load().adsbcu_emu_mark_device(None, ctypes.c_size_t(0))
