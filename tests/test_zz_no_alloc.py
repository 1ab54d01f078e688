"""adsbcu_1090_modeac_block_static: the Mode A/C block with its scratch in the library's file-scope device array
(no allocation, no copy, no wait inside the call) against adsbcu_1090_modeac_block on the same blocks.  Integer
work: the replies must be identical, field by field.  Collected last on purpose."""
import os

import numpy as np
import pytest

BLOCK = 131072      # ADSBCU_MODEAC_STATIC_MAX_SAMPLES == MODES_MAG_BUF_SAMPLES
EMU = os.environ.get("ADSBCU_TEST_EMU") == "1"


class DeviceBuffers:
    """Caller-owned device memory for the call (torch is only the allocator here; under the SIMT emulator of the CPU
    suite "device" memory is host memory)."""

    def __init__(self, mod, cap):
        self.mod, self.cap = mod, cap
        nbytes_out = cap * mod.MSG_DTYPE.itemsize
        if EMU:
            self.iq = np.zeros(2 * BLOCK + 64, np.uint8)
            self.ov = np.zeros(2 * 326, np.uint8)
            self.out = np.zeros(nbytes_out, np.uint8)
            self.res = np.zeros(2, np.uint32)
        else:
            import torch
            self.torch = torch
            self.iq = torch.zeros(2 * BLOCK + 64, dtype=torch.uint8, device="cuda:0")
            self.ov = torch.zeros(2 * 326, dtype=torch.uint8, device="cuda:0")
            self.out = torch.zeros(nbytes_out, dtype=torch.uint8, device="cuda:0")
            self.res = torch.zeros(2, dtype=torch.int32, device="cuda:0")

    @staticmethod
    def _ptr(b):
        return b.ctypes.data if isinstance(b, np.ndarray) else b.data_ptr()

    def _put(self, dst, src):
        if EMU:
            dst[:src.size] = src
        else:
            dst[:src.size].copy_(self.torch.from_numpy(src))

    def run(self, d, iq, ov, timestamp, cap=None):
        cap = self.cap if cap is None else cap
        iq = np.ascontiguousarray(iq, np.uint8)
        ov = np.ascontiguousarray(ov if ov is not None else np.zeros(0, np.uint8), np.uint8)
        self._put(self.iq, iq)
        self._put(self.ov, ov)
        if not EMU:
            self.out.fill_(0xEE)            # stale bytes must not survive in the records the call reports
            self.res.fill_(-1)
            self.torch.cuda.synchronize()
        d.modeac_block_static(self._ptr(self.iq), iq.size // 2, self._ptr(self.ov) if ov.size else 0, ov.size // 2, timestamp,
                              self._ptr(self.out), cap, self._ptr(self.res), 0)
        if EMU:
            res, raw = self.res.copy(), self.out.copy()
        else:
            self.torch.cuda.synchronize()
            res = self.res.cpu().numpy().view(np.uint32)
            raw = self.out.cpu().numpy()
        n = int(res[0])
        return raw.view(self.mod.MSG_DTYPE)[:min(n, cap)].copy(), n, int(res[1])


def blocks_of(iq, block=BLOCK):
    n = iq.size // 2
    for s in range(0, n, block):
        e = min(n, s + block)
        yield iq[2 * s:2 * e], iq[2 * max(0, s - 326):2 * s], 5 * s


@pytest.mark.gpu
def test_static_scratch_matches_allocating_entry_point(adsbcu_mod):
    from libadsb_b200 import synth
    iq = synth.config_ac(seconds=0.12, seed=21, sigma=2.0)[0]
    d = adsbcu_mod.Demod1090(0, 1)
    buf = DeviceBuffers(adsbcu_mod, 1 << 13)
    total = 0
    calls = list(blocks_of(iq))
    assert len(calls) >= 2
    # consecutive calls on the shared scratch: full blocks, then a short one (stale magnitudes and records behind it),
    # then the first block again
    calls += [(calls[1][0][:2 * 40000], calls[1][1], 77), calls[0]]
    for blk, ov, ts in calls:
        want = d.modeac_block(blk, ov if ov.size else None, ts)
        got, n, flags = buf.run(d, blk, ov, ts)
        assert flags == 0 and n == len(want)
        assert got.tobytes() == want.tobytes()
        total += n
    assert total > 20


@pytest.mark.gpu
def test_static_scratch_limits_and_flags(adsbcu_mod):
    from libadsb_b200 import synth
    iq = synth.config_ac(seconds=0.06, seed=22, sigma=2.0)[0][:2 * BLOCK]
    d = adsbcu_mod.Demod1090(0, 1)
    buf = DeviceBuffers(adsbcu_mod, 1 << 13)
    want = d.modeac_block(iq, None, 0)
    assert len(want) > 4
    got, n, flags = buf.run(d, iq, None, 0, cap=3)          # out[] too small: count reported, first cap written, flag set
    assert n == len(want) and flags == 2 and got.tobytes() == want[:3].tobytes()
    got, n, flags = buf.run(d, iq[:2], None, 0)              # fewer than two samples: nothing to do
    assert (n, flags, len(got)) == (0, 0, 0)
    with pytest.raises(adsbcu_mod.AdsbcuError, match="ADSBCU_MODEAC_STATIC_MAX_SAMPLES"):
        d.modeac_block_static(buf._ptr(buf.iq), BLOCK + 1, 0, 0, 0, buf._ptr(buf.out), 16, buf._ptr(buf.res), 0)
    with pytest.raises(adsbcu_mod.AdsbcuError):
        d.modeac_block_static(buf._ptr(buf.iq) + 2, 1000, 0, 0, 0, buf._ptr(buf.out), 16, buf._ptr(buf.res), 0)   # misaligned
    got, n, flags = buf.run(d, iq, None, 0)                  # and the scratch is fine afterwards
    assert flags == 0 and got.tobytes() == want.tobytes()
