"""Debug helper: per-phase cycle shares of the scan kernel as seen by thread 0 of every group.
Build first with: ADSBCU_CFLAGS=-DADSBCU_PHASE_TIMING python -c 'from libadsb_b200 import build; build.build(force=True)'"""
import ctypes as C, sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libadsb_b200 import adsbcu, synth
iq, _, _ = synth.config2(seconds=20.0, seed=0x10900002)
n = iq.size // 2
dev = torch.from_numpy(iq).cuda()
d = adsbcu.Demod1090(nfix_crc=int(os.environ.get("NFIX", "1")), max_feed_samples=n)
out = np.zeros(200000, dtype=adsbcu.MSG_DTYPE)
L = adsbcu.load()
buf = (C.c_ulonglong * 8)()
d.feed_device(dev.data_ptr(), n, out)
L.adsbcu_debug_phase_cycles(buf, 1)
d.reset(True); d.feed_device(dev.data_ptr(), n, out)
L.adsbcu_debug_phase_cycles(buf, 1)
v = np.array(list(buf), dtype=np.float64)
names = ["-", "wait for the group (top barrier)", "A magnitudes+edges+planes", "B edge patterns", "B2 preamble test", "C1 DF pre-filter",
         "C2 extract+CRC+classify", "D ordered emission"]
tiles = n / 5848.0
print(d.stats())
for nme, x in zip(names, v):
    print(f"{nme:34s} {x / v.sum() * 100:5.1f}%   {x / tiles:8.0f} cycles per tile")
