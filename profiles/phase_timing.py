"""Debug helper: per-phase cycle shares of the scan kernel (build with ADSBCU_CFLAGS=-DADSBCU_PHASE_TIMING)."""
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
v = np.array(list(buf)[:7], dtype=np.float64)
names = ["A mag+edges", "B patterns", "B2 preamble", "C3 classify (+C all if unsplit)", "D emit(+wait next)", "C1 DF pre-slice", "C2 unit slicing"]
# mark k measures the interval ending at barrier k, attributed to phase (k+4)%5: A ends at mark1 ...
print(d.stats()); d.h and None
for nme, x in zip(names, v):
    print(f"{nme:22s} {x / v.sum() * 100:5.1f}%")
