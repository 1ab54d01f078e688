#!/usr/bin/env python
"""Timeline probe of the order / resolve kernels (build with ADSBCU_CFLAGS=-DADSBCU_ORDER_TIMING, run with
ADSBCU_ORDER_TIMING_PRINT=1): python profiles/order_probe.py [seconds]"""
import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libadsb_b200 import adsbcu, synth
seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 12.0
iq = synth.config2(seconds=seconds, seed=0x10900002)[0]
n = iq.size // 2
dev = torch.from_numpy(iq).cuda()
out = np.zeros(max(4096, n // 500), dtype=adsbcu.MSG_DTYPE)
d = adsbcu.Demod1090(nfix_crc=1, fix_df=1, max_feed_samples=max(n, 1 << 20))
for rep in range(3):
    d.reset(True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    d.feed_device(dev.data_ptr(), n, out)
    t1 = time.perf_counter()
    print("feed cold %.3f ms" % ((t1 - t0) * 1e3), {k: round(v, 4) for k, v in d.stats().items() if k.startswith("ms_")}, flush=True)
