#!/usr/bin/env python
"""One resident UAT feed (default 2 s) for ncu launch lists: python profiles/prof_uat.py [seconds]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libadsb_b200 import adsbcu, synth978
seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
iq = synth978.config5(seconds=seconds)[0]
dev = torch.from_numpy(iq).cuda()
d = adsbcu.Demod1090()
out = np.zeros(1 << 16, adsbcu.UAT_DTYPE)
for _ in range(2):
    d.reset_uat()
    t0 = time.perf_counter()
    f = d.handle_data_uat_device(dev.data_ptr(), iq.size, out)
    dt = time.perf_counter() - t0
print(len(f), d.uat_stats(), "wall ms", dt * 1e3)
