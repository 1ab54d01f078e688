#!/usr/bin/env python
"""A few Path B HandleData calls (8 MiB and 262144-byte buffers) for ncu: python profiles/prof_pathb.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libadsb_b200 import adsbcu, synth
iq = synth.config_b(seconds=4.0, seed=0x1090B002)[0]
d = adsbcu.Demod1090()
out = np.zeros(1 << 17, dtype=adsbcu.MSG_B_DTYPE)
n = 0
for step in (8 << 20, 262144):
    d.reset_b()
    for k in range(0, min(iq.size, 2 * (8 << 20)), step):
        n += len(d.handle_data_b(iq[k:k + step], out=out))
print(n)
