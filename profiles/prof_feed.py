#!/usr/bin/env python
"""One resident feed of a config-2 capture (default 12 s = 28.8 M samples) for ncu runs:
   python profiles/prof_feed.py [seconds] [nfix] [workload]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from libadsb_b200 import adsbcu, build, synth  # noqa: E402

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 12.0
nfix = int(sys.argv[2]) if len(sys.argv) > 2 else 1
workload = sys.argv[3] if len(sys.argv) > 3 else "config2"
build.build()
iq = (synth.config2 if workload == "config2" else synth.config3)(seconds=seconds, seed=0x10900002)[0]
n = iq.size // 2
dev = torch.from_numpy(iq).cuda()
d = adsbcu.Demod1090(nfix_crc=nfix, fix_df=1, max_feed_samples=max(n, 1 << 20))
out = np.zeros(max(4096, n // 500), dtype=adsbcu.MSG_DTYPE)
for _ in range(2):
    d.reset(True)
    msgs = d.feed_device(dev.data_ptr(), n, out)
print(len(msgs), d.stats())
