import os, sys, numpy as np
sys.path.insert(0, '/root/repo')
import torch
from libadsb_b200 import adsbcu, synth
iq = synth.config2(seconds=12.0, seed=0x10900002)[0]
n = iq.size // 2
dev = torch.from_numpy(iq).cuda()
out = np.zeros(max(4096, n // 500), dtype=adsbcu.MSG_DTYPE)
d = adsbcu.Demod1090(nfix_crc=1, fix_df=1, max_feed_samples=max(n, 1 << 20))
for rep in range(3):
    d.reset(True)
    d.feed_device(dev.data_ptr(), n, out)
    print("feed cold ", {k: round(v, 4) for k, v in d.stats().items() if k.startswith("ms_")})
    d.reset(False)
    d.feed_device(dev.data_ptr(), n, out)
    print("feed warm ", {k: round(v, 4) for k, v in d.stats().items() if k.startswith("ms_")})
d2 = adsbcu.Demod1090(nfix_crc=1, fix_df=1, max_feed_samples=max(n, 1 << 20))
for rep in range(2):
    d2.reset(True); d2.seek(0)
    c = d2.scan_candidates(dev.data_ptr(), n)
    print("candidates", len(c), {k: round(v, 4) for k, v in d2.stats().items() if k.startswith("ms_")})
