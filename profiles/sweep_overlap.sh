#!/bin/bash
# A/B of ADSBCU_OVERLAP (filter/evaluate/signal of chunk c on a second, high-priority stream beside the scan of c+1)
mkdir -p gpurun_out
for v in 0 1 0 1; do
  ADSBCU_OVERLAP=$v python bench.py --no-cpu-baseline --steps 100 --warmup 3 > gpurun_out/bench_ov$v.json 2> gpurun_out/bench_ov$v.err
  python - <<P
import json
d = json.load(open("gpurun_out/bench_ov$v.json"))
print("overlap=$v", round(d["value"]), d["ms_per_step"], d["breakdown_ms"], "e2e", round(d["e2e"]["value"]))
P
done | tee gpurun_out/overlap_sweep.txt
ADSBCU_OVERLAP=1 python -m pytest tests/test_gpu_parity.py tests/test_sc16.py -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/overlap_sweep.txt
