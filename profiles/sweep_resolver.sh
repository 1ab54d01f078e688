#!/bin/bash
# ordered host pass on the calling thread (0) or on its own CUDA-free thread (1); chunk cap sweep; run on the GPU box
run() {
  echo "== $*"
  env "$@" python bench.py --steps 30 --warmup 3 --no-cpu-baseline | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1])
print(json.dumps({'value':round(d['value']),'ms_per_step':round(d['ms_per_step'],4),'e2e':round(d['e2e']['value']),'breakdown_ms':{k:round(v,3) for k,v in d['breakdown_ms'].items()}}))"
}
for tps in 6 8 10 12; do run ADSBCU_RESOLVER_THREAD=1 ADSBCU_WAIT=50 ADSBCU_TILES_PER_SEG=$tps; done
run ADSBCU_RESOLVER_THREAD=1 ADSBCU_WAIT=100 ADSBCU_TILES_PER_SEG=8
run ADSBCU_RESOLVER_THREAD=0
