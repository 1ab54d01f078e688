#!/bin/bash
# scan-kernel geometry: groups per SM x threads per group (tile = 24 samples x threads); builds each variant on the box
for geo in ${GEOS:-4x256 8x128}; do
  set -- ${geo/x/ }
  echo "== kGroups=$1 kGroupThreads=$2"
  ADSBCU_CFLAGS="-DADSBCU_GROUPS=$1 -DADSBCU_GROUP_THREADS=$2" python -c "from libadsb_b200 import build; build.build(force=True)" || continue
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
  python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1])
print(json.dumps({'value':round(d['value']),'e2e':round(d['e2e']['value']),'frac':round(d['roofline']['frac'],4),'breakdown_ms':{k:round(v,4) for k,v in d['breakdown_ms'].items()}}))"
done
python -c "from libadsb_b200 import build; build.build(force=True)"
