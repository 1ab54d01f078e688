#!/bin/bash
# front-kernel warps per CTA x message kernel on its own stream (beside the next chunk's front kernel): resident step, 144 M samples
out=gpurun_out/overlap2_sweep.txt
: > $out
for fw in 32 24 20 16; do
  ADSBCU_CFLAGS="-DADSBCU_FWARPS=$fw" python -c "from libadsb_b200 import build; build.build(force=True)" > /dev/null 2>&1 || { echo "build failed fw=$fw" >> $out; continue; }
  for ov in 0 1; do
    for mc in 6; do
      echo "== FWARPS=$fw MSG_OVERLAP=$ov" >> $out
      ADSBCU_MSG_OVERLAP=$ov python bench.py --steps 30 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1])
print('value', round(j['value']), 'ms', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['breakdown_ms'].items() if isinstance(v,float)}, j['message_set']['hash'])" >> $out
    done
  done
done
python -c "from libadsb_b200 import build; build.build(force=True)" > /dev/null 2>&1
cat $out
