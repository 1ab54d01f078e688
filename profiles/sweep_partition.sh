#!/bin/bash
# SMs set aside for the order / resolve / level kernels (green contexts) x message kernel beside the next front kernel
out=gpurun_out/partition_sweep.txt
: > $out
ADSBCU_DEBUG_CHUNKS=1 timeout 120 python profiles/prof_feed.py 2>&1 | grep -m1 "SM partition" >> $out
for part in 0 8 16; do
  for ov in 0 1; do
    echo "== SM_PARTITION=$part MSG_OVERLAP=$ov" >> $out
    ADSBCU_SM_PARTITION=$part ADSBCU_MSG_OVERLAP=$ov timeout 300 python bench.py --steps 30 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1])
print('value', round(j['value']), 'ms', round(j['ms_per_step'],4), 'e2e', round(j['e2e']['value']), {k:round(v,3) for k,v in j['breakdown_ms'].items() if isinstance(v,float)}, j['message_set']['hash'])" >> $out 2>&1
  done
done
cat $out
