#!/bin/bash
# message kernel with the software-pipelined opening loads: resident CTAs per SM (register budget) x grid
out=gpurun_out/msg_pipe_sweep.txt
: > $out
for mc in 6 5 4; do
  ADSBCU_CFLAGS="-DADSBCU_MSG_MINCTAS=$mc" python -c "from libadsb_b200 import build; build.build(force=True)" > /dev/null 2>&1 || { echo "build failed mc=$mc" >> $out; continue; }
  for ov in 1 0; do
    echo "== MSG_MINCTAS=$mc grid=$((148*mc)) MSG_OVERLAP=$ov" >> $out
    ADSBCU_MSG_GRID=$((148*mc)) ADSBCU_MSG_OVERLAP=$ov timeout 300 python bench.py --steps 30 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1])
print('value', round(j['value']), 'ms', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['breakdown_ms'].items() if isinstance(v,float)}, 'iso msg', round(j['kernels']['message_kernel']['ms_per_step'],3), 'front', round(j['kernels']['front_kernel']['ms_per_step'],3), j['message_set']['hash'])" >> $out 2>&1
  done
done
python -c "from libadsb_b200 import build; build.build(force=True)" > /dev/null 2>&1
cat $out
