#!/bin/bash
# usage: bash profiles/run_prof.sh <tag> : launch list + ncu --set full of the main kernels of one 12 s resident feed
# (the plain run first: ncu only after the same command exited 0 without it)
tag=${1:-x}
export ADSBCU_DEBUG_CHUNKS=1
python profiles/prof_feed.py > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$tag.csv python profiles/prof_feed.py > /dev/null 2>&1
for k in front_kernel message_kernel order_kernel resolve_kernel level_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o gpurun_out/prof_${k}_$tag python profiles/prof_feed.py > /dev/null 2>&1
done
cat gpurun_out/plain_$tag.log
