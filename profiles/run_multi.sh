#!/bin/bash
# usage: bash profiles/run_multi.sh N [tag] : default workload (weak: N x 60 s, one capture) and config 4 (strong: 24 h) on N GPUs
N=$1; T=${2:-v23}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_${T}_config2_n$N.json 2> gpurun_out/bench_${T}_config2_n$N.err; echo c2 rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload config4 --steps 3 > gpurun_out/bench_${T}_config4_n$N.json 2> gpurun_out/bench_${T}_config4_n$N.err; echo c4 rc=$?
grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/bench_${T}_config4_n$N.err | tail -3
