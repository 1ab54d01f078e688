N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_v16_config2_n$N.json 2> gpurun_out/bench_v16_config2_n$N.err; echo c2 rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload config4 --steps 3 > gpurun_out/bench_v16_config4_n$N.json 2> gpurun_out/bench_v16_config4_n$N.err; echo c4 rc=$?
tail -2 gpurun_out/bench_v16_config4_n$N.err
