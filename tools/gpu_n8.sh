#!/bin/bash
N=${1:-8}
set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit $?"
tail -1 gpurun_out/bench_n$N.json | cut -c1-300; tail -3 gpurun_out/bench_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload sweep --batch 131072 --steps 5 --warmup 3 > gpurun_out/sweep_n$N.json 2> gpurun_out/sweep_n$N.err; echo "sweep exit $?"
tail -1 gpurun_out/sweep_n$N.json | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --workload rollout --rollout-steps 200 > gpurun_out/rollout_n$N.json 2> gpurun_out/rollout_n$N.err; echo "rollout exit $?"
tail -1 gpurun_out/rollout_n$N.json | cut -c1-300
