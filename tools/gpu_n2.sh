#!/bin/bash
# 2-GPU check: headline bench and config-5 rollout under torchrun, reference arm under torchrun
set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 exit $?"
tail -1 gpurun_out/bench_n2.json | cut -c1-400; tail -3 gpurun_out/bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload rollout --rollout-steps 200 > gpurun_out/rollout_n2.json 2> gpurun_out/rollout_n2.err; echo "rollout n2 exit $?"
tail -1 gpurun_out/rollout_n2.json | cut -c1-600; tail -3 gpurun_out/rollout_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/ref_n2.json 2> gpurun_out/ref_n2.err; echo "ref n2 exit $?"
tail -1 gpurun_out/ref_n2.json | cut -c1-300
