#!/bin/bash
# warm start + rollout round: new GPU tests, all GPU tests, bench (headline + rollout)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_warm_rollout.py -m gpu -x -q > gpurun_out/pytest_warm.log 2>&1; echo "exit $?" >> gpurun_out/pytest_warm.log
tail -30 gpurun_out/pytest_warm.log
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json | cut -c1-600; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --workload rollout --rollout-steps 200 > gpurun_out/rollout.json 2> gpurun_out/rollout.err; echo "rollout exit $?"
cat gpurun_out/rollout.json; tail -3 gpurun_out/rollout.err
