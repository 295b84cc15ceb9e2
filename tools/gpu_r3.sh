#!/bin/bash
# workloads round: all GPU tests, c3 / ref_h30 / sweep bench lines, ncu capture of a warm-started launch
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for wl in c3 ref_h30 sweep; do
  timeout 900 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl exit $?"
  tail -1 gpurun_out/bench_$wl.json | cut -c1-200; tail -2 gpurun_out/bench_$wl.err
done
python bench.py --steps 3 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_warp -s 11 -c 1 -f -o gpurun_out/prof_warm python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_warm.log 2>&1
ls -la gpurun_out/
