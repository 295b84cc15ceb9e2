#!/bin/bash
# end-of-round evidence for the final binary: smoke, all GPU tests, bench lines of every workload,
# ncu launch list + full capture of solve_warp_kernel
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
for wl in c3 ref_h30 sweep; do python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl exit $?"; done
python bench.py --workload rollout --rollout-steps 1000 > gpurun_out/rollout.json 2> gpurun_out/rollout.err; echo "rollout exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ref.json 2> gpurun_out/ref.err; echo "ref exit $?"
python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_warp -s 3 -c 1 -f -o gpurun_out/prof_warp python bench.py --steps 3 --warmup 3 --no-warm > gpurun_out/ncu_full.log 2>&1
ls gpurun_out | wc -l
