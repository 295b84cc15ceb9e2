set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/q_pytest.log 2>&1; tail -3 gpurun_out/q_pytest.log
for wl in ref_h30 sweep c3; do timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/q_$wl.json 2> gpurun_out/q_$wl.err; echo "$wl exit $?"; cat gpurun_out/q_$wl.json | cut -c1-400; done
timeout 300 python bench.py --workload rollout --rollout-steps 500 > gpurun_out/q_rollout.json 2>gpurun_out/q_rollout.err; cut -c1-400 gpurun_out/q_rollout.json
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/q_bench.json 2>gpurun_out/q_bench.err; cut -c1-300 gpurun_out/q_bench.json
