# GPU tests (all) + headline bench line; extra workloads with: bash tools/gpu_r2_quick.sh c3 ref_h30 ...
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/q_pytest.log 2>&1; grep -E "parity-full|passed|failed|Error" gpurun_out/q_pytest.log | cut -c1-700
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/q_bench.json 2>gpurun_out/q_bench.err; cut -c1-260 gpurun_out/q_bench.json
for wl in "$@"; do timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/q_$wl.json 2> gpurun_out/q_$wl.err; echo "$wl exit $?"; cut -c1-260 gpurun_out/q_$wl.json; done
