set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/b_pytest.log 2>&1; tail -3 gpurun_out/b_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/b_bench.json 2>gpurun_out/b_bench.err; cut -c1-300 gpurun_out/b_bench.json
for wl in c3 ref_h30; do timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/b_$wl.json 2> gpurun_out/b_$wl.err; echo "$wl exit $?"; cut -c1-300 gpurun_out/b_$wl.json; done
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv; nproc
