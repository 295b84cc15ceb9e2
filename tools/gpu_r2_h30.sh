# H=30 (reference operating point) check: parity per gait, then throughput of each gait and the all-gait bench line.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity_fullsize.py tests/test_gpu_parity.py -m gpu -q -x -s -k "h30 or reference or c3" > gpurun_out/h30_pytest.log 2>&1; grep -E "parity-full|passed|failed|Error" gpurun_out/h30_pytest.log | cut -c1-500
for g in trot tws crawl stand; do timeout 200 python tools_profile.py 4096 reference_a1 3 $g 2>&1 | tail -1 | cut -c1-200; done
timeout 200 python tools_profile.py 16384 northstar_go1_h20 3 2>&1 | tail -1 | cut -c1-200
timeout 300 python bench.py --workload ref_h30 --steps 5 --warmup 3 2>/dev/null | cut -c1-200
