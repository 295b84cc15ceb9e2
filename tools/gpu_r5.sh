#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/gpu_sweep_fail.py 65536 2>&1 | grep -E "histogram|status 3" 
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e'], d['warm_start'])
PY
for t in 256 512; do
LMPC_CLS3_THREADS=$t LMPC_DEBUG_STATS=1 python tools_profile.py 2048 reference_a1 2 stand 2>&1 | tail -1
done
timeout 600 python bench.py --workload rollout --rollout-steps 200 2>/dev/null | cut -c1-900
