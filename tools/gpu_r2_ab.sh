# A/B of library builds on the tile classes + the host-lane e2e path.
# SKIP_LANES=1 skips the host-lane part.
# Usage: bash tools/gpu_r2_ab.sh "<lib suffixes, '' = liblmpc.so>"   e.g. "_f0 '' _f2"
set -x
mkdir -p gpurun_out
L=legged_mpc_control_b200
for sfx in "$@"; do
  export LMPC_LIB_PATH=$PWD/$L/liblmpc$sfx.so
  echo "=== lib $LMPC_LIB_PATH"
  timeout 900 python -m pytest tests/test_gpu_parity_fullsize.py tests/test_gpu_parity.py -m gpu -q -x -k "h30 or reference or c3" 2>&1 | tail -2
  for g in trot tws crawl stand; do timeout 200 python tools_profile.py 4096 reference_a1 3 $g 2>&1 | tail -1 | cut -c1-160; done
  timeout 200 python tools_profile.py 16384 northstar_go1_h20 3 2>&1 | tail -1 | cut -c1-160
done
unset LMPC_LIB_PATH
[ -n "$SKIP_LANES" ] && exit 0
timeout 600 python -m pytest tests/test_gpu_host_lanes.py -m gpu -q -x 2>&1 | tail -3
for lanes in 1 2 4 8; do LMPC_HOST_LANES=$lanes timeout 300 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('lanes $lanes value', d['value'], 'e2e', d['e2e'], 'launches', d.get('gpu_launches'))"; done
