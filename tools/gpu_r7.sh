#!/bin/bash
mkdir -p gpurun_out
for nw in 1 2 4; do
  echo "== NW $nw"
  LMPC_TILE_NW=$nw python tools_profile.py 16384 northstar_go1_h20 2 2>&1 | tail -1 | cut -c1-120
  LMPC_TILE_NW=$nw python tools_profile.py 4096 reference_a1 2 2>&1 | tail -1 | cut -c1-120
done
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
LMPC_TILE_NW=2 python -m pytest tests/test_gpu_parity.py tests/test_gpu_warm_rollout.py -m gpu -q -x 2>&1 | tail -2
LMPC_TILE_NW=4 python -m pytest tests/test_gpu_parity.py tests/test_gpu_warm_rollout.py -m gpu -q -x 2>&1 | tail -2
