# A/B of library builds on the one-warp class (headline, sweep): parity subset + device-timed throughput.
# Usage: bash tools/gpu_r2_ab_warp.sh "<lib suffixes, '' = liblmpc.so>"
set -x
L=legged_mpc_control_b200
for sfx in "$@"; do
  export LMPC_LIB_PATH=$PWD/$L/liblmpc$sfx.so
  echo "=== lib $LMPC_LIB_PATH"
  timeout 900 python -m pytest tests/test_gpu_parity_fullsize.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x -k "headline or c2 or c1 or sweep or alloc or standing" 2>&1 | tail -2
  for wl in headline sweep; do timeout 200 python tools/dbg/qb.py $wl 8 2>&1 | tail -1 | cut -c1-200; done
  timeout 200 python tools/dbg/qb.py headline 8 2>&1 | tail -1 | cut -c1-200
done
