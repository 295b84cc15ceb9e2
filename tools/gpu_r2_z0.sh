# Sweep of the interior point's initial complementarity (LMPC_Z0_SCALE, read at lmpc_create) on every workload.
# Usage: bash tools/gpu_r2_z0.sh "<scales>"
for z in "$@"; do
  export LMPC_Z0_SCALE=$z
  for wl in headline sweep c3; do QB_TAG="z0=$z" timeout 300 python tools/dbg/qb.py $wl 4 2>&1 | tail -1 | cut -c1-200; done
  QB_TAG="z0=$z" timeout 300 python tools/dbg/qb.py ref_h30 2 2>&1 | tail -1 | cut -c1-200
done
