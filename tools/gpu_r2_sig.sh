# Sweep of the centering exponent and the step fraction (LMPC_SIGMA_POW, LMPC_STEP_FRAC; read at lmpc_create).
for sp in 3 2; do for fr in 0.999 0.9999; do
  export LMPC_SIGMA_POW=$sp LMPC_STEP_FRAC=$fr
  for wl in headline sweep c3; do QB_TAG="sigma^$sp frac $fr" timeout 300 python tools/dbg/qb.py $wl 4 2>&1 | tail -1 | cut -c1-220; done
  QB_TAG="sigma^$sp frac $fr" timeout 300 python tools/dbg/qb.py ref_h30 2 2>&1 | tail -1 | cut -c1-220
done; done
