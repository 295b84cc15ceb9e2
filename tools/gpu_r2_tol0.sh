# Sweep of the interior point's hand-over tolerances (LMPC_TOL0, LMPC_RDTOL0; read at lmpc_create).
for t in 1e-7 1e-6 1e-5; do for r in 1e-9 1e-7; do
  export LMPC_TOL0=$t LMPC_RDTOL0=$r
  for wl in headline c3; do QB_TAG="tol0 $t rdtol0 $r" timeout 300 python tools/dbg/qb.py $wl 3 2>&1 | tail -1 | cut -c1-220; done
  timeout 200 python tools_profile.py 4096 reference_a1 3 trot 2>&1 | tail -1 | sed -E "s/'flops_executed': [0-9.e+]+//" | cut -c1-250
done; done
