import sys, os, numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
from legged_mpc_control_b200 import abi, presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine
from oracle import binding as ob
from parity_util import elem_err, norm_err, active_mismatch
p = presets.preset("northstar_a1")
B = 65536
b = synth.make_batch(B, p.horizon, p.dt, seed=3)
ref = ob.solve_batch(p, b, eliminate_swing=True, want_kkt=True)
for dr in sys.argv[1:]:
    os.environ["LMPC_DUAL_REL"] = dr
    eng = ConvexMpcEngine(p, B)
    got = eng.compute_grfs(b); got = eng.compute_grfs(b)
    print("=== dual_rel", dr, "kernel ms", eng.last_kernel_ms, eng.last_stats(), "status", np.bincount(got["status"], minlength=4))
    for k in ("grf", "U"):
        ee = elem_err(got[k], ref[k]); ne = norm_err(got[k], ref[k])
        print(k, "elem q", np.quantile(ee, [0.5, 0.99, 0.9999, 1.0]), "norm q", np.quantile(ne, [0.5, 0.99, 0.9999, 1.0]))
    print("active", active_mismatch(p, got["active_set"], ref["active_set"], ref["U"]))
    eng.close()
