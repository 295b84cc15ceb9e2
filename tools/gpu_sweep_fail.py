"""Find and dissect problems of the config-4 sweep that do not reach status 0/1."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from legged_mpc_control_b200 import abi, presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine
from oracle import binding as ob

p = presets.preset("northstar_a1")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
b = synth.make_batch(B, p.horizon, p.dt, seed=4, gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK),
                     per_problem_params=True, base_params=p, gait_speed=(2.0, 4.5))
eng = ConvexMpcEngine(p, B)
out = eng.compute_grfs(b)
print("status histogram", np.bincount(out["status"]), "max fact", out["factorizations"].max())
bad = np.nonzero(out["status"] > 0)[0]
print("not exact:", bad[:20])
for i in bad[:5]:
    one = {k: (v[i:i + 1] if isinstance(v, np.ndarray) and v.shape[0] == B else v) for k, v in b.items()}
    ref = ob.solve_batch(p, one, eliminate_swing=True)
    print(i, "status", out["status"][i], "fact", out["factorizations"][i], "polish", out["polish_rounds"][i],
          "oracle fails", ref["failures"], "kkt", ref["kkt"][0])
    print("  mu", one["mu"], "q", one["q"][0], "r", one["r"][0])
    print("  oracle grf", ref["grf"][0]); print("  device grf", out["grf"][i])
    print("  x0", one["x0"][0]); print("  contact", one["contact"][0].reshape(-1, 4).T)
for it in (80, 200):
    p2 = p.copy(); p2.max_iter = it
    e2 = ConvexMpcEngine(p2, B)
    o2 = e2.compute_grfs(b)
    print("max_iter", it, "status histogram", np.bincount(o2["status"]))
    e2.close()
