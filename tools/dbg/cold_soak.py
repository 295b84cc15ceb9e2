"""Cold-start robustness of the small-complementarity start + retry pass: N distinct 65 536-problem batches of the
headline and the per-problem-parameter sweep distributions, every answer must be status <= 1.
python tools/dbg/cold_soak.py [batches per workload]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from legged_mpc_control_b200 import presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
B = 65536
p = presets.preset("northstar_a1")
eng = ConvexMpcEngine(p, B)
res = []
for wl in ("headline", "sweep"):
    tot = {"workload": wl, "problems": 0, "status0": 0, "status1": 0, "failed": 0, "factorizations": 0.0}
    for k in range(nb):
        b = synth.make_batch(B, p.horizon, p.dt, seed=1000 + k, per_problem_params=(wl == "sweep"), base_params=p)
        out = eng.compute_grfs(b, want_U=False, copy=False)
        st = np.asarray(out["status"])
        tot["problems"] += B; tot["status0"] += int((st == 0).sum()); tot["status1"] += int((st == 1).sum()); tot["failed"] += int((st > 1).sum())
        tot["factorizations"] += float(eng.last_stats()["mean_factorizations"]) / nb
    res.append(tot)
print(json.dumps(res))
