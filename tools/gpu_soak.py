"""Closed-loop soak: long rollouts over every gait, both presets; counts failed solves and checks
that no robot falls or goes non-finite.  python tools/gpu_soak.py > gpurun_out/soak.json"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from legged_mpc_control_b200 import abi, presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine

res = []
for name, B, steps, gaits in (("northstar_a1", 16384, 5000, (abi.GAIT_TROT, abi.GAIT_PRONK, abi.GAIT_CRAWL, abi.GAIT_TROT_WITH_STAND)),
                              # stand: with the yaml weights (no position / yaw weight) fixed feet make the SRBD loop a slow inverted
                              # pendulum (the numpy twin shows the same growth); bound at H=10 is beyond this controller: short runs only
                              ("northstar_a1", 4096, 300, (abi.GAIT_STAND, abi.GAIT_BOUND)),
                              ("reference_a1", 4096, 2000, (abi.GAIT_TROT,)),
                              ("northstar_go1_h20", 8192, 2000, (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK))):
    p = presets.preset(name)
    st = synth.make_rollout_state(B, seed=77, gaits=gaits, robot="go1" if "go1" in name else "a1")
    fr = st.pop("foot_rel")
    eng = ConvexMpcEngine(p, B)
    tot = {"solves": 0, "not_converged": 0, "warm_hits": 0, "factorizations": 0, "ms": 0.0}
    first = True
    for c in range(0, steps, 500):
        o = eng.rollout(st, min(500, steps - c), fr, warm_start=1 if first else 2)
        first = False
        for k in ("solves", "not_converged", "warm_hits", "factorizations"):
            tot[k] += o[k]
        tot["ms"] += o["device_ms"]
    x = st["x"]
    by_gait = {}
    for g in np.unique(st["gait"]):
        m = st["gait"] == g
        by_gait[int(g)] = {"height_err_max": float(np.abs(x[m, 5] - st["cmd"][m, 2]).max()), "tilt_max": float(np.abs(x[m, 0:2]).max()),
                           "vel_err_max": float(np.abs(x[m, 9:11]).max())}
    res.append({"preset": name, "robots": B, "steps": steps, "finite": bool(np.isfinite(x).all()), **tot,
                "robot_steps_per_s": tot["solves"] / (tot["ms"] * 1e-3), "mean_factorizations": tot["factorizations"] / tot["solves"],
                "warm_hit_fraction": tot["warm_hits"] / tot["solves"], "by_gait": by_gait})
    eng.close()
print(json.dumps(res))
