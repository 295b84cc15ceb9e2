import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from legged_mpc_control_b200 import presets, synth, abi
from legged_mpc_control_b200.engine import ConvexMpcEngine
p = presets.preset("northstar_a1"); B = 32768
b = synth.make_batch(B, p.horizon, p.dt, seed=3)
eng = ConvexMpcEngine(p, B)
eng.set_warm_start(True, 0)
for rep in range(2):
    eng.reset_warm_start()
    out = eng.compute_grfs(b, want_U=False)
    print("kernel ms", eng.last_kernel_ms, eng.last_stats(), "hits", eng.last_warm_hits, "status", np.bincount(out["status"]), "polish hist", np.bincount(out["polish_rounds"])[:14])
