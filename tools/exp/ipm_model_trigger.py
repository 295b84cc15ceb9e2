import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from ipm_model_tapia import problems
from legged_mpc_control_b200 import abi

def trace(H, g, A, d, u0, act_ref, zscale=1e-4, tol=1e-7, max_iter=60):
    n, m = H.shape[0], A.shape[0]
    u = u0.copy(); s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    z = zscale * gi * s.mean() / s
    faces = []; mus = []
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        faces.append(F.face_of(s, z)); mus.append(mub)
        if np.abs(rd).max() <= 1e-9 * max(1, np.abs(g).max()) and mub <= tol:
            return it, faces, mus
        w = z / s
        M = H + A.T @ (w[:, None] * A)
        du = F.chol_solve(M, -(H @ u + g), False)
        dsa = -A @ du; dza = -z - w * dsa
        al = min(1.0, 1.0 / max(1e-300, max((-dsa / s).max(), (-dza / z).max())))
        sig = (((s + al * dsa) @ (z + al * dza)) / (s @ z)) ** 3
        rc = sig * mub - dsa * dza - s * z
        du = F.chol_solve(M, -rd - A.T @ (rc / s), False)
        ds = -A @ du; dz = rc / s - w * ds
        tau = max((-ds / s).max(), (-dz / z).max())
        al = min(1.0, 0.999 / tau) if tau > 0 else 1.0
        u = u + al * du; z = z + al * dz
    return max_iter, faces, mus

for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 96, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 8, 6), ("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), 12, 2)):
    rows = []
    for H, g, A, d, uc, act in problems(name, robot, gaits, nprob, seed):
        it, faces, mus = trace(H, g, A, d, uc, act)
        rows.append((it, faces, mus, act))
    print(name, "iterations to mu<=1e-7:", np.mean([r[0] for r in rows]))
    for mutrig in (1e-3, 1e-4, 1e-5, 1e-6):
        saved = []; good = 0; tried = 0
        for it, faces, mus, act in rows:
            k = next((j for j in range(1, len(faces)) if mus[j] <= mutrig and np.array_equal(faces[j], faces[j - 1])), None)
            if k is None or k >= it: saved.append(0.0); continue
            tried += 1
            if np.array_equal(faces[k], act): good += 1; saved.append(it - k)
            else: saved.append(-1.2)   # a wasted polish round, then on to the normal hand-over
        print(f"   trigger: face unchanged since the last iteration and mu <= {mutrig:g}: tried on {tried}/{len(rows)}, right face {good}/{tried}, mean iterations saved {np.mean(saved):+.2f}")
