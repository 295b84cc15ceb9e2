import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from ipm_model_tapia import problems
from legged_mpc_control_b200 import abi

def steplen(s, z, ds, dz, frac):
    tau = max((-ds / s).max(), (-dz / z).max())
    return min(1.0, frac / tau) if tau > 0 else 1.0

def run(H, g, A, d, u0, ncorr, zscale=1e-4, tol=1e-7, frac=0.999, max_iter=60, dalpha=0.3, bmin=0.1, bmax=10.0, gamma=0.1):
    n, m = H.shape[0], A.shape[0]
    u = u0.copy(); s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    z = zscale * gi * s.mean() / s
    solves = 0
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        if np.abs(rd).max() <= 1e-7 * max(1, np.abs(g).max()) and mub <= tol:
            return it, solves
        w = z / s
        M = H + A.T @ (w[:, None] * A)
        du = F.chol_solve(M, -(H @ u + g), False); solves += 1
        dsa = -A @ du; dza = -z - w * dsa
        al = min(1.0, 1.0 / max(1e-300, max((-dsa / s).max(), (-dza / z).max())))
        sig = (((s + al * dsa) @ (z + al * dza)) / (s @ z)) ** 3
        rc = sig * mub - dsa * dza - s * z
        du = F.chol_solve(M, -rd - A.T @ (rc / s), False); solves += 1
        ds = -A @ du; dz = rc / s - w * ds
        al = steplen(s, z, ds, dz, frac)
        mut = sig * mub
        for c in range(ncorr):
            if al >= 0.999: break
            at = min(1.0, al + dalpha)
            v = (s + at * ds) * (z + at * dz)
            vt = np.clip(v, bmin * mut, bmax * mut)
            rc2 = np.maximum(vt - v, -bmax * mut)
            duc = F.chol_solve(M, -A.T @ (rc2 / s), False); solves += 1
            dsc = -A @ duc; dzc = rc2 / s - w * dsc
            al2 = steplen(s, z, ds + dsc, dz + dzc, frac)
            if al2 >= al + gamma * dalpha:
                du, ds, dz, al = du + duc, ds + dsc, dz + dzc, al2
            else:
                break
        u = u + al * du; z = z + al * dz
    return max_iter, solves

for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 48, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 5, 6), ("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), 8, 2)):
    P = list(problems(name, robot, gaits, nprob, seed))
    for nc in (0, 1, 2):
        r = [run(H, g, A, d, uc, nc) for H, g, A, d, uc, act in P]
        print(name, "correctors", nc, "iterations %.2f  solves %.2f" % (np.mean([x[0] for x in r]), np.mean([x[1] for x in r])))
