import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from ipm_model_tapia import problems
from legged_mpc_control_b200 import abi

def run1(H, g, A, d, u0, mode, zscale=1e-4, tol=1e-7, frac=0.999, max_iter=80):
    """mode: 'mehrotra' (2 solves/iter), ('fixed', sigma), 'lagged' (sigma from last step's achieved reduction, 1 solve/iter)"""
    n, m = H.shape[0], A.shape[0]
    u = u0.copy(); s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    z = zscale * gi * s.mean() / s
    solves = 0; sig_prev = 0.3
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        if np.abs(rd).max() <= 1e-7 * max(1, np.abs(g).max()) and mub <= tol:
            return it, solves
        w = z / s
        M = H + A.T @ (w[:, None] * A)
        if mode == 'mehrotra':
            du = F.chol_solve(M, -(H @ u + g), False); solves += 1
            dsa = -A @ du; dza = -z - w * dsa
            al = min(1.0, 1.0 / max(1e-300, max((-dsa / s).max(), (-dza / z).max())))
            sig = (((s + al * dsa) @ (z + al * dza)) / (s @ z)) ** 3
            rc = sig * mub - dsa * dza - s * z
        else:
            sig = mode[1] if mode[0] == 'fixed' else sig_prev
            rc = sig * mub - s * z
        du = F.chol_solve(M, -rd - A.T @ (rc / s), False); solves += 1
        ds = -A @ du; dz = rc / s - w * ds
        tau = max((-ds / s).max(), (-dz / z).max())
        al = min(1.0, frac / tau) if tau > 0 else 1.0
        u = u + al * du; z = z + al * dz
        if mode[0] == 'lagged':
            mu_new = ((d - A @ u) @ z) / m
            sig_prev = float(np.clip((mu_new / mub) ** 2 if al > 0.9 else 0.5, 1e-3, 0.8))
    return max_iter, solves

for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 48, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 5, 6)):
    P = list(problems(name, robot, gaits, nprob, seed))
    for mode in ('mehrotra', ('fixed', 0.3), ('fixed', 0.1), ('fixed', 0.03), ('lagged',)):
        r = [run1(H, g, A, d, uc, mode) for H, g, A, d, uc, act in P]
        print(name, mode, "iterations %.2f  solves %.2f  max it %d" % (np.mean([x[0] for x in r]), np.mean([x[1] for x in r]), max(x[0] for x in r)))
