import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from ipm_model_tapia import problems
from legged_mpc_control_b200 import abi

def final_iterate(H, g, A, d, u0, zscale=1e-4, tol=1e-7, frac=0.999, max_iter=60):
    n, m = H.shape[0], A.shape[0]
    u = u0.copy(); s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    z = zscale * gi * s.mean() / s
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        if np.abs(rd).max() <= 1e-7 * max(1, np.abs(g).max()) and mub <= tol:
            return s, z, mub
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
        al = min(1.0, frac / tau) if tau > 0 else 1.0
        u = u + al * du; z = z + al * dz
    return d - A @ u, z, mub

rules = {
  "kernel: s < z*max(s)/max(z)": lambda s, z, mu: s < z * (s.max() / z.max()),
  "s < z*mean(s)/mean(z)": lambda s, z, mu: s < z * (s.mean() / z.mean()),
  "s/max(s) < sqrt(mu_rel)": lambda s, z, mu: s / s.max() < np.sqrt(mu / (s.max() * z.max())),
  "s < z*sqrt(smax*smin/(zmax*zmin))": lambda s, z, mu: s < z * np.sqrt(s.max() * s.min() / (z.max() * z.min())),
}
for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 96, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 10, 6), ("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), 12, 2)):
    P = list(problems(name, robot, gaits, nprob, seed))
    fin = [(final_iterate(H, g, A, d, uc), act) for H, g, A, d, uc, act in P]
    for rn, rf in rules.items():
        ok = np.mean([np.array_equal(rf(s, z, mu), act) for (s, z, mu), act in fin])
        wrong = np.mean([(rf(s, z, mu) != act).sum() for (s, z, mu), act in fin])
        print(f"{name:20s} {rn:40s} exact face {100*ok:5.1f} %   mean wrong rows {wrong:.2f}")
