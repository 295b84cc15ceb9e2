import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from legged_mpc_control_b200 import synth, abi, engine
from oracle import binding as ob

def ipm(H, g, A, d, u0, zmode, tol=1e-7, max_iter=60, frac=0.999):
    n, m = H.shape[0], A.shape[0]
    u = u0.copy()
    s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    if zmode == "kernel":
        z = gi * s.mean() / s
    else:  # smaller complementarity: mu0 = zmode * gi * mean(s)
        z = zmode * gi * s.mean() / s
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        if np.abs(rd).max() <= 1e-9 * max(1, np.abs(g).max()) and mub <= tol:
            return it
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
    return max_iter

for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 48, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 6, 6)):
    p = engine.c_preset(name)
    b = synth.make_batch(nprob, p.horizon, p.dt, seed, robot=robot, gaits=gaits)
    res = {}
    for i in range(nprob):
        H, g, keep = F.reduced_qp(p, b, i)
        A, d = F.constraints(H.shape[0] // 3, p.mu, p.fz_max)
        sc = 1.0 / np.abs(np.diag(H)).max(); H = H * sc; g = g * sc
        n = H.shape[0]
        uc = np.zeros(n); uc[2::3] = 0.25 * 12.0 * 9.8 / 2.0
        uu = np.linalg.solve(H, -g)
        def toward(theta_frac):
            dirn = uu - uc
            s0 = d - A @ uc; ds = -A @ dirn
            neg = ds < 0
            amax = min(1.0, (s0[neg] / -ds[neg]).min()) if neg.any() else 1.0
            return uc + theta_frac * amax * dirn
        for tag, u0, zm in (("kernel", uc, "kernel"), ("z x0.1", uc, 0.1), ("z x0.01", uc, 0.01),
                            ("unc 0.5", toward(0.5), "kernel"), ("unc 0.9", toward(0.9), "kernel"), ("unc 0.9 z0.1", toward(0.9), 0.1), ("unc 0.99 z0.1", toward(0.99), 0.1)):
            res.setdefault(tag, []).append(ipm(H, g, A, d, u0, zm))
    print(name, {k: round(float(np.mean(v)), 2) for k, v in res.items()})

print("--- z0 scale scan")
for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 64, 3), ("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), 12, 2), ("reference_a1", "a1", (abi.GAIT_TROT,), 6, 6)):
    p = engine.c_preset(name)
    b = synth.make_batch(nprob, p.horizon, p.dt, seed, robot=robot, gaits=gaits)
    res = {}
    for i in range(nprob):
        H, g, keep = F.reduced_qp(p, b, i)
        A, d = F.constraints(H.shape[0] // 3, p.mu, p.fz_max)
        sc = 1.0 / np.abs(np.diag(H)).max(); H = H * sc; g = g * sc
        n = H.shape[0]
        c = b["contact"][i].reshape(p.horizon, 4)
        cs_ = (c != 0).sum(1).astype(np.int64); ns = np.repeat(cs_, cs_)  # stance legs of the block's step, per block
        fz = np.clip(p.mass * 9.8 / np.maximum(ns, 1), 0.05 * p.fz_max, 0.9 * p.fz_max)
        uc = np.zeros(n); uc[2::3] = fz
        for zm in (1.0, 0.3, 0.1, 0.03, 0.01, 3e-3, 1e-3, 1e-4, 1e-5):
            res.setdefault(zm, []).append(ipm(H, g, A, d, uc, zm))
    print(name, {k: (round(float(np.mean(v)), 2), int(np.max(v))) for k, v in res.items()})
