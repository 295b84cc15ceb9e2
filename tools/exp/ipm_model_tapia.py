import sys, numpy as np
import os; _here = os.path.dirname(os.path.abspath(__file__)); sys.path.insert(0, _here); sys.path.insert(0, os.path.dirname(os.path.dirname(_here)))
import fp32_ipm_numerics as F
from legged_mpc_control_b200 import synth, abi, engine
from oracle import binding as ob

def run(H, g, A, d, u0, act_ref, zscale=1e-4, tol=1e-7, sigpow=3, frac=0.999, max_iter=60):
    n, m = H.shape[0], A.shape[0]
    u = u0.copy(); s = d - A @ u
    gi = max(np.abs(H @ u + g).max(), 1e-6)
    z = zscale * gi * s.mean() / s
    hist = []   # per iteration: (mu, wrong_rule, wrong_tapia)
    prev_s = None; prev_z = None
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        f_rule = F.face_of(s, z)
        if prev_s is not None:
            f_tap = (s / prev_s < 0.5) & (z / prev_z > 0.5)   # Tapia: slack collapsing, multiplier staying
        else:
            f_tap = np.zeros(m, bool)
        hist.append((mub, int((f_rule != act_ref).sum()), int((f_tap != act_ref).sum())))
        if np.abs(rd).max() <= 1e-9 * max(1, np.abs(g).max()) and mub <= tol:
            return it, hist
        prev_s, prev_z = s.copy(), z.copy()
        w = z / s
        M = H + A.T @ (w[:, None] * A)
        du = F.chol_solve(M, -(H @ u + g), False)
        dsa = -A @ du; dza = -z - w * dsa
        al = min(1.0, 1.0 / max(1e-300, max((-dsa / s).max(), (-dza / z).max())))
        sig = (((s + al * dsa) @ (z + al * dza)) / (s @ z)) ** sigpow
        rc = sig * mub - dsa * dza - s * z
        du = F.chol_solve(M, -rd - A.T @ (rc / s), False)
        ds = -A @ du; dz = rc / s - w * ds
        tau = max((-ds / s).max(), (-dz / z).max())
        al = min(1.0, frac / tau) if tau > 0 else 1.0
        u = u + al * du; z = z + al * dz
    return max_iter, hist

def problems(name, robot, gaits, nprob, seed):
    p = engine.c_preset(name)
    b = synth.make_batch(nprob, p.horizon, p.dt, seed, robot=robot, gaits=gaits)
    for i in range(nprob):
        H, g, keep = F.reduced_qp(p, b, i)
        A, d = F.constraints(H.shape[0] // 3, p.mu, p.fz_max)
        ref = ob.solve_one(p, b, i, eliminate_swing=True)["U"][keep]
        act = (d - A @ ref) <= 1e-7 * max(np.abs(ref).max(), 1)
        sc = 1.0 / np.abs(np.diag(H)).max(); H = H * sc; g = g * sc
        c = b["contact"][i].reshape(p.horizon, 4)
        cs_ = (c != 0).sum(1).astype(np.int64); ns = np.repeat(cs_, cs_)
        fz = np.clip(p.mass * 9.8 / np.maximum(ns, 1), 0.05 * p.fz_max, 0.9 * p.fz_max)
        uc = np.zeros(H.shape[0]); uc[2::3] = fz
        yield H, g, A, d, uc, act

if __name__ == "__main__":
    for name, robot, gaits, nprob, seed in (("northstar_a1", "a1", (abi.GAIT_TROT,), 64, 3), ("reference_a1", "a1", (abi.GAIT_TROT,), 6, 6)):
        its = []; firstrule = []; firsttap = []
        P = list(problems(name, robot, gaits, nprob, seed))
        for H, g, A, d, uc, act in P:
            it, hist = run(H, g, A, d, uc, act)
            its.append(it)
            def first_stable(col):
                k = len(hist)
                for j in range(len(hist) - 1, -1, -1):
                    if hist[j][col] == 0: k = j
                    else: break
                return k
            firstrule.append(first_stable(1)); firsttap.append(first_stable(2))
        print(name, "iterations", np.mean(its), "face correct (kernel rule) from", np.mean(firstrule), " (Tapia) from", np.mean(firsttap))
        for sp in (2, 3, 4):
            for fr in (0.99, 0.999, 0.9999):
                v = [run(H, g, A, d, uc, act, sigpow=sp, frac=fr)[0] for H, g, A, d, uc, act in P]
                print(f"   sigma^{sp} frac {fr}: {np.mean(v):.2f} (max {np.max(v)})", end="")
            print()
