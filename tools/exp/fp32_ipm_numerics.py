"""Numerical experiment behind DESIGN.md §6 "FP32 variant": does a Mehrotra interior point whose normal
matrix M = H + C'WC is factored in FP32 (Jacobi-scaled, residuals and iterates in FP64 — the design SURVEY
§7.3-3 prescribes) reach the accuracy the north star asks of an FP32 path (GRF within 1e-3 relative), and how
many iterations does it need compared with an FP64 factor?  CPU only (numpy + the oracle for H, g and the
exact minimiser).  python tools/exp/fp32_ipm_numerics.py [n_problems]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from legged_mpc_control_b200 import synth, abi, engine
from oracle import binding as ob


def reduced_qp(p, b, i):
    Hm, g, _ = ob.condensed_qp(p, b, i)
    c = b["contact"][i].reshape(p.horizon, 4)
    keep = np.repeat(c.reshape(-1) != 0, 3)
    return Hm[np.ix_(keep, keep)], g[keep], keep


def constraints(nb, mu, fzmax):
    A = np.zeros((5 * nb, 3 * nb)); d = np.zeros(5 * nb)
    for k in range(nb):
        r, c = 5 * k, 3 * k
        A[r, c] = 1; A[r, c + 2] = -mu            # fx - mu fz <= 0
        A[r + 1, c] = -1; A[r + 1, c + 2] = -mu   # -fx - mu fz <= 0
        A[r + 2, c + 1] = 1; A[r + 2, c + 2] = -mu
        A[r + 3, c + 1] = -1; A[r + 3, c + 2] = -mu
        A[r + 4, c + 2] = 1; d[r + 4] = fzmax      # fz <= fzmax
    return A, d


def chol_solve(M, rhs, f32, refine=0):
    dg = 1.0 / np.sqrt(np.diag(M))
    Ms = M * dg[:, None] * dg[None, :]
    if f32:
        try:
            L = np.linalg.cholesky(Ms.astype(np.float32))
        except np.linalg.LinAlgError:
            return None
        def sol(r):
            y = np.linalg.solve(L, (r * dg).astype(np.float32))      # LAPACK sgesv on a triangular matrix: FP32 substitution accuracy
            return np.linalg.solve(L.T, y).astype(np.float64) * dg
    else:
        L = np.linalg.cholesky(Ms)
        def sol(r):
            return np.linalg.solve(L.T, np.linalg.solve(L, r * dg)) * dg
    x = sol(rhs)
    for _ in range(refine):                       # FP64 residual on M itself, FP32 correction
        x = x + sol(rhs - M @ x)
    return x


def ipm(H, g, A, d, f32, refine=0, tol=1e-7, max_iter=60):
    n, m = H.shape[0], A.shape[0]
    sc = 1.0 / np.abs(np.diag(H)).max()
    H = H * sc; g = g * sc
    u = np.zeros(n); u[2::3] = 0.25 * 12.0 * 9.8 / 2.0
    s = d - A @ u
    z = max(np.abs(H @ u + g).max(), 1e-6) * s.mean() / s
    for it in range(max_iter):
        s = d - A @ u
        rd = H @ u + g + A.T @ z
        mub = s @ z / m
        if np.abs(rd).max() <= 1e-9 * max(1, np.abs(g).max()) and mub <= tol:
            return u, it, True, s, z
        w = z / s
        M = H + A.T @ (w[:, None] * A)
        du = chol_solve(M, -(H @ u + g), f32, refine)       # predictor (z eliminated)
        if du is None or not np.all(np.isfinite(du)):
            return u, it, False, s, z
        dsa = -A @ du
        dza = -z - w * dsa
        al = min(1.0, 1.0 / max(1e-300, max((-dsa / s).max(), (-dza / z).max())))
        sig = (((s + al * dsa) @ (z + al * dza)) / (s @ z)) ** 3
        rc = sig * mub - dsa * dza - s * z
        du = chol_solve(M, -rd - A.T @ (rc / s), f32, refine)  # corrector
        if du is None or not np.all(np.isfinite(du)):
            return u, it, False, s, z
        ds = -A @ du
        dz = rc / s - w * ds
        tau = max((-ds / s).max(), (-dz / z).max())
        al = min(1.0, 0.99 / tau) if tau > 0 else 1.0
        u = u + al * du; z = z + al * dz
    return u, max_iter, False, d - A @ u, z


def face_of(s, z):
    """The kernel's face rule (lmpc_tile_solve.cuh stage 4b): row active iff s < z * max(s) / max(z)."""
    return s < z * (s.max() / max(z.max(), 1e-300))


def polish(H, g, A, d, act, f32, refine):
    """Equality-constrained QP on the active rows through an orthonormal null-space basis; the reduced Hessian
    is factored in FP32 (or FP64) and the solve refined `refine` times with FP64 residuals."""
    Aa = A[act]
    if Aa.shape[0] == 0:
        Z = np.eye(H.shape[0]); p0 = np.zeros(H.shape[0])
    else:
        U_, sv, Vt = np.linalg.svd(Aa, full_matrices=True)
        r = int((sv > 1e-10 * sv.max()).sum())
        Z = Vt[r:].T
        p0 = np.linalg.lstsq(Aa, d[act], rcond=None)[0]
    Hr = Z.T @ H @ Z
    v = chol_solve(Hr, -(Z.T @ (H @ p0 + g)), f32, refine)
    if v is None:
        return None
    return p0 + Z @ v


def run(name, robot, gaits, nprob, seed):
    p = engine.c_preset(name)
    b = synth.make_batch(nprob, p.horizon, p.dt, seed, robot=robot, gaits=gaits)
    variants = (("f64", False, 0), ("f32", True, 0), ("f32+1ref", True, 1), ("f32+2ref", True, 2), ("f32+4ref", True, 4))
    rows = []
    for i in range(nprob):
        H, g, keep = reduced_qp(p, b, i)
        nb = H.shape[0] // 3
        A, d = constraints(nb, p.mu, p.fz_max)
        ref = ob.solve_one(p, b, i, eliminate_swing=True)["U"][keep]
        scale = max(np.abs(ref).max(), 1.0)
        act_ref = (d - A @ ref) <= 1e-7 * scale
        dg = 1 / np.sqrt(np.diag(H))
        res = {"n": H.shape[0], "cond_scaled": np.linalg.cond(H * dg[:, None] * dg[None, :])}
        for tag, f32, refine in variants:
            u, it, ok, sl, z = ipm(H, g, A, d, f32, refine)
            face_ok = bool(ok) and np.array_equal(face_of(sl, z), act_ref)
            up = polish(H, g, A, d, act_ref, f32, refine)   # polish accuracy on the RIGHT face, this precision
            rel = np.inf if up is None else (np.abs(up - ref) / np.maximum(np.abs(ref), 1e-3 * scale)).max()
            res[tag] = (ok, it, face_ok, rel)
        rows.append(res)
    print(f"== {name} gaits {gaits}: {nprob} problems, n = {rows[0]['n']}, cond(Jacobi-scaled H) median {np.median([r['cond_scaled'] for r in rows]):.1e} max {np.max([r['cond_scaled'] for r in rows]):.1e}")
    for tag, _, _ in variants:
        ok = np.mean([r[tag][0] for r in rows]); it = np.mean([r[tag][1] for r in rows]); fo = np.mean([r[tag][2] for r in rows])
        rel = np.array([r[tag][3] for r in rows])
        print(f"   {tag:9s} interior point reaches mu<=1e-7, rd<=1e-9: {100*ok:5.1f} %  (mean {it:4.1f} iterations)  face read off correctly: {100*fo:5.1f} %"
              f"  | polish on the exact face, element-wise rel err vs exact: median {np.median(rel):.1e} max {rel.max():.1e}  within 1e-3: {100*np.mean(rel <= 1e-3):5.1f} %  within 1e-6: {100*np.mean(rel <= 1e-6):5.1f} %")


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    run("northstar_a1", "a1", (abi.GAIT_TROT,), n, 3)
    run("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), max(4, n // 4), 2)
    run("reference_a1", "a1", (abi.GAIT_TROT,), max(4, n // 8), 6)
