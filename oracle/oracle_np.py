"""numpy twin of the oracle's matrix construction + KKT certificates.  TEST INFRASTRUCTURE ONLY.

Written independently of lmpc_oracle.c (different language, vectorised, closed-form where the C
code is generic) so that the two restatements check each other; neither is the product.
Citations are to /root/reference/src/legged_ctrl/src/mpc_ctrl/convex_mpc/ConvexQPSolver.cpp.
PARITY UNPINNED (see oracle/lmpc_oracle.h).
"""
import numpy as np
from scipy.linalg import expm
from scipy.optimize import lsq_linear, nnls

from legged_mpc_control_b200 import abi

INF = 1e30


def _params(p):
    return dict(H=p.horizon, dt=p.dt, mu=p.mu, fz=p.fz_max, m=p.mass, g=p.gravity,
                I=np.array(list(p.inertia)).reshape(3, 3, order="F"), q=np.array(list(p.q)[:12]),
                r=np.array(list(p.r)), zoh=p.discretisation == abi.DISC_ZOH,
                cs=2.0 if p.cost_scaling == abi.COST_NS_TWO else 1.0, off=p.xref_index_offset)


def skew(v):  # Utils::skew, utils/Utils.cpp:89-95
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


def continuous(pp, rot, foot, yaw):
    """Ac (:214-225), Bc (:198-208), affine gravity c."""
    R = np.asarray(rot).reshape(3, 3, order="F")
    Ac = np.zeros((12, 12)); Bc = np.zeros((12, 12)); c = np.zeros(12)
    cy, sy = np.cos(yaw), np.sin(yaw)
    Ac[0:3, 6:9] = [[cy, sy, 0], [-sy, cy, 0], [0, 0, 1]]
    Ac[3:6, 9:12] = np.eye(3)
    Iw_inv = np.linalg.inv(R @ pp["I"] @ R.T)
    feet = np.asarray(foot).reshape(4, 3)
    for leg in range(4):
        Bc[6:9, 3 * leg:3 * leg + 3] = Iw_inv @ skew(feet[leg])
        Bc[9:12, 3 * leg:3 * leg + 3] = np.eye(3) / pp["m"]
    c[11] = -pp["g"]
    return Ac, Bc, c


def discretise(pp, Ac, Bc, c):
    dt = pp["dt"]
    if not pp["zoh"]:
        return np.eye(12) + Ac * dt, Bc * dt, c * dt  # :227, :211, :175
    M = np.zeros((25, 25))
    M[:12, :12] = Ac; M[:12, 12:24] = Bc; M[:12, 24] = c
    E = expm(M * dt)
    return E[:12, :12], E[:12, 12:24], E[:12, 24]


def xref_traj(pp, x0, rot, cmd):
    R = np.asarray(rot).reshape(3, 3, order="F")
    vw = R @ np.asarray(cmd[3:6])  # :260
    H = pp["H"]
    xr = np.zeros((H, 12))
    for i in range(H):
        t = pp["dt"] * (i + pp["off"])
        xr[i] = [cmd[0], cmd[1], x0[2] + cmd[8] * t, x0[3] + vw[0] * t, x0[4] + vw[1] * t, cmd[2],
                 cmd[6], cmd[7], cmd[8], vw[0], vw[1], 0.0]  # :264-276
    return xr


def sparse_qp(p, x0, rot, foot, cmd, contact):
    """Reference QP (P diag, q, A, l, u), z = [u0;x1;u1;x2;...]  (:16-196, :254-346)."""
    pp = _params(p); H = pp["H"]
    nz, nd, nf = 24 * H, 12 * H, 16 * H
    P = np.zeros(nz); q = np.zeros(nz); A = np.zeros((32 * H, nz)); l = np.zeros(32 * H); u = np.zeros(32 * H)
    xr = xref_traj(pp, x0, rot, cmd)
    contact = np.asarray(contact).reshape(H, 4)
    for i in range(H):
        P[24 * i:24 * i + 12] = pp["cs"] * pp["r"]; P[24 * i + 12:24 * i + 24] = pp["cs"] * pp["q"]
        Ad, Bd, cd = discretise(pp, *continuous(pp, rot, foot, xr[i, 2]))
        A[12 * i:12 * i + 12, 24 * i:24 * i + 12] = Bd
        A[12 * i:12 * i + 12, 24 * i + 12:24 * i + 24] = -np.eye(12)
        if i == 0:
            l[0:12] = u[0:12] = -Ad @ x0 - cd  # :294-297
        else:
            A[12 * i:12 * i + 12, 24 * (i - 1) + 12:24 * (i - 1) + 24] = Ad
            l[12 * i:12 * i + 12] = u[12 * i:12 * i + 12] = -cd  # :175-176
        q[24 * i + 12:24 * i + 24] = -pp["cs"] * pp["q"] * xr[i]  # :308
        for j in range(4):
            fr = nd + 16 * i + 4 * j; cx = 24 * i + 3 * j
            A[fr + 0, cx] = 1; A[fr + 0, cx + 2] = pp["mu"]; l[fr + 0], u[fr + 0] = 0, INF
            A[fr + 1, cx] = 1; A[fr + 1, cx + 2] = -pp["mu"]; l[fr + 1], u[fr + 1] = -INF, 0
            A[fr + 2, cx + 1] = 1; A[fr + 2, cx + 2] = pp["mu"]; l[fr + 2], u[fr + 2] = 0, INF
            A[fr + 3, cx + 1] = 1; A[fr + 3, cx + 2] = -pp["mu"]; l[fr + 3], u[fr + 3] = -INF, 0
            br = nd + nf + 4 * i + j
            A[br, cx + 2] = 1; l[br] = 0; u[br] = float(bool(contact[i, j])) * pp["fz"]  # :336,341
    return P, q, A, l, u


def condensed_qp(p, x0, rot, foot, cmd):
    """Condensed (Hm, g, e): U-only QP obtained by eliminating x with the dynamics rows."""
    pp = _params(p); H = pp["H"]; n = 12 * H
    xr = xref_traj(pp, x0, rot, cmd)
    mats = [discretise(pp, *continuous(pp, rot, foot, xr[k, 2])) for k in range(H)]
    Bq = np.zeros((n, n)); e = np.zeros((H, 12))
    xf = np.asarray(x0, float).copy()
    for k in range(H):
        Ad, Bd, cd = mats[k]
        xf = Ad @ xf + cd
        e[k] = xf - xr[k]
        Bq[12 * k:12 * k + 12, 12 * k:12 * k + 12] = Bd
        if k > 0:
            Bq[12 * k:12 * k + 12, :12 * k] = Ad @ Bq[12 * (k - 1):12 * k, :12 * k]
    Qb = np.tile(pp["q"], H); Rb = np.tile(pp["r"], H)
    Hm = pp["cs"] * (Bq.T @ (Qb[:, None] * Bq) + np.diag(Rb))
    g = pp["cs"] * (Bq.T @ (Qb * e.reshape(-1)))
    return Hm, g, e, Bq


def condensed_rows(p, contact):
    """One-sided rows C U >= b in the bitmask order of include/lmpc.h (24H rows)."""
    H = p.horizon; n = 12 * H; mu = p.mu
    contact = np.asarray(contact).reshape(H, 4)
    Cm = np.zeros((24 * H, n)); b = np.zeros(24 * H)
    for k in range(H):
        for leg in range(4):
            c = 12 * k + 3 * leg; r = 16 * k + 4 * leg
            Cm[r + 0, c] = 1; Cm[r + 0, c + 2] = mu
            Cm[r + 1, c] = -1; Cm[r + 1, c + 2] = mu
            Cm[r + 2, c + 1] = 1; Cm[r + 2, c + 2] = mu
            Cm[r + 3, c + 1] = -1; Cm[r + 3, c + 2] = mu
            Cm[16 * H + 4 * k + leg, c + 2] = 1
            Cm[20 * H + 4 * k + leg, c + 2] = -1
            b[20 * H + 4 * k + leg] = -float(bool(contact[k, leg])) * p.fz_max
    return Cm, b


def kkt_certificate(Hm, g, Cm, b, U, act_tol=1e-6):
    """Optimality certificate for min 1/2 U'HU + g'U s.t. Cm U >= b, independent of any solver:
    finds lambda >= 0 supported on the primal-active rows by NNLS.  Returns dict of residuals."""
    grad = Hm @ U + g
    res = Cm @ U - b
    act = np.nonzero(np.abs(res) <= act_tol)[0]
    if len(act):
        lam, rn = nnls(Cm[act].T, grad, maxiter=50 * len(act) + 100)
    else:
        lam, rn = np.zeros(0), np.linalg.norm(grad)
    return {"stationarity": rn / max(1.0, np.abs(g).max()), "primal_violation": max(0.0, -res.min()),
            "n_active": len(act), "lambda": lam, "active": act}


def sparse_kkt_certificate(P, q, A, l, u, z, act_tol=1e-6):
    """Certificate that z solves the reference's SPARSE QP: stationarity P z + q + A'y = 0 with
    y free on equality rows, y <= 0 on rows active at their lower bound, y >= 0 at the upper."""
    Az = A @ z
    viol = max(0.0, (l - Az).max(), (Az - u).max())
    eq = np.abs(u - l) <= 0
    at_l = (np.abs(Az - l) <= act_tol) & ~eq
    at_u = (np.abs(Az - u) <= act_tol) & ~eq
    rows = np.nonzero(eq | at_l | at_u)[0]
    lo = np.where(eq[rows], -np.inf, np.where(at_u[rows] & ~at_l[rows], 0.0, -np.inf))
    hi = np.where(eq[rows], np.inf, np.where(at_l[rows] & ~at_u[rows], 0.0, np.inf))
    rhs = -(P * z + q)
    sol = lsq_linear(A[rows].T, rhs, bounds=(lo, hi), method="bvls", tol=1e-14, max_iter=2000)
    return {"stationarity": np.linalg.norm(A[rows].T @ sol.x - rhs) / max(1.0, np.abs(q).max()),
            "primal_violation": viol}


def rollout_z(p, x0, rot, foot, cmd, U):
    """z = [u0;x1;...] obtained by rolling the discrete dynamics forward with U."""
    pp = _params(p); H = pp["H"]
    xr = xref_traj(pp, x0, rot, cmd)
    z = np.zeros(24 * H); x = np.asarray(x0, float).copy()
    for k in range(H):
        Ad, Bd, cd = discretise(pp, *continuous(pp, rot, foot, xr[k, 2]))
        uk = U[12 * k:12 * k + 12]
        x = Ad @ x + Bd @ uk + cd
        z[24 * k:24 * k + 12] = uk; z[24 * k + 12:24 * k + 24] = x
    return z
