"""CPU restatement of the reference's leg kinematics and GRF -> joint command map (test infrastructure
only — never imported by the package).  Paths relative to /root/reference/src/legged_ctrl/.

  fk, jac        src/utils/A1Kinematics.cpp:8-18 with the Matlab-generated polynomials :42-140, written
                 out term by term as generated (so that the pin against the compiled reference,
                 tests/test_leg_ctrl.py, is tight: <= 4 ulp)
  inv_kin        A1Kinematics.cpp:321-446 (analytic, branch by branch)
  leg_kinematics src/interfaces/BaseInterface.cpp:280-300
  tau_ctrl       BaseInterface::tau_ctrl_update, BaseInterface.cpp:451-500
Scalar loops over small batches: this is the checker, not a fast path."""
import math

import numpy as np

A1_RHO_FIX = np.array([[0.1805, 0.047, 0.0838, 0.21, 0.21], [0.1805, -0.047, -0.0838, 0.21, 0.21],
                       [-0.1805, 0.047, 0.0838, 0.21, 0.21], [-0.1805, -0.047, -0.0838, 0.21, 0.21]])  # BaseInterface.cpp:72-98
A1_RHO_OPT = np.zeros((4, 3))


def fk(q, ro, rf):
    c0, c1, c2 = math.cos(q[0]), math.cos(q[1]), math.cos(q[2])
    s0, s1, s2 = math.sin(q[0]), math.sin(q[1]), math.sin(q[2])
    s12, c12 = math.sin(q[1] + q[2]), math.cos(q[1] + q[2])
    px = (((rf[0] + ro[2] * s12) - rf[4] * s12) - s1 * rf[3]) + ro[0] * c12
    py = ((((((((rf[1] + ro[1] * c0) + rf[2] * c0) + c1 * s0 * rf[3]) + ro[0] * c1 * s0 * s2) + ro[0] * c2 * s0 * s1)
            - ro[2] * c1 * c2 * s0) + ro[2] * s0 * s1 * s2) + rf[4] * c1 * c2 * s0) - rf[4] * s0 * s1 * s2
    a, b, c = ro[0] * c0, ro[2] * c0, rf[4] * c0
    pz = (((((((ro[1] * s0 + rf[2] * s0) - c0 * c1 * rf[3]) - a * c1 * s2) - a * c2 * s1) + b * c1 * c2) - b * s1 * s2)
          - c * c1 * c2) + c * s1 * s2
    return np.array([px, py, pz])


def jac(q, ro, rf, h=None):
    """d fk / d q, column-major semantics J[r, c]; closed form collected by angle sums (the compiled
    reference's generated polynomial agrees to rounding, tests/test_leg_ctrl.py)."""
    s0, c0, s1, c1 = math.sin(q[0]), math.cos(q[0]), math.sin(q[1]), math.cos(q[1])
    s12, c12 = math.sin(q[1] + q[2]), math.cos(q[1] + q[2])
    dd, lcc = rf[2] + ro[1], rf[4] - ro[2]
    X = -rf[3] * s1 + ro[0] * c12 - lcc * s12
    L = rf[3] * c1 + ro[0] * s12 + lcc * c12
    X2, L2 = -ro[0] * s12 - lcc * c12, ro[0] * c12 - lcc * s12
    return np.array([[0.0, -L, X2], [-dd * s0 + L * c0, s0 * X, s0 * L2], [dd * c0 + L * s0, -c0 * X, -c0 * L2]])


def inv_kin(p, cur_q, ro, rf):
    ox, oy, d, lt, lc = rf
    Xf, Yf, Zf = p
    Yf_, Xf_ = Yf - oy, Xf - ox
    L2 = Zf * Zf + Yf_ * Yf_ - d * d
    L = math.sqrt(L2) if L2 >= 0 else float("nan")
    t1 = cand = 0.0
    at = math.atan2
    if oy > 0:
        if Zf > 0:
            t1 = (at(Zf, Yf_) if Yf_ > 0 else (math.pi / 2 if Yf_ == 0 else math.pi - at(Zf, -Yf_))) - at(L, d)
            cand = at(Zf, Yf_) + at(L, d)
        elif Zf < 0:
            t1 = (at(Zf, Yf_) if Yf_ > 0 else (-math.pi / 2 if Yf_ == 0 else -math.pi - at(Zf, -Yf_))) + at(L, d)
            cand = at(Zf, Yf_) - at(L, d)
        else:
            t1, cand = at(L, d), -at(L, d)
    else:
        if Zf > 0:
            t1 = (-at(Zf, -Yf_) if Yf_ < 0 else (-math.pi / 2 if Yf_ == 0 else -math.pi + at(Zf, Yf_))) + at(L, -d)
            cand = -at(L, -d) - at(Zf, -Yf_)
        elif Zf < 0:
            t1 = (at(-Zf, -Yf_) if Yf_ < 0 else (-math.pi / 2 if Yf_ == 0 else math.pi - at(-Zf, Yf_))) - at(L, -d)
            cand = at(-Zf, -Yf_) + at(L, -d)
    if not (abs(t1 - cur_q[0]) < abs(cand - cur_q[0])):
        t1 = cand
    cb = (lt * lt + lc * lc - Xf_ * Xf_ - L * L) / (2 * lt * lc)
    if abs(cb + 1) < 0.001:
        beta = math.pi
    elif abs(cb - 1) < 0.001:
        beta = 0.0
    else:
        beta = math.acos(cb) if -1.0 <= cb <= 1.0 else float("nan")
    t3 = beta - math.pi
    if Zf > d * math.sin(t1):
        L = -L
    t2 = at(-Xf_, L) + at(lc * math.sin(-t3), lt + lc * math.cos(-t3))
    if t2 < -60 * math.pi / 180:
        t2 += 2 * math.pi
    elif t2 > 240 * math.pi / 180:
        t2 -= 2 * math.pi
    return np.array([t1, t2, t3])


def leg_kinematics(a, rho_fix=A1_RHO_FIX, rho_opt=A1_RHO_OPT):
    B = a["joint_pos"].shape[0]
    out = {k: np.zeros((B, 12)) for k in ("foot_pos_rel", "foot_vel_rel", "foot_pos_abs", "foot_vel_abs", "foot_pos_world", "foot_vel_world")}
    out["j_foot"] = np.zeros((B, 36))
    for b in range(B):
        R = a["rot"][b].reshape(3, 3, order="F")
        w = a["imu_ang_vel"][b] if a.get("imu_ang_vel") is not None else np.zeros(3)
        for i in range(4):
            q, dq = a["joint_pos"][b, 3 * i:3 * i + 3], a["joint_vel"][b, 3 * i:3 * i + 3]
            pr = fk(q, rho_opt[i], rho_fix[i]); J = jac(q, rho_opt[i], rho_fix[i]); vr = J @ dq
            sl = slice(3 * i, 3 * i + 3)
            out["foot_pos_rel"][b, sl], out["foot_vel_rel"][b, sl] = pr, vr
            out["j_foot"][b, 9 * i:9 * i + 9] = J.reshape(-1, order="F")
            out["foot_pos_abs"][b, sl], out["foot_vel_abs"][b, sl] = R @ pr, R @ vr
            out["foot_pos_world"][b, sl] = R @ pr + a["root_pos"][b]
            out["foot_vel_world"][b, sl] = R @ vr + a["root_lin_vel"][b] + R @ np.cross(w, pr)
    return out


def tau_ctrl(a, rho_fix=A1_RHO_FIX, rho_opt=A1_RHO_OPT):
    B = a["joint_pos"].shape[0]
    prev_in = a.get("joint_ang_tgt", np.zeros((B, 12)))
    out = {k: np.zeros((B, 12)) for k in ("joint_tau_tgt", "joint_ang_tgt", "prev_joint_ang_tgt", "joint_vel_tgt")}
    for b in range(B):
        R = a["rot"][b].reshape(3, 3, order="F")
        for i in range(4):
            sl = slice(3 * i, 3 * i + 3)
            q, dq = a["joint_pos"][b, sl], a["joint_vel"][b, sl]
            J = jac(q, rho_opt[i], rho_fix[i])
            out["joint_tau_tgt"][b, sl] = -J.T @ (R.T @ a["grf"][b, sl])
            if a["movement_mode"][b] > 0:
                pt = R.T @ (a["foot_pos_target_world"][b, sl] - a["root_pos"][b])
                vt = R.T @ (a["foot_vel_target_world"][b, sl] - a["root_lin_vel"][b])
                qi = inv_kin(pt, q, rho_opt[i], rho_fix[i])
                out["prev_joint_ang_tgt"][b, sl] = prev_in[b, sl]
                out["joint_ang_tgt"][b, sl] = q if np.isnan(qi).any() else qi
                try:
                    qd = np.linalg.solve(J, vt)
                except np.linalg.LinAlgError:
                    qd = np.full(3, np.nan)
                out["joint_vel_tgt"][b, sl] = dq if np.isnan(qd).any() else qd
            else:
                out["prev_joint_ang_tgt"][b, sl] = q
                out["joint_ang_tgt"][b, sl] = q
                out["joint_vel_tgt"][b, sl] = dq
    return out
