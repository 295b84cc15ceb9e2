"""ctypes binding of oracle/_ref/libref_convex_mpc.so: the reference's own convex-MPC sources,
UNMODIFIED, compiled against oracle/shim (see oracle/Makefile target `ref`, oracle/ref_harness.cpp).
Test infrastructure only.  `available()` is False where neither the prebuilt library nor
/root/reference exists."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libref_convex_mpc.so")
REF_ROOT = "/root/reference/src/legged_ctrl"
_LIB = None

SOLVE_CB = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                       C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double))


def build() -> bool:
    if os.path.isdir(os.path.join(REF_ROOT, "src")):
        subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)
    return os.path.exists(_SO)


def available() -> bool:
    return os.path.exists(_SO) or build()


def lib():
    global _LIB
    if _LIB is None:
        if not available():
            raise RuntimeError("oracle/_ref is not built and /root/reference is not mounted")
        L = C.CDLL(_SO)
        L.ref_mpc_dt.restype = C.c_double
        L.ref_osqp_infty.restype = C.c_double
        L.ref_fsm_new.restype = C.c_void_p
        L.ref_fsm_new.argtypes = [C.c_int, C.c_int, C.c_double]
        L.ref_fsm_free.argtypes = [C.c_void_p]
        L.ref_fsm_reset.argtypes = [C.c_void_p]
        L.ref_fsm_set_phase.argtypes = [C.c_void_p, C.c_double]
        L.ref_fsm_predict.argtypes = [C.c_void_p, C.c_double]
        L.ref_fsm_update.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_tick_new.restype = C.c_void_p
        L.ref_tick_new.argtypes = [C.c_char_p]
        L.ref_tick_free.argtypes = [C.c_void_p]
        L.ref_tick_set_gait.argtypes = [C.c_void_p, C.c_int]
        L.ref_bezier_foot_curve.argtypes = [C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(a, n=None):
    a = np.ascontiguousarray(a, np.float64).reshape(-1)
    assert n is None or a.size == n, (a.size, n)
    return a


def horizon() -> int:
    return lib().ref_horizon()


def mpc_dt() -> float:
    return lib().ref_mpc_dt()


def yaml_path(name: str) -> str:
    return os.path.join(REF_ROOT, "config", name)


def load_params(yaml: str):
    """LeggedParam::load (LeggedState.cpp:20-209) on one of the reference's yaml files."""
    q = np.zeros(12); r = np.zeros(12); inertia = np.zeros(9); foot = np.zeros(12); gains = np.zeros(27)
    mass = C.c_double(0); speed = C.c_double(0)
    rc = lib().ref_load_params(yaml.encode(), _p(q), _p(r), C.byref(mass), _p(inertia), C.byref(speed), _p(foot), _p(gains))
    if rc != 0:
        raise RuntimeError(f"ref_load_params({yaml}) -> {rc}")
    return {"q": q, "r": r, "mass": mass.value, "inertia": inertia, "gait_counter_speed": speed.value,
            "default_foot_pos_rel": foot, "kp_foot": gains[:12], "kd_foot": gains[12:24], "km_foot": gains[24:]}


def qp_build(q, r, mass, inertia, x0, rot, foot_abs, cmd, plan_contacts, gait, phase, gait_speed, z_solution=None):
    """The (P, q, A, l, u) ConvexQPSolver hands to OSQP for one robot (ConvexQPSolver.cpp:315-318).
    Returns dict(Pdiag[24H], q[24H], A[32H,24H], l, u, v_d_world[3], grf[12])."""
    L = lib(); nz, nr = L.ref_nz(), L.ref_nrow()
    Pd = np.zeros(nz); qv = np.zeros(nz); A = np.zeros((nr, nz), order="F"); lo = np.zeros(nr); up = np.zeros(nr)
    vdw = np.zeros(3); grf = np.zeros(12)
    pc = np.ascontiguousarray(plan_contacts, np.uint8).reshape(4)
    z = None if z_solution is None else _f(z_solution, nz)
    rc = L.ref_qp_build(_p(_f(q, 12)), _p(_f(r, 12)), C.c_double(mass), _p(_f(inertia, 9)), _p(_f(x0, 12)), _p(_f(rot, 9)),
                        _p(_f(foot_abs, 12)), _p(_f(cmd, 9)), _p(pc), C.c_int(int(gait)), _p(_f(phase, 4)),
                        C.c_double(gait_speed), _p(z), _p(Pd), _p(qv), _p(A), _p(lo), _p(up), _p(vdw), _p(grf))
    if rc != 0:
        raise RuntimeError(f"ref_qp_build -> {rc}")
    return {"Pdiag": Pd, "q": qv, "A": A, "l": lo, "u": up, "v_d_world": vdw, "grf": grf}


def osqp_settings():
    o = np.zeros(4); lib().ref_osqp_settings(_p(o))
    return {"eps_abs": o[0], "eps_rel": o[1], "warm_start": bool(o[2]), "verbose": bool(o[3])}


class Fsm:
    """One LeggedContactFSM (LeggedContactFSM.cpp), reset to phase 0."""

    def __init__(self, leg, gait, gait_speed):
        self.h = lib().ref_fsm_new(leg, gait, gait_speed)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_fsm_free(self.h); self.h = None

    def reset(self):
        lib().ref_fsm_reset(self.h)

    def set_phase(self, phase):
        lib().ref_fsm_set_phase(self.h, float(phase))

    def predict(self, delta_t) -> int:
        return lib().ref_fsm_predict(self.h, float(delta_t))

    def update(self, dt, foot_cur=(0, 0, 0), foot_target=(0, 0, 0), force_flag=False):
        info = np.zeros(6); pos = np.zeros(3); vel = np.zeros(3)
        lib().ref_fsm_update(self.h, float(dt), _p(_f(foot_cur, 3)), _p(_f(foot_target, 3)), int(bool(force_flag)), _p(info), _p(pos), _p(vel))
        return {"phase": info[0], "state": int(info[1]), "pattern_index": int(info[2]), "start": info[3], "end": info[4],
                "percent": info[5], "pos_target": pos, "vel_target": vel}


class Tick:
    """ConvexMpc(state) + ConvexMpc::update (ConvexMpc.cpp:6-62) on a LeggedState whose params come from
    LeggedParam::load of `yaml`; OSQP's answer is supplied by `solver(Pdiag,q,A,l,u) -> z`."""

    def __init__(self, yaml, gait=None):
        self.h = lib().ref_tick_new(yaml.encode())
        if not self.h:
            raise RuntimeError(f"ref_tick_new({yaml}) failed")
        if gait is not None:
            lib().ref_tick_set_gait(self.h, int(gait))

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_tick_free(self.h); self.h = None

    def update(self, t, dt, x0, rot, foot_abs, foot_world, foot_target_world, joy, contact_flag=(0, 0, 0, 0),
               roll_pitch_d=(0.0, 0.0), movement_mode=1.0, estimation_inited=True, solver=None, want_qp=False):
        L = lib(); nz, nr = L.ref_nz(), L.ref_nrow()
        ost = np.zeros(18); oin = np.zeros(24); pc = np.zeros(4, np.uint8); ph = np.zeros(4); vdw = np.zeros(3)
        eul = np.zeros(3); lvr = np.zeros(2); inited = C.c_int(0)
        Pd = qv = A = lo = up = None
        if want_qp:
            Pd = np.zeros(nz); qv = np.zeros(nz); A = np.zeros((nr, nz), order="F"); lo = np.zeros(nr); up = np.zeros(nr)

        def _cb(n, m, P_, q_, A_, l_, u_, z_):
            try:
                asn = lambda p, k: np.ctypeslib.as_array(p, shape=(k,))  # noqa: E731
                z = solver(asn(P_, n).copy(), asn(q_, n).copy(), asn(A_, n * m).reshape(m, n, order="F").copy(),
                           asn(l_, m).copy(), asn(u_, m).copy())
                if z is None:
                    return 1
                asn(z_, n)[:] = z
                return 0
            except Exception as e:  # noqa: BLE001
                self.cb_error = e
                return 1

        self.cb_error = None
        cb = SOLVE_CB(_cb) if solver is not None else C.cast(None, SOLVE_CB)
        rc = L.ref_tick_update(C.c_void_p(self.h), C.c_double(t), C.c_double(dt), C.c_int(int(estimation_inited)),
                               C.c_double(movement_mode), _p(_f(joy, 4)), _p(_f(roll_pitch_d, 2)), _p(_f(x0, 12)), _p(_f(rot, 9)),
                               _p(_f(foot_abs, 12)), _p(_f(foot_world, 12)), _p(_f(foot_target_world, 12)), _p(_f(contact_flag, 4)),
                               cb, _p(ost), _p(oin), _p(pc), _p(ph), _p(vdw), _p(eul), _p(lvr), _p(Pd), _p(qv), _p(A), _p(lo), _p(up),
                               C.byref(inited))
        if self.cb_error is not None:
            raise self.cb_error
        out = {"rc": rc, "optimized_state": ost, "optimized_input": oin, "plan_contacts": pc, "gait_phase": ph,
               "v_d_world": vdw, "euler_d": eul, "lin_vel_d_rel": lvr, "mpc_solver_inited": bool(inited.value)}
        if want_qp:
            out.update({"Pdiag": Pd, "q": qv, "A": A, "l": lo, "u": up})
        return out


def skew(v):
    o = np.zeros(9); lib().ref_skew(_p(_f(v, 3)), _p(o)); return o.reshape(3, 3, order="F")


def a1_fk(q, rho_opt, rho_fix):
    o = np.zeros(3); lib().ref_a1_fk(_p(_f(q, 3)), _p(_f(rho_opt, 3)), _p(_f(rho_fix, 5)), _p(o)); return o


def a1_jac(q, rho_opt, rho_fix):
    o = np.zeros(9); lib().ref_a1_jac(_p(_f(q, 3)), _p(_f(rho_opt, 3)), _p(_f(rho_fix, 5)), _p(o)); return o.reshape(3, 3, order="F")


def a1_inv_kin(p, cur_q, rho_opt, rho_fix):
    o = np.zeros(3); lib().ref_a1_inv_kin(_p(_f(p, 3)), _p(_f(cur_q, 3)), _p(_f(rho_opt, 3)), _p(_f(rho_fix, 5)), _p(o)); return o


def bezier_foot_curve(t, start, final, terrain_pitch=0.0):
    o = np.zeros(6); lib().ref_bezier_foot_curve(float(t), _p(_f(start, 3)), _p(_f(final, 3)), float(terrain_pitch), _p(o)); return o


def condense_captured(Pdiag, q, A, l, u, H=None):
    """Eliminate x from a captured reference QP (z = [u_0;x_1;...;u_{H-1};x_H]) using ITS OWN equality
    rows (the first 12H): returns (Hm, g, C, lo, up, X0, S) with cost 1/2 U'Hm U + g'U, rows lo <= C U <= up,
    X = X0 + S U.  Pure linear algebra on the captured numbers; no model knowledge beyond the variable order."""
    nz = Pdiag.size
    H = nz // 24 if H is None else H
    ui = np.concatenate([np.arange(24 * k, 24 * k + 12) for k in range(H)])
    xi = np.concatenate([np.arange(24 * k + 12, 24 * k + 24) for k in range(H)])
    E = A[:12 * H]
    assert np.array_equal(l[:12 * H], u[:12 * H])
    Ex, Eu, b = E[:, xi], E[:, ui], u[:12 * H]
    X0 = np.linalg.solve(Ex, b)
    S = -np.linalg.solve(Ex, Eu)
    Q, R = Pdiag[xi], Pdiag[ui]
    Hm = S.T @ (Q[:, None] * S) + np.diag(R)
    g = S.T @ (Q * X0 + q[xi]) + q[ui]
    Ci = A[12 * H:]
    assert not Ci[:, xi].any()
    return Hm, g, Ci[:, ui], l[12 * H:], u[12 * H:], X0, S
