"""ctypes binding of oracle/liblmpc_oracle.so (test infrastructure; see lmpc_oracle.h)."""
import ctypes as C
import os
import subprocess

import numpy as np

from legged_mpc_control_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_dp = C.POINTER(C.c_double)


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liblmpc_oracle.so")
    src = [os.path.join(_HERE, f) for f in ("lmpc_oracle.c", "lmpc_oracle.h")] + [
        os.path.join(_HERE, "..", "include", "lmpc.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liblmpc_oracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.orc_solve.restype = C.c_int
        _LIB.orc_solve_batch.restype = C.c_int
        _LIB.orc_gi_dense.restype = C.c_int
        _LIB.orc_gait_predict_contact.restype = C.c_int
        _LIB.orc_gait_predict_contact.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _one(batch, i, H):
    g = lambda k, n: np.ascontiguousarray(np.asarray(batch[k], np.float64).reshape(-1, n)[i])  # noqa: E731
    return (g("x0", 12), g("rot", 9), g("foot_pos_abs", 12), g("cmd", 9),
            np.ascontiguousarray(np.asarray(batch["contact"], np.uint8).reshape(-1, 4 * H)[i]))


def step_matrices(params, rot, foot, yaw_ref):
    Ad = np.zeros((12, 12), order="F"); Bd = np.zeros((12, 12), order="F"); cd = np.zeros(12)
    rot = np.ascontiguousarray(rot, np.float64); foot = np.ascontiguousarray(foot, np.float64)
    lib().orc_step_matrices(C.byref(params), _p(rot), _p(foot), C.c_double(yaw_ref), _p(Ad), _p(Bd), _p(cd))
    return Ad, Bd, cd


def reference_trajectory(params, x0, rot, cmd):
    H = params.horizon
    xref = np.zeros((H, 12))
    x0, rot, cmd = (np.ascontiguousarray(a, np.float64) for a in (x0, rot, cmd))
    lib().orc_reference_trajectory(C.byref(params), _p(x0), _p(rot), _p(cmd), _p(xref))
    return xref


def sparse_qp(params, batch, i=0):
    """(Pdiag, q, A, l, u) of the reference's sparse QP for problem i."""
    H = params.horizon
    x0, rot, foot, cmd, contact = _one(batch, i, H)
    nz, nrow = 24 * H, 32 * H
    Pd = np.zeros(nz); q = np.zeros(nz); A = np.zeros((nrow, nz), order="F"); l = np.zeros(nrow); u = np.zeros(nrow)
    lib().orc_sparse_qp(C.byref(params), _p(x0), _p(rot), _p(foot), _p(cmd), _p(contact), _p(Pd), _p(q), _p(A), _p(l), _p(u))
    return Pd, q, A, l, u


def condensed_qp(params, batch, i=0):
    """(Hm, g, e) of the condensed QP for problem i."""
    H = params.horizon
    x0, rot, foot, cmd, _ = _one(batch, i, H)
    n = 12 * H
    Hm = np.zeros((n, n), order="F"); g = np.zeros(n); e = np.zeros((H, 12))
    lib().orc_condensed_qp(C.byref(params), _p(x0), _p(rot), _p(foot), _p(cmd), _p(Hm), _p(g), _p(e))
    return Hm, g, e


def solve_one(params, batch, i=0, eliminate_swing=False):
    H = params.horizon
    x0, rot, foot, cmd, contact = _one(batch, i, H)
    U = np.zeros(12 * H); grf = np.zeros(12); act = np.zeros(abi.active_words(H), np.uint32)
    lam = np.zeros(24 * H); it = C.c_int(0); kkt = np.zeros(4)
    rc = lib().orc_solve(C.byref(params), _p(x0), _p(rot), _p(foot), _p(cmd), _p(contact),
                         C.c_int(1 if eliminate_swing else 0), _p(U), _p(grf), _p(act), _p(lam), C.byref(it), _p(kkt))
    return {"rc": rc, "U": U, "grf": grf, "active_set": act, "lambda": lam, "iters": it.value, "kkt": kkt}


def solve_batch(params, batch, eliminate_swing=True, nthreads=None, want_U=True, want_kkt=True):
    H = params.horizon
    sin, keep, nb = abi.host_batch_in(batch, H)
    sout, out = abi.host_batch_out(nb, H, want_U)
    kkt = np.zeros((nb, 4)) if want_kkt else None
    if nthreads is None:
        nthreads = os.cpu_count() or 1
    fails = lib().orc_solve_batch(C.byref(params), C.c_int(nb), C.byref(sin), C.byref(sout),
                                  _p(kkt) if want_kkt else None, C.c_int(1 if eliminate_swing else 0), C.c_int(nthreads))
    out["failures"] = fails
    if want_kkt:
        out["kkt"] = kkt
    del keep
    return out


def gi_dense(G, g, Cm, b):
    """min 1/2 x'Gx + g'x s.t. Cm x >= b."""
    n = G.shape[0]; m = Cm.shape[0]
    Gf = np.array(G, np.float64, order="F"); gf = np.ascontiguousarray(g, np.float64)
    Cf = np.array(Cm, np.float64, order="F").reshape(m, n, order="F"); bf = np.ascontiguousarray(b, np.float64)
    x = np.zeros(n); lam = np.zeros(max(m, 1)); it = C.c_int(0)
    rc = lib().orc_gi_dense(C.c_int(n), C.c_int(m), _p(Gf), _p(gf), _p(Cf), _p(bf), _p(x), _p(lam), C.byref(it))
    return rc, x, lam[:m], it.value


def gait_predict_contact(gait, leg, phase, speed, delta_t):
    return lib().orc_gait_predict_contact(gait, leg, float(phase), float(speed), float(delta_t))
