"""ctypes mirror of include/lmpc.h (struct layouts and constants only; no code path).

Kept byte-for-byte in step with the header; tests/test_host.py::test_ctypes_mirror_matches_library_layout checks sizeof/offsets against the
built library (lmpc_abi_version + a round trip through lmpc_params_preset).
"""
import ctypes as C

import numpy as np

ABI_VERSION = 2
MAX_HORIZON = 32
NUM_LEG = 4

OK, ERR_INVALID_ARG, ERR_CUDA, ERR_NO_DEVICE, ERR_NOT_CONVERGED, ERR_ALLOC = range(6)
PROB_OK, PROB_OK_IPM_ONLY, PROB_MAX_ITER, PROB_NON_FINITE, PROB_TOO_LARGE = range(5)
NOALLOC_MAX_BATCH = 4096  # include/lmpc.h LMPC_NOALLOC_MAX_BATCH
DISC_EULER, DISC_ZOH = 0, 1
COST_REF_HALF, COST_NS_TWO = 0, 1
PREC_F64, PREC_F32 = 0, 1
SOLVER_IPM, SOLVER_ACTIVE_SET = 0, 1
GAIT_TROT, GAIT_TROT_WITH_STAND, GAIT_CRAWL, GAIT_STAND, GAIT_BOUND, GAIT_PRONK = range(6)


def active_words(horizon: int) -> int:
    """LMPC_ACTIVE_WORDS(H)."""
    return (24 * horizon + 31) // 32


class LmpcParams(C.Structure):
    _fields_ = [
        ("horizon", C.c_int32),
        ("discretisation", C.c_int32),
        ("cost_scaling", C.c_int32),
        ("xref_index_offset", C.c_int32),
        ("precision", C.c_int32),
        ("solver", C.c_int32),
        ("max_iter", C.c_int32),
        ("reserved0", C.c_int32),
        ("dt", C.c_double),
        ("mu", C.c_double),
        ("fz_max", C.c_double),
        ("mass", C.c_double),
        ("gravity", C.c_double),
        ("active_tol", C.c_double),
        ("inertia", C.c_double * 9),
        ("q", C.c_double * 13),
        ("r", C.c_double * 12),
    ]

    def copy(self) -> "LmpcParams":
        out = LmpcParams()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(LmpcParams))
        return out


class LmpcBatchIn(C.Structure):
    _fields_ = [
        ("x0", C.c_void_p),
        ("rot", C.c_void_p),
        ("foot_pos_abs", C.c_void_p),
        ("cmd", C.c_void_p),
        ("contact", C.c_void_p),
        ("q", C.c_void_p),
        ("r", C.c_void_p),
        ("mu", C.c_void_p),
        ("fz_max", C.c_void_p),
        ("mass", C.c_void_p),
        ("inertia", C.c_void_p),
        ("on_device", C.c_int32),
        ("reserved0", C.c_int32),
    ]


class LmpcBatchOut(C.Structure):
    _fields_ = [
        ("grf", C.c_void_p),
        ("status", C.c_void_p),
        ("iters", C.c_void_p),
        ("active_set", C.c_void_p),
        ("U", C.c_void_p),
        ("on_device", C.c_int32),
        ("reserved0", C.c_int32),
    ]


class LmpcRolloutState(C.Structure):
    _fields_ = [
        ("x", C.c_void_p),
        ("foot_world", C.c_void_p),
        ("phase", C.c_void_p),
        ("gait", C.c_void_p),
        ("gait_speed", C.c_void_p),
        ("cmd", C.c_void_p),
        ("on_device", C.c_int32),
        ("reserved0", C.c_int32),
        ("fsm_state", C.c_void_p),
        ("fsm_index", C.c_void_p),
        ("fsm_start", C.c_void_p),
        ("fsm_end", C.c_void_p),
    ]


class LmpcFsmIo(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("gait_phase", "fsm_state", "fsm_index", "fsm_start", "fsm_end", "gait", "gait_speed",
                                          "foot_contact_flag", "plan_contacts", "contact_table")] + [("on_device", C.c_int32), ("reserved0", C.c_int32)]


class LmpcLegIo(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "joint_pos", "joint_vel", "rot", "root_pos", "root_lin_vel", "imu_ang_vel", "grf", "foot_pos_target_world", "foot_vel_target_world",
        "movement_mode", "foot_pos_rel", "j_foot", "foot_vel_rel", "foot_pos_abs", "foot_vel_abs", "foot_pos_world", "foot_vel_world",
        "joint_tau_tgt", "joint_ang_tgt", "prev_joint_ang_tgt", "joint_vel_tgt")] + [("on_device", C.c_int32), ("reserved0", C.c_int32)]


class LmpcLegGeometry(C.Structure):
    _fields_ = [("rho_fix", (C.c_double * 5) * 4), ("rho_opt", (C.c_double * 3) * 4)]


class LmpcRolloutCfg(C.Structure):
    _fields_ = [
        ("steps", C.c_int32),
        ("warm_start", C.c_int32),
        ("default_foot_pos_rel", C.c_double * 12),
        ("foot_delta_x_limit", C.c_double),
        ("foot_delta_y_limit", C.c_double),
        ("ground_z", C.c_double),
    ]


class LmpcRolloutResult(C.Structure):
    _fields_ = [
        ("grf", C.c_void_p),
        ("x_log", C.c_void_p),
        ("grf_log", C.c_void_p),
        ("solves", C.c_int64),
        ("factorizations", C.c_int64),
        ("polish_rounds", C.c_int64),
        ("not_converged", C.c_int64),
        ("warm_hits", C.c_int64),
        ("flops_executed", C.c_double),
        ("mean_n_eff", C.c_double),
        ("device_ms", C.c_float),
        ("reserved0", C.c_int32),
    ]


IN_FIELDS = (
    # name, per-problem elements (None -> 4*H), dtype, required
    ("x0", 12, np.float64, True),
    ("rot", 9, np.float64, True),
    ("foot_pos_abs", 12, np.float64, True),
    ("cmd", 9, np.float64, True),
    ("contact", None, np.uint8, True),
    ("q", 12, np.float64, False),
    ("r", 12, np.float64, False),
    ("mu", 1, np.float64, False),
    ("fz_max", 1, np.float64, False),
    ("mass", 1, np.float64, False),
    ("inertia", 9, np.float64, False),
)


def host_batch_in(batch: dict, horizon: int):
    """Build an LmpcBatchIn over numpy arrays (host pointers).  Returns (struct, keepalive)."""
    s = LmpcBatchIn()
    keep = []
    nb = None
    for name, per, dt, req in IN_FIELDS:
        a = batch.get(name)
        if a is None:
            if req:
                raise ValueError(f"batch field '{name}' is required")
            setattr(s, name, None)
            continue
        per = 4 * horizon if per is None else per
        a = np.ascontiguousarray(a, dtype=dt)
        if a.size % per:
            raise ValueError(f"batch field '{name}' has {a.size} elements, not a multiple of {per}")
        b = a.size // per
        if nb is None:
            nb = b
        elif b != nb:
            raise ValueError(f"batch field '{name}' holds {b} problems, expected {nb}")
        keep.append(a)
        setattr(s, name, a.ctypes.data)
    s.on_device = 0
    return s, keep, nb


def out_spec(nb: int, horizon: int, want_U: bool = True):
    """(name, shape, dtype) of every lmpc_batch_out array."""
    spec = [("grf", (nb, 12), np.float64), ("status", (nb,), np.int32), ("iters", (nb,), np.int32),
            ("active_set", (nb, active_words(horizon)), np.uint32)]
    if want_U:
        spec.append(("U", (nb, 12 * horizon), np.float64))
    return spec


def host_batch_out(nb: int, horizon: int, want_U: bool = True, arrays: dict = None):
    """Allocate numpy outputs (or use `arrays`, e.g. pinned views) and an LmpcBatchOut over them."""
    if arrays is None:
        out = {name: np.zeros(shape, dt) for name, shape, dt in out_spec(nb, horizon, want_U)}
        out["status"][:] = -1
    else:
        out = dict(arrays)
    s = LmpcBatchOut()
    s.grf = out["grf"].ctypes.data
    s.status = out["status"].ctypes.data
    s.iters = out["iters"].ctypes.data
    s.active_set = out["active_set"].ctypes.data
    s.U = out["U"].ctypes.data if want_U else None
    s.on_device = 0
    return s, out
