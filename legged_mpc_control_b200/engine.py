"""Python face of the C ABI (include/lmpc.h): ConvexMpcEngine.

Mirrors the reference's solver object for this path — ConvexQPSolver constructed once from the
weights (ConvexQPSolver.cpp:16-196), then per tick calc_mpc_reference + update_cons_matrix +
compute_grfs (ConvexMpc.cpp:70-72) — for a whole batch of robots at once.  All arithmetic runs in
liblmpc.so's CUDA kernels; this file only marshals pointers.
"""
import ctypes as C

import numpy as np

from . import _lib, abi


class LmpcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"lmpc status {code}: {msg}")
        self.code = code


class ConvexMpcEngine:
    """Batched replacement of legged::ConvexQPSolver (one engine per GPU)."""

    def __init__(self, params: abi.LmpcParams, max_batch: int, device: int = 0):
        self._L = _lib.lib()
        self.params = params.copy()
        self.max_batch = int(max_batch)
        self.device = int(device)
        self._h = C.c_void_p()
        self._pin_ptrs, self._pin_arrays, self._pin_key = [], None, None
        st = self._L.lmpc_create(C.byref(self.params), self.max_batch, self.device, C.byref(self._h))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())

    # -- reference-shaped entry point: host arrays in, first-step GRFs (+ diagnostics) out
    def _pinned_out(self, nb: int, want_U: bool) -> dict:
        """Engine-owned page-locked output arrays (lmpc_alloc_pinned), reused across calls."""
        key = (nb, bool(want_U))
        if self._pin_key != key:
            self._free_pinned()
            arrays = {}
            for name, shape, dt in abi.out_spec(nb, self.params.horizon, want_U):
                nbytes = int(np.prod(shape)) * np.dtype(dt).itemsize
                ptr = self._L.lmpc_alloc_pinned(nbytes)
                if not ptr:
                    raise LmpcError(abi.ERR_ALLOC, _lib.last_error())
                self._pin_ptrs.append(ptr)
                arrays[name] = np.frombuffer((C.c_char * nbytes).from_address(ptr), dtype=dt).reshape(shape)
            self._pin_arrays, self._pin_key = arrays, key
        return self._pin_arrays

    def _free_pinned(self):
        self._pin_arrays, self._pin_key = None, None
        for ptr in self._pin_ptrs:
            self._L.lmpc_free_pinned(ptr)
        self._pin_ptrs = []

    def compute_grfs(self, batch: dict, want_U: bool = True, strict: bool = False, copy: bool = True) -> dict:
        """batch: dict of numpy arrays in the lmpc_batch_in layout.  Returns dict with grf [B,12],
        status, iters (low 16 bits factorizations, high 16 polish rounds), active_set, U.
        copy=False returns views of the engine's page-locked output buffers (valid until the next
        call on this engine) — the zero-overhead form a control loop uses."""
        H = self.params.horizon
        sin, keep, nb = abi.host_batch_in(batch, H)
        sout, out = abi.host_batch_out(nb, H, want_U, arrays=self._pinned_out(nb, want_U))
        st = self._L.lmpc_solve(self._h, nb, C.byref(sin), C.byref(sout))
        del keep
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED) or (strict and st != abi.OK):
            raise LmpcError(st, _lib.last_error())
        if copy:
            out = {k: v.copy() for k, v in out.items()}
        out["rc"] = st
        out["factorizations"] = out["iters"] & 0xFFFF
        out["polish_rounds"] = out["iters"] >> 16
        return out

    # -- device-resident path (torch tensors or raw device pointers); used by bench `value`
    def solve_device(self, ptrs_in: dict, ptrs_out: dict, batch: int, stream: int = 0, wait: bool = True) -> int:
        """ptrs_*: dict name -> int device pointer (e.g. tensor.data_ptr()).  Required in: x0, rot,
        foot_pos_abs, cmd, contact; out: grf (+ optional status, iters, active_set, U)."""
        sin = abi.LmpcBatchIn()
        for name, *_ in abi.IN_FIELDS:
            setattr(sin, name, ptrs_in.get(name))
        sin.on_device = 1
        sout = abi.LmpcBatchOut()
        for name in ("grf", "status", "iters", "active_set", "U"):
            setattr(sout, name, ptrs_out.get(name))
        sout.on_device = 1
        st = self._L.lmpc_solve_async(self._h, int(batch), C.byref(sin), C.byref(sout), C.c_void_p(stream) if stream else None)
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())
        if wait:
            return self.wait()
        return st

    def solve_device_noalloc(self, ptrs_in: dict, ptrs_out: dict, batch: int, stream: int = 0, blocking: bool = True) -> int:
        """lmpc_solve_noalloc (include/lmpc.h): the one-warp solve on the library's static device
        scratch — no allocation on the call path.  Device pointers as in solve_device; batch <=
        abi.NOALLOC_MAX_BATCH; problems with n > 63 come back with status PROB_TOO_LARGE.  Returns
        OK or ERR_NOT_CONVERGED (blocking), raises on every other error."""
        sin = abi.LmpcBatchIn()
        for name, *_ in abi.IN_FIELDS:
            setattr(sin, name, ptrs_in.get(name))
        sin.on_device = 1
        sout = abi.LmpcBatchOut()
        for name in ("grf", "status", "iters", "active_set", "U"):
            setattr(sout, name, ptrs_out.get(name))
        sout.on_device = 1
        st = self._L.lmpc_solve_noalloc(self._h, int(batch), C.byref(sin), C.byref(sout),
                                        C.c_void_p(stream) if stream else None, 1 if blocking else 0)
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED):
            raise LmpcError(st, _lib.last_error())
        return st

    def wait(self) -> int:
        st = self._L.lmpc_wait(self._h)
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED):
            raise LmpcError(st, _lib.last_error())
        return st

    def set_warm_start(self, enable: bool, horizon_shift: int = 1):
        """OSQP warm-start analogue (ConvexQPSolver.cpp:185): reuse the previous solve's pyramid faces
        of the same batch index, shifted `horizon_shift` steps (lmpc_set_warm_start)."""
        st = self._L.lmpc_set_warm_start(self._h, 1 if enable else 0, int(horizon_shift))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())

    def reset_warm_start(self):
        self._L.lmpc_reset_warm_start(self._h)

    @property
    def last_warm_hits(self) -> int:
        return int(self._L.lmpc_last_warm_hits(self._h))

    # -- closed-loop batched rollout (BASELINE config 5), state resident on the device between steps
    def rollout(self, state: dict, steps: int, foot_rel, warm_start: int = 1, ground_z: float = 0.0,
                log: bool = False, on_device: bool = False) -> dict:
        """state: x [B,12], foot_world [B,12], phase [B,4], gait [B] int32, gait_speed [B], cmd [B,9]
        — numpy arrays (updated in place) or, with on_device=True, a dict of device pointers plus
        'batch'.  Returns the counters of lmpc_rollout_result (+ logs / last GRFs for host arrays)."""
        cfg = abi.LmpcRolloutCfg()
        cfg.steps, cfg.warm_start, cfg.ground_z = int(steps), int(warm_start), float(ground_z)
        fr = np.ascontiguousarray(foot_rel, np.float64).reshape(-1)
        for i in range(12):
            cfg.default_foot_pos_rel[i] = fr[i]
        st = abi.LmpcRolloutState()
        res = abi.LmpcRolloutResult()
        out = {}
        if on_device:
            B = int(state["batch"])
            for k in ("x", "foot_world", "phase", "gait", "gait_speed", "cmd", "fsm_state", "fsm_index", "fsm_start", "fsm_end"):
                if k in state:
                    setattr(st, k, state[k])
            st.on_device = 1
            res.grf, res.x_log, res.grf_log = state.get("grf"), state.get("x_log"), state.get("grf_log")
        else:
            spec = (("x", np.float64, 12), ("foot_world", np.float64, 12), ("phase", np.float64, 4), ("gait", np.int32, 1),
                    ("gait_speed", np.float64, 1), ("cmd", np.float64, 9))
            if "fsm_state" in state:  # the reference's contact state machine instead of the stateless phase wrap
                spec = spec + (("fsm_state", np.int32, 4), ("fsm_index", np.int32, 4), ("fsm_start", np.float64, 4), ("fsm_end", np.float64, 4))
            B = np.asarray(state["x"]).size // 12
            for k, dt, per in spec:
                a = state[k]
                if not (isinstance(a, np.ndarray) and a.dtype == dt and a.flags.c_contiguous and a.size == B * per):
                    raise ValueError(f"rollout state '{k}' must be a C-contiguous {np.dtype(dt).name} array of {B}x{per}")
                setattr(st, k, a.ctypes.data)
            st.on_device = 0
            out["grf"] = np.zeros((B, 12))
            res.grf = out["grf"].ctypes.data
            if log:
                out["x_log"] = np.zeros((steps, B, 12))
                out["grf_log"] = np.zeros((steps, B, 12))
                res.x_log, res.grf_log = out["x_log"].ctypes.data, out["grf_log"].ctypes.data
        rc = self._L.lmpc_rollout(self._h, B, C.byref(cfg), C.byref(st), C.byref(res))
        if rc not in (abi.OK, abi.ERR_NOT_CONVERGED):
            raise LmpcError(rc, _lib.last_error())
        out.update(rc=rc, solves=res.solves, factorizations=res.factorizations, polish_rounds=res.polish_rounds,
                   not_converged=res.not_converged, warm_hits=res.warm_hits, flops_executed=res.flops_executed,
                   mean_n_eff=res.mean_n_eff, device_ms=res.device_ms)
        return out

    # -- batched LeggedContactFSM (reset / update) on the device; numpy arrays updated in place
    def _fsm_io(self, fsm: dict, foot_contact_flag=None, want_table=False):
        B = np.asarray(fsm["gait"]).size
        io = abi.LmpcFsmIo()
        spec = (("gait_phase", np.float64, 4), ("fsm_state", np.int32, 4), ("fsm_index", np.int32, 4), ("fsm_start", np.float64, 4),
                ("fsm_end", np.float64, 4), ("gait", np.int32, 1), ("gait_speed", np.float64, 1))
        for k, dt, per in spec:
            a = fsm[k]
            if not (isinstance(a, np.ndarray) and a.dtype == dt and a.flags.c_contiguous and a.size == B * per):
                raise ValueError(f"fsm['{k}'] must be a C-contiguous {np.dtype(dt).name} array of {B}x{per}")
            setattr(io, k, a.ctypes.data)
        keep = []
        if foot_contact_flag is not None:
            f = np.ascontiguousarray(foot_contact_flag, np.float64).reshape(B, 4)
            keep.append(f)
            io.foot_contact_flag = f.ctypes.data
        out = {"plan_contacts": np.zeros((B, 4), np.uint8)}
        io.plan_contacts = out["plan_contacts"].ctypes.data
        if want_table:
            out["contact"] = np.zeros((B, self.params.horizon, 4), np.uint8)
            io.contact_table = out["contact"].ctypes.data
        return io, B, out, keep

    def fsm_reset(self, fsm: dict) -> dict:
        io, B, out, keep = self._fsm_io(fsm)
        st = self._L.lmpc_fsm_reset(self._h, B, C.byref(io))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())
        return out

    def fsm_update(self, fsm: dict, dt: float, foot_contact_flag=None, want_table: bool = False) -> dict:
        """One LeggedContactFSM::update(dt, ., ., flag) for every robot-leg; returns plan_contacts [B,4] (+ the
        solver's contact table [B,H,4] with want_table)."""
        io, B, out, keep = self._fsm_io(fsm, foot_contact_flag, want_table)
        st = self._L.lmpc_fsm_update(self._h, B, float(dt), C.byref(io))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())
        return out

    # -- batched leg kinematics and GRF -> joint command map
    @staticmethod
    def a1_geometry() -> abi.LmpcLegGeometry:
        g = abi.LmpcLegGeometry()
        _lib.lib().lmpc_leg_geometry_a1(C.byref(g))
        return g

    def _leg_call(self, fn, arrays: dict, outputs, geometry=None):
        B = np.asarray(arrays["joint_pos"]).size // 12
        io = abi.LmpcLegIo()
        keep = []
        per = {"joint_pos": 12, "joint_vel": 12, "rot": 9, "root_pos": 3, "root_lin_vel": 3, "imu_ang_vel": 3, "grf": 12,
               "foot_pos_target_world": 12, "foot_vel_target_world": 12, "movement_mode": 1}
        for k, n in per.items():
            if k in arrays and arrays[k] is not None:
                a = np.ascontiguousarray(arrays[k], np.float64).reshape(B, n)
                keep.append(a)
                setattr(io, k, a.ctypes.data)
        out = {}
        for k, n in outputs:
            out[k] = np.array(arrays[k], np.float64).reshape(B, n).copy() if k in arrays else np.zeros((B, n))
            setattr(io, k, out[k].ctypes.data)
        g = geometry if geometry is not None else self.a1_geometry()
        st = fn(self._h, B, C.byref(g), C.byref(io))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())
        return out

    def leg_kinematics(self, arrays: dict, geometry=None) -> dict:
        """BaseInterface.cpp:280-300 for B robots: fk, Jacobian blocks, foot position / velocity in the body,
        body-aligned ('abs') and world frames."""
        outs = (("foot_pos_rel", 12), ("j_foot", 36), ("foot_vel_rel", 12), ("foot_pos_abs", 12), ("foot_vel_abs", 12),
                ("foot_pos_world", 12), ("foot_vel_world", 12))
        return self._leg_call(self._L.lmpc_leg_kinematics, arrays, outs, geometry)

    def tau_ctrl(self, arrays: dict, geometry=None) -> dict:
        """BaseInterface::tau_ctrl_update (BaseInterface.cpp:451-500) for B robots: joint torques -J'R'f and the
        joint angle / velocity targets.  arrays['joint_ang_tgt'] (optional) is the previous target."""
        outs = (("joint_tau_tgt", 12), ("joint_ang_tgt", 12), ("prev_joint_ang_tgt", 12), ("joint_vel_tgt", 12))
        return self._leg_call(self._L.lmpc_tau_ctrl, arrays, outs, geometry)

    @property
    def launch_count(self) -> int:
        return int(self._L.lmpc_launch_count(self._h))

    @property
    def last_kernel_ms(self) -> float:
        return float(self._L.lmpc_last_kernel_ms(self._h))

    def last_stats(self) -> dict:
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self._L.lmpc_last_stats(self._h, C.byref(a), C.byref(b), C.byref(c))
        return {"mean_n_eff": a.value, "mean_factorizations": b.value, "flops_executed": c.value}

    def close(self):
        self._free_pinned()
        if self._h:
            self._L.lmpc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedConvexMpcEngine:
    """One process, several GPUs (lmpc_create_sharded): contiguous index ranges, one engine and one
    host thread per device inside liblmpc.so, no collective, host arrays in and out."""

    def __init__(self, params: abi.LmpcParams, max_batch: int, devices):
        self._L = _lib.lib()
        self.params = params.copy()
        self.devices = [int(d) for d in devices]
        self._s = C.c_void_p()
        self._pin_ptrs, self._pin_arrays, self._pin_key = [], None, None
        arr = (C.c_int32 * len(self.devices))(*self.devices)
        st = self._L.lmpc_create_sharded(C.byref(self.params), int(max_batch), arr, len(self.devices), C.byref(self._s))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())

    _pinned_out = ConvexMpcEngine._pinned_out
    _free_pinned = ConvexMpcEngine._free_pinned

    def compute_grfs(self, batch: dict, want_U: bool = False, copy: bool = True) -> dict:
        H = self.params.horizon
        sin, keep, nb = abi.host_batch_in(batch, H)
        sout, out = abi.host_batch_out(nb, H, want_U, arrays=self._pinned_out(nb, want_U))
        st = self._L.lmpc_solve_sharded(self._s, nb, C.byref(sin), C.byref(sout))
        del keep
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED):
            raise LmpcError(st, _lib.last_error())
        if copy:
            out = {k: v.copy() for k, v in out.items()}
        out["rc"] = st
        out["not_converged"] = int(self._L.lmpc_sharded_last_not_converged(self._s))
        return out

    def set_warm_start(self, enable: bool, horizon_shift: int = 1):
        st = self._L.lmpc_set_warm_start_sharded(self._s, 1 if enable else 0, int(horizon_shift))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())

    @property
    def launch_count(self) -> int:
        return sum(int(self._L.lmpc_launch_count(self._L.lmpc_sharded_handle(self._s, g))) for g in range(len(self.devices)))

    def close(self):
        self._free_pinned()
        if self._s:
            self._L.lmpc_destroy_sharded(self._s)
            self._s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def microbench_fp64(kind: int, device: int = 0, reps: int = 5) -> float:
    """TFLOP/s of the FP64 micro-kernel `kind` (include/lmpc_bench.h)."""
    out = C.c_double()
    rc = _lib.lib().lmpc_microbench_fp64(device, kind, reps, C.byref(out))
    if rc != 0:
        raise LmpcError(rc, "lmpc_microbench_fp64 failed (no usable CUDA device?)")
    return out.value


def gait_contact_table(horizon: int, dt: float, gait, phase, gait_speed) -> np.ndarray:
    """C twin of gait.contact_table (lmpc_gait_contact_table)."""
    gait = np.ascontiguousarray(gait, np.int32).reshape(-1)
    B = gait.shape[0]
    phase = np.ascontiguousarray(phase, np.float64).reshape(B, 4)
    gs = np.ascontiguousarray(gait_speed, np.float64).reshape(B)
    out = np.zeros((B, horizon, 4), np.uint8)
    st = _lib.lib().lmpc_gait_contact_table(B, horizon, float(dt), gait.ctypes.data, phase.ctypes.data, gs.ctypes.data, out.ctypes.data)
    if st != abi.OK:
        raise LmpcError(st, "lmpc_gait_contact_table: invalid argument")
    return out


def c_preset(name: str) -> abi.LmpcParams:
    p = abi.LmpcParams()
    st = _lib.lib().lmpc_params_preset(name.encode(), C.byref(p))
    if st != abi.OK:
        raise LmpcError(st, f"unknown preset {name}")
    return p
