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
        st = self._L.lmpc_create(C.byref(self.params), self.max_batch, self.device, C.byref(self._h))
        if st != abi.OK:
            raise LmpcError(st, _lib.last_error())

    # -- reference-shaped entry point: host arrays in, first-step GRFs (+ diagnostics) out
    def compute_grfs(self, batch: dict, want_U: bool = True, strict: bool = False) -> dict:
        """batch: dict of numpy arrays in the lmpc_batch_in layout.  Returns dict with grf [B,12],
        status, iters (low 16 bits factorizations, high 16 polish rounds), active_set, U."""
        H = self.params.horizon
        sin, keep, nb = abi.host_batch_in(batch, H)
        sout, out = abi.host_batch_out(nb, H, want_U)
        st = self._L.lmpc_solve(self._h, nb, C.byref(sin), C.byref(sout))
        del keep
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED) or (strict and st != abi.OK):
            raise LmpcError(st, _lib.last_error())
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

    def wait(self) -> int:
        st = self._L.lmpc_wait(self._h)
        if st not in (abi.OK, abi.ERR_NOT_CONVERGED):
            raise LmpcError(st, _lib.last_error())
        return st

    def reset_warm_start(self):
        self._L.lmpc_reset_warm_start(self._h)

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
        if self._h:
            self._L.lmpc_destroy(self._h)
            self._h = C.c_void_p()

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
