"""Loader for the in-tree CUDA library liblmpc.so.  There is no fallback: if the library is
missing or cannot be loaded the import of anything that computes fails loudly."""
import ctypes as C
import os

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# LMPC_LIB_PATH: another in-tree build of the same library (kernel A/B experiments, tools/gpu_r2_ab.sh)
LIB_PATH = os.environ.get("LMPC_LIB_PATH") or os.path.join(_HERE, "liblmpc.so")
_LIB = None

EXPORTS = (
    "lmpc_params_preset", "lmpc_create", "lmpc_solve", "lmpc_solve_async", "lmpc_wait",
    "lmpc_reset_warm_start", "lmpc_destroy", "lmpc_last_error_string", "lmpc_abi_version",
    "lmpc_launch_count", "lmpc_last_kernel_ms", "lmpc_last_stats", "lmpc_gait_predict_contact",
    "lmpc_gait_contact_table", "lmpc_microbench_fp64", "lmpc_set_warm_start", "lmpc_last_warm_hits",
    "lmpc_rollout", "lmpc_alloc_pinned", "lmpc_free_pinned", "lmpc_last_not_converged",
    "lmpc_create_sharded", "lmpc_solve_sharded", "lmpc_set_warm_start_sharded", "lmpc_sharded_num_devices",
    "lmpc_sharded_handle", "lmpc_sharded_last_not_converged", "lmpc_destroy_sharded", "lmpc_abi_sizeof",
    "lmpc_fsm_reset", "lmpc_fsm_update", "lmpc_leg_geometry_a1", "lmpc_leg_kinematics", "lmpc_tau_ctrl",
)


class LmpcLibraryMissing(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise LmpcLibraryMissing(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    L.lmpc_last_error_string.restype = C.c_char_p
    L.lmpc_abi_version.restype = C.c_int32
    L.lmpc_launch_count.restype = C.c_int64
    L.lmpc_launch_count.argtypes = [C.c_void_p]
    L.lmpc_last_kernel_ms.restype = C.c_float
    L.lmpc_last_kernel_ms.argtypes = [C.c_void_p]
    L.lmpc_params_preset.argtypes = [C.c_char_p, C.POINTER(abi.LmpcParams)]
    L.lmpc_create.argtypes = [C.POINTER(abi.LmpcParams), C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]
    L.lmpc_solve.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcBatchIn), C.POINTER(abi.LmpcBatchOut)]
    L.lmpc_solve_async.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcBatchIn), C.POINTER(abi.LmpcBatchOut), C.c_void_p]
    L.lmpc_wait.argtypes = [C.c_void_p]
    L.lmpc_solve_noalloc.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcBatchIn), C.POINTER(abi.LmpcBatchOut), C.c_void_p, C.c_int32]
    L.lmpc_reset_warm_start.argtypes = [C.c_void_p]
    L.lmpc_set_warm_start.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
    L.lmpc_last_warm_hits.argtypes = [C.c_void_p]
    L.lmpc_last_warm_hits.restype = C.c_int64
    L.lmpc_rollout.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcRolloutCfg), C.POINTER(abi.LmpcRolloutState),
                               C.POINTER(abi.LmpcRolloutResult)]
    L.lmpc_alloc_pinned.argtypes = [C.c_size_t]
    L.lmpc_alloc_pinned.restype = C.c_void_p
    L.lmpc_free_pinned.argtypes = [C.c_void_p]
    L.lmpc_free_pinned.restype = None
    L.lmpc_destroy.argtypes = [C.c_void_p]
    L.lmpc_destroy.restype = None
    L.lmpc_last_stats.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.lmpc_gait_predict_contact.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double]
    L.lmpc_gait_predict_contact.restype = C.c_int32
    L.lmpc_gait_contact_table.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.lmpc_microbench_fp64.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double)]
    L.lmpc_last_not_converged.argtypes = [C.c_void_p]
    L.lmpc_last_not_converged.restype = C.c_int64
    L.lmpc_create_sharded.argtypes = [C.POINTER(abi.LmpcParams), C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_void_p)]
    L.lmpc_solve_sharded.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcBatchIn), C.POINTER(abi.LmpcBatchOut)]
    L.lmpc_set_warm_start_sharded.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
    L.lmpc_sharded_num_devices.argtypes = [C.c_void_p]
    L.lmpc_sharded_handle.argtypes = [C.c_void_p, C.c_int32]
    L.lmpc_sharded_handle.restype = C.c_void_p
    L.lmpc_sharded_last_not_converged.argtypes = [C.c_void_p]
    L.lmpc_sharded_last_not_converged.restype = C.c_int64
    L.lmpc_destroy_sharded.argtypes = [C.c_void_p]
    L.lmpc_destroy_sharded.restype = None
    L.lmpc_fsm_reset.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcFsmIo)]
    L.lmpc_fsm_update.argtypes = [C.c_void_p, C.c_int32, C.c_double, C.POINTER(abi.LmpcFsmIo)]
    L.lmpc_leg_geometry_a1.argtypes = [C.POINTER(abi.LmpcLegGeometry)]
    L.lmpc_leg_geometry_a1.restype = None
    L.lmpc_leg_kinematics.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcLegGeometry), C.POINTER(abi.LmpcLegIo)]
    L.lmpc_tau_ctrl.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.LmpcLegGeometry), C.POINTER(abi.LmpcLegIo)]
    L.lmpc_abi_sizeof.argtypes = [C.c_int32, C.c_int32]
    L.lmpc_abi_sizeof.restype = C.c_int64
    if L.lmpc_abi_version() != abi.ABI_VERSION:
        raise LmpcLibraryMissing(f"liblmpc.so ABI {L.lmpc_abi_version()} != python binding {abi.ABI_VERSION}")
    _LIB = L
    return L


def last_error() -> str:
    return lib().lmpc_last_error_string().decode()
