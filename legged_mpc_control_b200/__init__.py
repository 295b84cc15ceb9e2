"""legged_mpc_control_b200 — B200-native batched convex-MPC engine (one hot path of
zha0ming1e/legged_mpc_control: SRBD build -> discretise -> condense -> QP -> 4x3 GRF).

The compute lives in liblmpc.so (hand-written CUDA for sm_100a behind the C ABI of
include/lmpc.h).  This package is the thin Python face used by the tests and bench; importing
`engine` without the built library raises — there is no CPU or PyTorch fallback.
"""
from . import abi, gait, presets, synth  # noqa: F401  (pure host-side helpers)

__all__ = ["abi", "gait", "presets", "synth", "ConvexMpcEngine"]


def __getattr__(name):
    if name in ("ConvexMpcEngine", "engine"):
        import importlib
        engine = importlib.import_module(__name__ + ".engine")
        return engine if name == "engine" else engine.ConvexMpcEngine
    raise AttributeError(name)
