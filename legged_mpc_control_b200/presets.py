"""Named parameter presets (values are DATA taken from the reference's yaml/param loader).

Sources, relative to /root/reference/src/legged_ctrl/:
  A1 sim weights/mass/inertia   config/gazebo_a1_convex.yaml:40-72,133-140
  Go1 sim weights               config/gazebo_go1_convex.yaml:39-67 (mass/inertia fall back to
                                the loader defaults 13.0 / A1 inertia, src/LeggedState.cpp:146-160)
  mu = 0.3, dt, f_max = 180     src/mpc_ctrl/convex_mpc/ConvexQPSolver.cpp:25-26,171
  H = 30                        include/LeggedParams.h:13
The same tables exist in C (lmpc_params_preset in csrc/lmpc_host.cpp); tests compare the two.
"""
from . import abi

A1_Q = [60.0, 100.0, 0.0, 0.0, 0.0, 450.0, 0.15, 0.15, 100.0, 3.0, 3.0, 5.0, 0.0]
GO1_Q = [50.0, 100.0, 0.0, 0.0, 0.0, 3500.0, 0.01, 0.01, 10.0, 15.0, 15.0, 20.0, 0.0]
R_DEFAULT = [1e-4] * 12
A1_INERTIA = [0.0158533, 0.0, 0.0, 0.0, 0.0377999, 0.0, 0.0, 0.0, 0.0456542]  # column-major
A1_FOOTHOLDS = [[0.17, 0.17, -0.3], [0.17, -0.17, -0.3], [-0.17, 0.17, -0.3], [-0.17, -0.17, -0.3]]
GO1_FOOTHOLDS = [[0.17, 0.12, -0.3], [0.17, -0.17, -0.3], [-0.17, 0.17, -0.3], [-0.17, -0.12, -0.3]]
A1_GAIT_SPEED = 3.5   # config/gazebo_a1_convex.yaml:36
GO1_GAIT_SPEED = 4.0  # config/gazebo_go1_convex.yaml:35

_PRESETS = {
    # name: (H, dt, disc, scaling, xref_offset, q)
    "reference_a1": (30, 0.01, abi.DISC_EULER, abi.COST_REF_HALF, 0, A1_Q),
    "reference_go1": (30, 0.01, abi.DISC_EULER, abi.COST_REF_HALF, 0, GO1_Q),
    "northstar_a1": (10, 0.025, abi.DISC_ZOH, abi.COST_NS_TWO, 1, A1_Q),
    "northstar_go1_h20": (20, 0.025, abi.DISC_ZOH, abi.COST_NS_TWO, 1, GO1_Q),
}


def preset_names():
    return list(_PRESETS)


def preset(name: str) -> abi.LmpcParams:
    """Python twin of lmpc_params_preset()."""
    if name not in _PRESETS:
        raise KeyError(f"unknown preset '{name}' (have {preset_names()})")
    H, dt, disc, scal, off, q = _PRESETS[name]
    p = abi.LmpcParams()
    p.horizon = H
    p.discretisation = disc
    p.cost_scaling = scal
    p.xref_index_offset = off
    p.precision = abi.PREC_F64
    p.solver = abi.SOLVER_IPM
    p.max_iter = 0
    p.dt = dt
    p.mu = 0.3
    p.fz_max = 180.0
    p.mass = 13.0
    p.gravity = 9.8
    p.active_tol = 1e-6
    for i, v in enumerate(A1_INERTIA):
        p.inertia[i] = v
    for i, v in enumerate(q):
        p.q[i] = v
    for i, v in enumerate(R_DEFAULT):
        p.r[i] = v
    return p


def footholds(robot: str):
    return {"a1": A1_FOOTHOLDS, "go1": GO1_FOOTHOLDS}[robot]
