"""CPU twin of lmpc_rollout (test infrastructure only — never imported by the package).

Closed loop per control step (period = params.dt), following the reference pieces the device
kernels cite: contact table = LeggedContactFSM::predict_contact_state(i*dt) (ConvexQPSolver.cpp:
329-346), Raibert footholds for swing legs (src/interfaces/BaseInterface.cpp:358-399), QP solve by
the oracle's exact active-set solver, plant x+ = A_d x + B_d u + c_d from the oracle's own step
matrices (ConvexQPSolver.cpp:198-228), phase advance (LeggedContactFSM.cpp:50).
"""
import numpy as np

from legged_mpc_control_b200 import gait as gaitmod
from legged_mpc_control_b200.synth import _rot_zyx

from . import binding as ob


def prepare(params, state, foot_rel, ground_z=0.0, dlim=(0.8, 0.8)):
    """Mutates state['foot_world'] for swing legs; returns the solver batch dict."""
    H, dt = params.horizon, params.dt
    x = state["x"]
    B = x.shape[0]
    cmd = state["cmd"]
    contact = gaitmod.contact_table(H, dt, state["gait"], state["phase"], state["gait_speed"])  # [B][H][4]
    if "fsm_state" in state:  # row 0 = plan_contacts = the state machine's own state (ConvexMpc.cpp:101-104)
        contact[:, 0, :] = state["fsm_state"]
    cy, sy = np.cos(x[:, 2]), np.sin(x[:, 2])
    vdx = cy * cmd[:, 3] - sy * cmd[:, 4]
    vdy = sy * cmd[:, 3] + cy * cmd[:, 4]
    kk = np.sqrt(np.abs(x[:, 5]) / 9.8)
    half = (1.0 / state["gait_speed"] / 2.0) / 2.0
    dx = np.clip(kk * (x[:, 9] - vdx) + half * vdx, -dlim[0], dlim[0])
    dy = np.clip(kk * (x[:, 10] - vdy) + half * vdy, -dlim[1], dlim[1])
    fr = np.asarray(foot_rel, np.float64).reshape(4, 3)
    fw = state["foot_world"].reshape(B, 4, 3)
    for leg in range(4):
        sw = contact[:, 0, leg] == 0
        tx = cy * fr[leg, 0] - sy * fr[leg, 1] + dx + x[:, 3]
        ty = sy * fr[leg, 0] + cy * fr[leg, 1] + dy + x[:, 4]
        fw[sw, leg, 0] = tx[sw]
        fw[sw, leg, 1] = ty[sw]
        fw[sw, leg, 2] = ground_z
    foot_abs = fw - x[:, None, 3:6]
    R = _rot_zyx(x[:, 0], x[:, 1], x[:, 2])
    rot = np.ascontiguousarray(np.transpose(R, (0, 2, 1))).reshape(B, 9)
    return {"x0": np.ascontiguousarray(x), "rot": rot, "foot_pos_abs": np.ascontiguousarray(foot_abs.reshape(B, 12)),
            "cmd": np.ascontiguousarray(cmd), "contact": np.ascontiguousarray(contact.reshape(B, 4 * H))}


def rollout(params, state, steps, foot_rel, ground_z=0.0, grf_override=None):
    """state: dict of numpy arrays (x, foot_world, phase, gait, gait_speed, cmd), mutated in place.
    grf_override [steps][B][12]: apply these forces instead of solving (plant-only check).
    Returns x_log, grf_log [steps][B][12]."""
    B = state["x"].shape[0]
    x_log = np.zeros((steps, B, 12))
    g_log = np.zeros((steps, B, 12))
    for t in range(steps):
        batch = prepare(params, state, foot_rel, ground_z)
        if grf_override is None:
            sol = ob.solve_batch(params, batch, eliminate_swing=True, want_U=False, want_kkt=False)
            assert sol["failures"] == 0
            grf = sol["grf"]
        else:
            grf = grf_override[t]
        for i in range(B):
            Ad, Bd, cd = ob.step_matrices(params, batch["rot"][i], batch["foot_pos_abs"][i], state["x"][i, 2])
            state["x"][i] = Ad @ state["x"][i] + Bd @ grf[i] + cd
        if "fsm_state" in state:  # LeggedContactFSM::update per robot-leg (no contact sensing in this plant)
            for i in range(B):
                for leg in range(4):
                    f = gaitmod.ContactFsm(int(state["gait"][i]), leg, float(state["gait_speed"][i]))
                    f.phase, f.index, f.state = float(state["phase"][i, leg]), int(state["fsm_index"][i, leg]), int(state["fsm_state"][i, leg])
                    f.start, f.end = float(state["fsm_start"][i, leg]), float(state["fsm_end"][i, leg])
                    f.update(params.dt, False)
                    state["phase"][i, leg], state["fsm_index"][i, leg], state["fsm_state"][i, leg] = f.phase, f.index, f.state
                    state["fsm_start"][i, leg], state["fsm_end"][i, leg] = f.start, f.end
        else:
            ph = state["phase"] + (state["gait_speed"] * params.dt)[:, None]
            while (ph > 1.0).any():
                ph[ph > 1.0] -= 1.0
            state["phase"][:] = ph
        x_log[t] = state["x"]
        g_log[t] = grf
    return x_log, g_log
