"""the batched contact state machine (LeggedContactFSM::update) and the GRF ->
joint command stage (A1 leg kinematics + BaseInterface::tau_ctrl_update).

CPU: the numpy restatement oracle/leg_ctrl_np.py against the reference's own A1Kinematics.cpp compiled in
oracle/_ref (fk, Jacobian, analytic inverse kinematics), wherever that library is present.
GPU: the device kernels (through the C ABI) against the numpy restatements / the host FSM mirror, which
tests/test_ref_pin.py pins to the compiled reference."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from legged_mpc_control_b200 import abi, gait as gaitmod, presets, synth  # noqa: E402
from oracle import leg_ctrl_np as lc  # noqa: E402
from oracle import ref_binding as rb  # noqa: E402

REF_GAITS = (abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND)


def _rand_joint(rng, n):
    q = np.empty((n, 12))
    for i in range(4):
        q[:, 3 * i] = rng.uniform(-0.6, 0.6, n)
        q[:, 3 * i + 1] = rng.uniform(0.2, 1.4, n)
        q[:, 3 * i + 2] = rng.uniform(-2.4, -0.9, n)
    return q


@pytest.mark.skipif(not rb.available(), reason="oracle/_ref not built (reference not mounted and no prebuilt library)")
def test_numpy_kinematics_match_compiled_reference():
    """fk / jac / inv_kin of oracle/leg_ctrl_np.py vs A1Kinematics.cpp (compiled, unmodified) for the four legs."""
    rng = np.random.default_rng(1)
    q = _rand_joint(rng, 300)
    for n in range(300):
        for i in range(4):
            qi = q[n, 3 * i:3 * i + 3]
            ro = rng.uniform(-0.01, 0.01, 3) if n % 3 == 0 else np.zeros(3)  # calibration offsets exercised too
            rf = lc.A1_RHO_FIX[i]
            assert np.abs(lc.fk(qi, ro, rf) - rb.a1_fk(qi, ro, rf)).max() <= 1e-15
            assert np.abs(lc.jac(qi, ro, rf) - rb.a1_jac(qi, ro, rf)).max() <= 5e-16
            # inverse kinematics of a reachable point, starting from a perturbed current angle
            p = rb.a1_fk(qi, np.zeros(3), rf)
            cur = qi + rng.uniform(-0.2, 0.2, 3)
            a, b = lc.inv_kin(p, cur, np.zeros(3), rf), rb.a1_inv_kin(p, cur, np.zeros(3), rf)
            assert np.abs(a - b).max() <= 1e-13, (n, i, a, b)
    # unreachable foot -> NaN in both
    far = np.array([0.18, 0.13, -0.9])
    assert np.isnan(lc.inv_kin(far, np.zeros(3), np.zeros(3), lc.A1_RHO_FIX[0])).any()
    assert np.isnan(rb.a1_inv_kin(far, np.zeros(3), np.zeros(3), lc.A1_RHO_FIX[0])).any()


def test_c_geometry_equals_reference_constants(lmpc_lib):
    import ctypes as C
    g = abi.LmpcLegGeometry()
    lmpc_lib.lmpc_leg_geometry_a1(C.byref(g))
    assert np.array_equal(np.array([[g.rho_fix[i][k] for k in range(5)] for i in range(4)]), lc.A1_RHO_FIX)
    assert not np.array([[g.rho_opt[i][k] for k in range(3)] for i in range(4)]).any()


def _leg_inputs(B, seed):
    rng = np.random.default_rng(seed)
    p = presets.preset("reference_a1")
    b = synth.make_batch(B, p.horizon, p.dt, seed=seed)
    q = _rand_joint(rng, B)
    a = {"joint_pos": q, "joint_vel": rng.normal(0, 2.0, (B, 12)), "rot": b["rot"], "root_pos": b["x0"][:, 3:6].copy(),
         "root_lin_vel": b["x0"][:, 9:12].copy(), "imu_ang_vel": rng.normal(0, 1.0, (B, 3)), "grf": rng.uniform(-60, 120, (B, 12)),
         "movement_mode": (rng.uniform(size=B) < 0.8).astype(np.float64), "joint_ang_tgt": rng.normal(0, 1.0, (B, 12))}
    # foot targets near the current feet (reachable), some far away (inv_kin -> NaN -> keep current angle)
    kin = lc.leg_kinematics(a)
    a["foot_pos_target_world"] = kin["foot_pos_world"] + rng.normal(0, 0.02, (B, 12))
    a["foot_pos_target_world"][::17, 2] -= 2.0
    a["foot_vel_target_world"] = rng.normal(0, 0.5, (B, 12))
    return a, kin


@pytest.mark.gpu
def test_leg_kinematics_and_tau_ctrl_match_numpy_restatement():
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    B = 2048
    a, kin_ref = _leg_inputs(B, 3)
    eng = ConvexMpcEngine(presets.preset("reference_a1"), B)
    kin = eng.leg_kinematics(a)
    for k, v in kin_ref.items():
        assert np.abs(kin[k] - v).max() <= 1e-12, k
    got, ref = eng.tau_ctrl(a), lc.tau_ctrl(a)
    assert np.abs(got["joint_tau_tgt"] - ref["joint_tau_tgt"]).max() <= 1e-11 * np.abs(ref["joint_tau_tgt"]).max()
    assert np.array_equal(got["prev_joint_ang_tgt"], ref["prev_joint_ang_tgt"])
    assert np.abs(got["joint_ang_tgt"] - ref["joint_ang_tgt"]).max() <= 1e-10
    # J^-1 amplifies rounding near the knee singularity: compare relative to the size of the answer
    assert (np.abs(got["joint_vel_tgt"] - ref["joint_vel_tgt"]) / np.maximum(1.0, np.abs(ref["joint_vel_tgt"]))).max() <= 1e-9
    far = np.zeros(B, bool); far[::17] = True; far &= a["movement_mode"] > 0
    assert far.sum() >= B // 30 and np.array_equal(got["joint_ang_tgt"][far, :3], a["joint_pos"][far, :3])  # NaN -> current angle (BaseInterface.cpp:472-478)
    assert not np.array_equal(got["joint_ang_tgt"][far, 3:], a["joint_pos"][far, 3:])
    stand = a["movement_mode"] == 0
    assert np.array_equal(got["joint_vel_tgt"][stand], a["joint_vel"][stand]) and np.array_equal(got["joint_ang_tgt"][stand], a["joint_pos"][stand])
    eng.close()


@pytest.mark.gpu
def test_device_fsm_equals_host_mirror():
    """lmpc_fsm_reset / lmpc_fsm_update: 4096 robots x 4 legs over 400 ticks of every reference gait with random
    early-contact flags — phase, state, pattern index, start/end times bit-identical to gait.ContactFsm (which
    tests/test_ref_pin.py pins to the compiled LeggedContactFSM.cpp), and the emitted contact table."""
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    p = presets.preset("reference_a1")
    B, T, dt = 512, 400, 0.01
    rng = np.random.default_rng(9)
    g = np.array([REF_GAITS[i % 4] for i in range(B)], np.int32)
    speed = rng.uniform(2.0, 4.5, B)
    fsm = {"gait_phase": np.zeros((B, 4)), "fsm_state": np.zeros((B, 4), np.int32), "fsm_index": np.zeros((B, 4), np.int32),
           "fsm_start": np.zeros((B, 4)), "fsm_end": np.zeros((B, 4)), "gait": g, "gait_speed": speed}
    eng = ConvexMpcEngine(p, B)
    out = eng.fsm_reset(fsm)
    mirror = [[gaitmod.ContactFsm(int(g[i]), leg, float(speed[i])) for leg in range(4)] for i in range(B)]
    assert np.array_equal(out["plan_contacts"], np.array([[m.state for m in r] for r in mirror], np.uint8))
    for t in range(T):
        flags = (rng.uniform(size=(B, 4)) < 0.3).astype(np.float64)
        want_table = t % 50 == 7
        out = eng.fsm_update(fsm, dt, flags, want_table=want_table)
        for i in range(B):
            for leg in range(4):
                mirror[i][leg].update(dt, bool(flags[i, leg]))
        for name, attr in (("gait_phase", "phase"), ("fsm_state", "state"), ("fsm_index", "index"), ("fsm_start", "start"), ("fsm_end", "end")):
            ref = np.array([[getattr(m, attr) for m in r] for r in mirror])
            assert np.array_equal(fsm[name], ref.astype(fsm[name].dtype)), (t, name)
        assert np.array_equal(out["plan_contacts"], fsm["fsm_state"].astype(np.uint8))
        if want_table:
            tab = gaitmod.contact_table(p.horizon, p.dt, g, fsm["gait_phase"], speed)
            tab[:, 0, :] = fsm["fsm_state"]
            assert np.array_equal(out["contact"], tab)
    eng.close()


@pytest.mark.gpu
def test_rollout_with_state_machine_matches_cpu_twin(oracle):
    """lmpc_rollout driven by the reference's contact STATE MACHINE (fsm_* arrays) instead of the stateless phase
    wrap: 12 robots x 40 steps over trot / trot-with-stand / crawl against the numpy twin + oracle QP."""
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    from oracle import rollout_np
    p = presets.preset("northstar_a1")
    B, steps = 12, 40
    st = synth.make_rollout_state(B, seed=8, gaits=(abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL))
    fr = st.pop("foot_rel")
    st["phase"][:] = 0.0
    st["fsm_state"] = np.zeros((B, 4), np.int32); st["fsm_index"] = np.zeros((B, 4), np.int32)
    st["fsm_start"] = np.zeros((B, 4)); st["fsm_end"] = np.zeros((B, 4))
    for i in range(B):
        for leg in range(4):
            m = gaitmod.ContactFsm(int(st["gait"][i]), leg, float(st["gait_speed"][i]))
            st["fsm_state"][i, leg], st["fsm_end"][i, leg] = m.state, m.end
    ref_state = {k: np.array(v, copy=True) for k, v in st.items()}
    x_ref, g_ref = rollout_np.rollout(p, ref_state, steps, fr)
    eng = ConvexMpcEngine(p, B)
    out = eng.rollout(st, steps, fr, warm_start=1, log=True)
    assert out["not_converged"] == 0
    assert np.abs(out["x_log"] - x_ref).max() <= 1e-7
    scale = np.maximum(1.0, np.abs(g_ref).max(axis=2, keepdims=True))
    assert (np.abs(out["grf_log"] - g_ref) / scale).max() <= 1e-6
    for k in ("phase", "fsm_state", "fsm_index", "fsm_start", "fsm_end"):
        assert np.array_equal(st[k], ref_state[k]), k
    assert (st["fsm_index"] > 0).any() and len(np.unique(st["fsm_state"])) == 2  # transitions happened
    eng.close()
