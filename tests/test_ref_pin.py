"""Pins the oracle to the REFERENCE'S OWN CODE.

oracle/_ref/libref_convex_mpc.so is the unmodified reference (ConvexQPSolver.cpp, ConvexMpc.cpp,
LeggedContactFSM.cpp, LeggedState.cpp, Utils.cpp, A1Kinematics.cpp) compiled from /root/reference against
the header shims of oracle/shim with an OsqpEigen stub that CAPTURES the (P, q, A, l, u) handed to OSQP
(oracle/Makefile target `ref`, oracle/ref_harness.cpp).  These tests check, on seeded random states:

  * every number of the captured QP == oracle/lmpc_oracle.c::orc_sparse_qp (bit for bit; <= 2 ulp on
    the four B_d entries whose 3x3-inverse evaluation order differs),
  * LeggedContactFSM::predict_contact_state == orc_gait_predict_contact == the product's gait tables,
  * LeggedParam::load of the reference's yaml files == the product's presets,
  * the exact minimiser of the CAPTURED QP (condensed with its own equality rows, solver-independent of
    the oracle's matrix build) == the oracle's answer, and compute_grfs' u_0 slice / NaN->0 rule,
  * the whole ConvexMpc::update tick (filters, FSM update, grf_update, output packing).

They run in the CPU suite wherever the prebuilt library or /root/reference is present."""
import numpy as np
import pytest

from legged_mpc_control_b200 import abi, gait as gaitmod, presets, synth
from oracle import ref_binding as rb

pytestmark = pytest.mark.skipif(not rb.available(), reason="oracle/_ref not built and /root/reference not mounted")

REF_GAITS = (abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND)
H = 30


def _have_yaml():
    import os
    return os.path.isdir(rb.REF_ROOT)


def _ulp_close(a, b, ulps):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    tol = ulps * np.spacing(np.maximum(np.abs(a), np.abs(b)))
    return np.all(np.abs(a - b) <= tol)


def _A_close(capA, A):
    """Equal everywhere except the angular rows of B_d (rows 12k+6..8, u columns: I_w^-1 [r]x dt,
    ConvexQPSolver.cpp:205,211), where the 3x3 inverse's evaluation order may differ: there within 4 eps of the
    block's largest entry (sums with cancellation, so not ulp-relative per entry)."""
    d = capA != A
    if not d.any():
        return True, 0.0
    rows, cols = np.nonzero(d)
    ok = np.all(rows < 12 * H) and np.all((rows % 12 >= 6) & (rows % 12 <= 8)) and np.all(cols % 24 < 12)
    blk = np.abs(A[6:9, :12]).max()
    err = np.abs(capA - A).max()
    return bool(ok and err <= 4 * np.finfo(np.float64).eps * blk), err


def _capture(p, b, i):
    return rb.qp_build(list(p.q)[:12], list(p.r), p.mass, list(p.inertia), b["x0"][i], b["rot"][i], b["foot_pos_abs"][i],
                       b["cmd"][i], b["contact"][i].reshape(H, 4)[0], b["gait"][i], b["phase"][i], b["gait_speed"][i])


def test_reference_constants():
    assert rb.horizon() == 30 and rb.mpc_dt() == 0.01  # LeggedParams.h:7,13; ConvexQPSolver.cpp:26
    s = rb.osqp_settings()  # ConvexQPSolver.cpp:182-185
    assert s == {"eps_abs": 1e-3, "eps_rel": 1e-4, "warm_start": True, "verbose": False}
    p = presets.preset("reference_a1")
    assert (p.horizon, p.dt, p.discretisation, p.xref_index_offset) == (30, 0.01, abi.DISC_EULER, 0)


@pytest.mark.skipif(not _have_yaml(), reason="reference yaml files not mounted")
@pytest.mark.parametrize("yaml,preset,robot", [("gazebo_a1_convex.yaml", "reference_a1", "a1"),
                                                 ("gazebo_go1_convex.yaml", "reference_go1", "go1")])
def test_yaml_loader_matches_presets(yaml, preset, robot):
    """LeggedParam::load (LeggedState.cpp:20-209) on the shipped yaml == lmpc_params_preset / presets.py."""
    P = rb.load_params(rb.yaml_path(yaml))
    p = presets.preset(preset)
    assert np.array_equal(P["q"], list(p.q)[:12]) and np.array_equal(P["r"], list(p.r))
    assert P["mass"] == p.mass and np.array_equal(P["inertia"], list(p.inertia))
    assert np.array_equal(P["default_foot_pos_rel"].reshape(4, 3), np.asarray(presets.footholds(robot)))
    assert P["gait_counter_speed"] == (presets.A1_GAIT_SPEED if robot == "a1" else presets.GO1_GAIT_SPEED)


@pytest.mark.parametrize("g", REF_GAITS)
def test_captured_qp_equals_oracle_sparse_qp(oracle, g):
    """256 random states per gait: the QP the reference code builds (ConvexQPSolver.cpp:16-196, 254-346) ==
    orc_sparse_qp.  P, q, l, u bit for bit; A bit for bit except B_d's angular rows (I_w^-1 [r]x), where the
    evaluation order of the 3x3 inverse may differ by an ulp."""
    p = presets.preset("reference_a1")
    n = 256
    b = synth.make_batch(n, H, p.dt, seed=100 + g, gaits=(g,))
    # random per-robot Go1-like weights on half of them exercise the cost path too
    worst = 0.0
    for i in range(n):
        cap = _capture(p, b, i)
        Pd, q, A, l, u = oracle.sparse_qp(p, b, i)
        A = np.asarray(A)
        assert np.array_equal(cap["Pdiag"], Pd), i
        assert np.array_equal(cap["q"], q), i
        assert np.array_equal(cap["l"], l) and np.array_equal(cap["u"], u), i
        assert np.array_equal(cap["A"] != 0, A != 0), i
        ok, err = _A_close(cap["A"], A)
        assert ok, (i, err)
        worst = max(worst, err)
        # contact rows of the bound block come from predict_contact_state through the reference's own FSMs
        assert np.array_equal(cap["u"][28 * H:], 180.0 * b["contact"][i].astype(np.float64))
    assert worst < 2e-16


def test_captured_qp_go1_weights_and_side_effect(oracle):
    """Go1 yaml weights; and ctrl.root_lin_vel_d_world = R v_d_rel is written (ConvexQPSolver.cpp:260)."""
    p = presets.preset("reference_go1")
    b = synth.make_batch(32, H, p.dt, seed=5, robot="go1", gaits=REF_GAITS)
    for i in range(32):
        cap = _capture(p, b, i)
        Pd, q, A, l, u = oracle.sparse_qp(p, b, i)
        assert np.array_equal(cap["Pdiag"], Pd) and np.array_equal(cap["q"], q)
        assert np.array_equal(cap["l"], l) and np.array_equal(cap["u"], u) and _A_close(cap["A"], np.asarray(A))[0]
        R = b["rot"][i].reshape(3, 3, order="F")
        assert np.allclose(cap["v_d_world"], R @ b["cmd"][i, 3:6], rtol=0, atol=1e-15)


def test_predict_contact_state_grid(oracle):
    """LeggedContactFSM::predict_contact_state (LeggedContactFSM.cpp:280-294) on a phase x delta_t grid,
    against the oracle's restatement, the numpy tables and the product's C table."""
    from legged_mpc_control_b200 import engine
    rng = np.random.default_rng(0)
    for g in REF_GAITS:
        for speed in (3.5, 4.0, 2.0):
            fs = [rb.Fsm(leg, g, speed) for leg in range(4)]
            phases = np.concatenate([np.linspace(0, 1, 41), [0.1, 0.25, 0.5, 0.6, 0.75], rng.uniform(0, 1, 60)])
            for ph in phases:
                for leg in range(4):
                    fs[leg].set_phase(ph)
                ref = np.array([[fs[leg].predict(i * 0.01) for leg in range(4)] for i in range(H)], np.uint8)
                orc = np.array([[oracle.gait_predict_contact(g, leg, ph, speed, i * 0.01) for leg in range(4)] for i in range(H)], np.uint8)
                tab = gaitmod.contact_table(H, 0.01, [g], np.full((1, 4), ph), [speed])[0]
                ctab = engine.gait_contact_table(H, 0.01, [g], np.full((1, 4), ph), [speed])[0]
                assert np.array_equal(ref, orc) and np.array_equal(ref, tab) and np.array_equal(ref, ctab), (g, speed, ph)


def _solve_captured(oracle, cap):
    Hm, g, Cm, lo, up, X0, S = rb.condense_captured(cap["Pdiag"], cap["q"], cap["A"], cap["l"], cap["u"])
    rows, rhs = [], []
    for k in range(Cm.shape[0]):
        if lo[k] > -1e29:
            rows.append(Cm[k]); rhs.append(lo[k])
        if up[k] < 1e29:
            rows.append(-Cm[k]); rhs.append(-up[k])
    rc, U, lam, it = oracle.gi_dense(Hm, g, np.array(rows), np.array(rhs))
    assert rc == 0
    z = np.zeros(24 * H)
    X = X0 + S @ U
    for k in range(H):
        z[24 * k:24 * k + 12] = U[12 * k:12 * k + 12]
        z[24 * k + 12:24 * k + 24] = X[12 * k:12 * k + 12]
    return U, z


def test_minimiser_of_captured_qp_equals_oracle(oracle):
    """Exact minimiser of the reference-built QP vs the oracle's own pipeline, and the reference's output
    slice: compute_grfs returns z[0:12] (ConvexQPSolver.cpp:320-322) and zeros on NaN (:323-326)."""
    p = presets.preset("reference_a1")
    b = synth.make_batch(12, H, p.dt, seed=21, gaits=REF_GAITS)
    for i in range(12):
        cap = _capture(p, b, i)
        assert np.array_equal(cap["grf"], np.zeros(12))  # no solution supplied -> NaN -> zeros
        U, z = _solve_captured(oracle, cap)
        r = oracle.solve_one(p, b, i, eliminate_swing=True)
        scale = max(1.0, np.abs(r["U"]).max())
        assert np.abs(U - r["U"]).max() <= 1e-9 * scale, (i, np.abs(U - r["U"]).max())
        # the captured QP's own constraint rows hold at z (equalities to 1e-9, inequalities feasible)
        Az = cap["A"] @ z
        assert np.all(Az >= cap["l"] - 1e-7) and np.all(Az <= cap["u"] + 1e-7)
        cap2 = rb.qp_build(list(p.q)[:12], list(p.r), p.mass, list(p.inertia), b["x0"][i], b["rot"][i], b["foot_pos_abs"][i],
                           b["cmd"][i], b["contact"][i].reshape(H, 4)[0], b["gait"][i], b["phase"][i], b["gait_speed"][i], z_solution=z)
        assert np.array_equal(cap2["grf"], z[:12])


def test_fsm_update_state_machine():
    """LeggedContactFSM::update (LeggedContactFSM.cpp:38-84) + common_enter (:214-229) + percent_in_state
    (:269-278) against the host mirror gait.ContactFsm: phase, stance/swing state, pattern index, state
    start/end times over 600 ticks per (gait, leg), with random early-contact flags."""
    speed, dt = 3.5, 0.01
    rng = np.random.default_rng(4)
    for g in REF_GAITS:
        for leg in range(4):
            f = rb.Fsm(leg, g, speed)
            m = gaitmod.ContactFsm(g, leg, speed)
            for k in range(600):
                flag = bool(rng.random() < 0.3)
                r = f.update(dt, force_flag=flag)
                m.update(dt, flag)
                assert r["phase"] == m.phase and r["state"] == m.state and r["pattern_index"] == m.index, (g, leg, k, r, vars(m))
                assert r["start"] == m.start and r["end"] == m.end
                pm = m.percent()
                assert r["percent"] == pm or (np.isnan(pm) and np.isnan(r["percent"]))
                for i in (0, 3, 11, 29):
                    assert f.predict(i * dt) == m.predict(i * dt)
    # early contact: trot leg 1 starts in swing [0, 0.5); at > 90 % of the swing a contact flag ends it at once
    f = rb.Fsm(1, abi.GAIT_TROT, speed)
    for k in range(12):
        r = f.update(dt, force_flag=False)
    assert r["state"] == 0 and r["percent"] < 0.9
    r = f.update(dt, force_flag=True)  # 13 * 0.035 = 0.455 -> 91 %
    assert r["percent"] == 0.0 and r["state"] == 1 and r["pattern_index"] == 1 and abs(r["start"] - 0.455) < 1e-12
    # trot, no early contact: state always equals the pattern entry and the phase stays in [0, 1 + speed*dt)
    m = gaitmod.ContactFsm(abi.GAIT_TROT, 0, speed)
    for k in range(1000):
        m.update(dt)
        assert m.state == m.table[m.index][0] and -1e-12 <= m.phase < 1.0 + speed * dt


@pytest.mark.skipif(not _have_yaml(), reason="reference yaml files not mounted")
def test_full_tick_convex_mpc_update(oracle):
    """ConvexMpc::update (ConvexMpc.cpp:24-62): joystick filter -> refs, foot_update -> plan_contacts, grf_update ->
    optimized_input[0:12] = u_0 of the QP it built; packing of optimized_state/input; estimation guard."""
    p = presets.preset("reference_a1")
    tick = rb.Tick(rb.yaml_path("gazebo_a1_convex.yaml"))
    b = synth.make_batch(1, H, p.dt, seed=33)
    foot_world = np.arange(12) * 0.01
    tgt = foot_world + 0.5
    r0 = tick.update(0.0, 0.01, b["x0"][0], b["rot"][0], b["foot_pos_abs"][0], foot_world, tgt, joy=(0.6, 0.1, 0.2, 0.27),
                     estimation_inited=False)
    assert r0["rc"] == 0 and not r0["mpc_solver_inited"] and not r0["optimized_input"].any()  # :27-31
    joy = (0.6, 0.1, 0.2, 0.27)
    yaw_d, fx, fy = 0.0, [], []
    for k in range(40):
        caps = {}

        def solver(Pd, q, A, l, u):
            caps.update(Pdiag=Pd, q=q, A=A, l=l, u=u)
            return _solve_captured(oracle, caps)[1]

        r = tick.update(0.01 * (k + 1), 0.01, b["x0"][0], b["rot"][0], b["foot_pos_abs"][0], foot_world, tgt, joy=joy, solver=solver)
        yaw_d += joy[2] * 0.01  # :38
        fx.append(joy[0]); fy.append(joy[1])
        # MovingWindowFilter(30) divides by the window size even while filling (MovingWindowFilter.hpp:36)
        assert abs(r["lin_vel_d_rel"][0] - sum(fx[-30:]) / 30.0) < 1e-12 and abs(r["lin_vel_d_rel"][1] - sum(fy[-30:]) / 30.0) < 1e-12
        assert abs(r["euler_d"][2] - yaw_d) < 1e-12 and r["mpc_solver_inited"]
        # the same tick through the oracle: contact row 0 = plan_contacts, rows i>0 predicted from the advanced phase
        ph = r["gait_phase"]
        contact = gaitmod.contact_table(H, 0.01, [abi.GAIT_TROT], ph[None], [3.5])[0]
        contact[0] = r["plan_contacts"]
        one = {"x0": b["x0"][:1], "rot": b["rot"][:1], "foot_pos_abs": b["foot_pos_abs"][:1],
               "cmd": np.array([[0.0, 0.0, joy[3], r["lin_vel_d_rel"][0], r["lin_vel_d_rel"][1], 0.0, 0.0, 0.0, joy[2]]]),
               "contact": contact.reshape(1, -1)}
        # root_lin_vel_d_rel[2] and root_ang_vel_d_rel[0..1] are never written by the tick: zero here
        orc = oracle.solve_one(p, one, 0, eliminate_swing=True)
        assert np.abs(r["optimized_input"][:12] - orc["grf"]).max() <= 1e-8 * max(1.0, np.abs(orc["grf"]).max()), k
        Pd, q, A, l, u = oracle.sparse_qp(p, one, 0)
        assert np.array_equal(caps["q"], q) and np.array_equal(caps["u"], u) and _A_close(caps["A"], np.asarray(A))[0]
        assert np.array_equal(r["optimized_state"][:3], [0.0, 0.0, joy[3]]) and abs(r["optimized_state"][5] - yaw_d) < 1e-12  # :49-50
