"""CPU tests of the oracle (test infrastructure) — the reference holds no golden vectors for this
path (parity unpinned), so the oracle is pinned by (1) an independent numpy
restatement of the reference's matrix construction, (2) solver-independent KKT certificates on
both the condensed QP and the reference's own sparse QP, (3) analytic known-answer tests
(SURVEY.md §8c K1-K10), and (4) committed golden fixtures."""
import os

import numpy as np
import pytest
from scipy.linalg import expm

from legged_mpc_control_b200 import abi, gait, presets, synth
from oracle import oracle_np as on

PRESETS = ("northstar_a1", "reference_a1", "northstar_go1_h20")


def _one(oracle, b, i, H):
    return oracle._one(b, i, H)


@pytest.mark.parametrize("name", PRESETS)
def test_matrices_match_numpy_twin(oracle, name):
    """orc_step_matrices / orc_sparse_qp / orc_condensed_qp against the independent numpy twin
    (ConvexQPSolver.cpp:198-228, 254-346)."""
    p = presets.preset(name)
    H = p.horizon
    b = synth.make_batch(3, H, p.dt, seed=7, robot="go1" if "go1" in name else "a1",
                         gaits=(abi.GAIT_TROT, abi.GAIT_CRAWL))
    for i in range(3):
        x0, rot, foot, cmd, contact = _one(oracle, b, i, H)
        Ad, Bd, cd = oracle.step_matrices(p, rot, foot, 0.3)
        pp = on._params(p)
        Adn, Bdn, cdn = on.discretise(pp, *on.continuous(pp, rot, foot, 0.3))
        assert np.abs(Ad - Adn).max() < 1e-14 and np.abs(Bd - Bdn).max() < 1e-12 and np.abs(cd - cdn).max() < 1e-14
        Hm, g, e = oracle.condensed_qp(p, b, i)
        Hn, gn, en, _ = on.condensed_qp(p, x0, rot, foot, cmd)
        assert np.abs(Hm - Hn).max() <= 1e-13 * np.abs(Hn).max()
        assert np.abs(g - gn).max() <= 1e-12 * max(1.0, np.abs(gn).max())
        assert np.abs(e - en).max() < 1e-12
        got = oracle.sparse_qp(p, b, i)
        ref = on.sparse_qp(p, x0, rot, foot, cmd, contact)
        for a, r in zip(got, ref):
            assert np.abs(a - r).max() < 1e-13


def test_reference_sparse_structure(oracle):
    """Sizes, row order and constants of the reference's QP (ConvexQPSolver.cpp:17-23, 131-177)."""
    p = presets.preset("reference_a1")
    H = p.horizon
    b = synth.make_batch(1, H, p.dt, seed=1)
    Pd, q, A, l, u = oracle.sparse_qp(p, b, 0)
    assert A.shape == (32 * H, 24 * H) == (960, 720)
    assert np.allclose(Pd[:12], 1e-4) and np.allclose(Pd[12:24], list(p.q)[:12])
    # gravity on row 12i+11 of the dynamics rows: 9.8 * dt  (:175)
    assert np.allclose(l[12 + 11:12 * H:12], 9.8 * 0.01) and np.array_equal(l[:12 * H], u[:12 * H])
    # friction row 0 of step 2 leg 1: fx + mu fz in [0, inf)
    r = 12 * H + 16 * 2 + 4 * 1
    assert A[r, 24 * 2 + 3] == 1 and A[r, 24 * 2 + 5] == 0.3 and l[r] == 0 and u[r] >= 1e29
    assert A[r + 1, 24 * 2 + 5] == -0.3 and l[r + 1] <= -1e29 and u[r + 1] == 0
    # bound rows: 180 * contact  (:336,341)
    c = b["contact"].reshape(H, 4)
    assert np.array_equal(u[12 * H + 16 * H:], 180.0 * c.reshape(-1))
    # x_ref index quirk: the gradient of x_1 uses the CURRENT yaw/pos (i = 0)  (:267-269)
    x0 = b["x0"][0]
    assert np.isclose(q[12 + 2], -p.q[2] * x0[2]) and np.isclose(q[12 + 5], -p.q[5] * b["cmd"][0, 2])


def test_zoh_closed_form_and_nilpotency(oracle):
    """KAT K8: the ZOH matrices equal expm([[Ac,Bc,c],[0,0,0]] dt) and the closed form
    Bd = (dt I + dt^2/2 Ac) Bc.  KAT K9: Ac^2 = 0 so exp(Ac dt) = I + Ac dt exactly."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(4, p.horizon, p.dt, seed=3)
    pp = on._params(p)
    for i in range(4):
        x0, rot, foot, cmd, _ = _one(oracle, b, i, p.horizon)
        Ac, Bc, c = on.continuous(pp, rot, foot, x0[2])
        assert np.abs(Ac @ Ac).max() == 0.0
        Ad, Bd, cd = oracle.step_matrices(p, rot, foot, x0[2])
        M = np.zeros((25, 25)); M[:12, :12] = Ac; M[:12, 12:24] = Bc; M[:12, 24] = c
        E = expm(M * p.dt)
        assert np.abs(Ad - E[:12, :12]).max() < 1e-14
        assert np.abs(Bd - E[:12, 12:24]).max() < 1e-12
        assert np.abs(Bd - (p.dt * np.eye(12) + 0.5 * p.dt ** 2 * Ac) @ Bc).max() < 1e-12
        assert np.abs(cd - E[:12, 24]).max() < 1e-14


@pytest.mark.parametrize("name", PRESETS)
def test_kkt_certificates(oracle, name):
    """The oracle's U is the exact minimiser: NNLS finds multipliers >= 0 on the primal-active
    rows with stationarity <= 1e-9 (a certificate that does not depend on the oracle's solver)."""
    p = presets.preset(name)
    H = p.horizon
    b = synth.make_batch(6, H, p.dt, seed=11, robot="go1" if "go1" in name else "a1",
                         gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_STAND))
    for i in range(6):
        x0, rot, foot, cmd, contact = _one(oracle, b, i, H)
        r = oracle.solve_one(p, b, i, eliminate_swing=False)
        assert r["rc"] == 0
        assert r["kkt"][0] <= 1e-9 and r["kkt"][1] <= 1e-10 and r["kkt"][2] >= 0 and r["kkt"][3] <= 1e-9
        Hn, gn, _, _ = on.condensed_qp(p, x0, rot, foot, cmd)
        Cm, bb = on.condensed_rows(p, contact)
        cert = on.kkt_certificate(Hn, gn, Cm, bb, r["U"])
        assert cert["stationarity"] <= 1e-9 and cert["primal_violation"] <= 1e-9


def test_sparse_equals_condensed(oracle):
    """the reference's sparse 24H-variable QP and the condensed 12H-variable QP
    have the same minimiser.  z* rolled out from the condensed U* satisfies the KKT conditions of
    the sparse QP (free multipliers on the dynamics rows)."""
    for name, nprob in (("northstar_a1", 3), ("reference_a1", 1)):
        p = presets.preset(name)
        H = p.horizon
        b = synth.make_batch(nprob, H, p.dt, seed=5)
        for i in range(nprob):
            x0, rot, foot, cmd, contact = _one(oracle, b, i, H)
            U = oracle.solve_one(p, b, i, eliminate_swing=False)["U"]
            P, q, A, l, u = on.sparse_qp(p, x0, rot, foot, cmd, contact)
            z = on.rollout_z(p, x0, rot, foot, cmd, U)
            cert = on.sparse_kkt_certificate(P, q, A, l, u, z)
            assert cert["stationarity"] <= 1e-9 and cert["primal_violation"] <= 1e-9


def test_full_rows_equal_swing_eliminated(oracle):
    """Keeping the reference's full 24H one-sided row set (swing legs pinned by their [0,0] bound)
    or eliminating swing-leg variables first gives the same U and the same primal-active set."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(64, p.horizon, p.dt, seed=21, gaits=(abi.GAIT_TROT, abi.GAIT_CRAWL, abi.GAIT_PRONK))
    full = oracle.solve_batch(p, b, eliminate_swing=False)
    red = oracle.solve_batch(p, b, eliminate_swing=True)
    assert full["failures"] == 0 and red["failures"] == 0
    assert np.abs(full["U"] - red["U"]).max() < 1e-7
    assert np.array_equal(full["active_set"], red["active_set"])


def _solve(oracle, p, b):
    out = oracle.solve_batch(p, b, eliminate_swing=False)
    assert out["failures"] == 0
    return out


def test_kat_standing(oracle):
    """K1: all-stance symmetric stand, x0 = x_ref: fx = fy = 0, four equal fz close to m g / 4
    (fixture idea: the reference's test_wbc.cpp:126-129)."""
    for name in ("northstar_a1", "reference_a1"):
        p = presets.preset(name)
        out = _solve(oracle, p, synth.standing_problem(p.horizon))
        f = out["grf"][0].reshape(4, 3)
        assert np.abs(f[:, :2]).max() < 1e-8 and np.ptp(f[:, 2]) < 1e-8
        assert abs(f[0, 2] - p.mass * p.gravity / 4) < 1.5


def test_kat_swing_zero_and_all_swing(oracle):
    """K2: swing legs get exact zeros.  K3: all legs swinging -> U = 0."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(8, p.horizon, p.dt, seed=2)
    b["contact"][3, :] = 0
    out = _solve(oracle, p, b)
    U = out["U"].reshape(8, p.horizon, 4, 3)
    c = b["contact"].reshape(8, p.horizon, 4).astype(bool)
    assert np.abs(U[~c]).max() < 1e-9
    assert np.abs(U[3]).max() < 1e-9


def test_kat_mirror_and_yaw_symmetry(oracle):
    """K4: left/right mirror symmetry.  K5: rotating state, feet and world quantities about z by
    psi rotates the GRFs (checked on a problem whose friction rows are inactive, since the
    pyramid itself is not rotation invariant)."""
    p = presets.preset("northstar_a1")
    H = p.horizon
    b = synth.standing_problem(H)
    b["x0"][0, 5] = 0.26            # a little low: wants more fz, no lateral force
    b["x0"][0, 0] = 0.05            # roll
    out = _solve(oracle, p, b)
    m = {k: np.array(v, copy=True) for k, v in b.items()}
    m["x0"][0, 0] *= -1             # mirror roll
    outm = _solve(oracle, p, m)
    f, fm = out["grf"][0].reshape(4, 3), outm["grf"][0].reshape(4, 3)
    # mirror about the xz-plane swaps FL<->FR, RL<->RR and flips fy
    perm = [1, 0, 3, 2]
    assert np.abs(fm[perm] * [1, -1, 1] - f).max() < 1e-7
    psi = 0.9
    cz, sz = np.cos(psi), np.sin(psi)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    r = {k: np.array(v, copy=True) for k, v in b.items()}
    r["x0"][0, 2] += psi
    r["rot"] = np.ascontiguousarray(Rz.T).reshape(1, 9)  # column-major of Rz
    r["foot_pos_abs"] = (b["foot_pos_abs"].reshape(4, 3) @ Rz.T).reshape(1, 12)
    # roll about the body x axis: rebuild R = Rz * Rx(roll) for both problems
    roll = b["x0"][0, 0]
    Rx = np.array([[1, 0, 0], [0, np.cos(roll), -np.sin(roll)], [0, np.sin(roll), np.cos(roll)]])
    b2 = {k: np.array(v, copy=True) for k, v in b.items()}
    b2["rot"] = np.ascontiguousarray(Rx.T).reshape(1, 9)
    b2["foot_pos_abs"] = (np.asarray(presets.A1_FOOTHOLDS) @ Rx.T).reshape(1, 12)
    r["rot"] = np.ascontiguousarray((Rz @ Rx).T).reshape(1, 9)
    r["foot_pos_abs"] = (np.asarray(presets.A1_FOOTHOLDS) @ (Rz @ Rx).T).reshape(1, 12)
    o2, orr = _solve(oracle, p, b2), _solve(oracle, p, r)
    assert np.abs(orr["grf"][0].reshape(4, 3) - o2["grf"][0].reshape(4, 3) @ Rz.T).max() < 1e-6


def test_kat_friction_and_cap_active(oracle):
    """K6: a large lateral velocity error saturates the friction rows (|fx| = mu fz).
    K7: a large height deficit drives fz into its cap (180 N)."""
    p = presets.preset("northstar_a1")
    H = p.horizon
    b = synth.standing_problem(H)
    b["x0"][0, 9] = 2.0  # moving at 2 m/s in x, commanded 0 -> brake as hard as friction allows
    out = _solve(oracle, p, b)
    f = out["grf"][0].reshape(4, 3)
    assert np.allclose(f[:, 0], -p.mu * f[:, 2], atol=1e-7) and (f[:, 2] > 1).all()
    bits = np.unpackbits(out["active_set"][0].view(np.uint8), bitorder="little")
    assert bits[0] == 1 and bits[1] == 0  # row 0 (0 <= fx + mu fz) active on leg 0, step 0
    low = synth.standing_problem(H)
    low["x0"][0, 5] = 0.20; low["cmd"][0, 2] = 0.40; low["x0"][0, 11] = -1.0  # low and falling
    out = _solve(oracle, p, low)
    f = out["grf"][0].reshape(4, 3)
    assert np.allclose(f[:, 2], 180.0, atol=1e-7)
    bits = np.unpackbits(out["active_set"][0].view(np.uint8), bitorder="little")
    assert bits[20 * H:20 * H + 4].all()


def test_kat_scaling_invariance(oracle):
    """K10: REF_HALF and NS_TWO cost scalings give the same U (multipliers scale by 2)."""
    p = presets.preset("northstar_a1")
    p2 = p.copy(); p2.cost_scaling = abi.COST_REF_HALF
    b = synth.make_batch(16, p.horizon, p.dt, seed=8)
    a, c = _solve(oracle, p, b), _solve(oracle, p2, b)
    assert np.abs(a["U"] - c["U"]).max() < 1e-7 and np.array_equal(a["active_set"], c["active_set"])


def test_gi_dense_random_qp(oracle):
    """The generic Goldfarb-Idnani core on random strictly convex QPs, certified by NNLS."""
    rng = np.random.default_rng(0)
    for trial in range(20):
        n, m = int(rng.integers(2, 12)), int(rng.integers(1, 20))
        A = rng.standard_normal((n + 2, n)); G = A.T @ A + 0.1 * np.eye(n)
        g = rng.standard_normal(n); Cm = rng.standard_normal((m, n)); x_feas = rng.standard_normal(n)
        bb = Cm @ x_feas - rng.random(m)  # feasible by construction
        rc, x, lam, it = oracle.gi_dense(G, g, Cm, bb)
        assert rc == 0
        cert = on.kkt_certificate(G, g, Cm, bb, x, act_tol=1e-8)
        assert cert["stationarity"] < 1e-8 and cert["primal_violation"] < 1e-9
        assert (lam >= 0).all()
    # infeasible problem is reported
    rc, *_ = oracle.gi_dense(np.eye(2), np.zeros(2), np.array([[1.0, 0], [-1.0, 0]]), np.array([1.0, 1.0]))
    assert rc == 2


def test_golden_fixtures_pin_the_oracle(oracle):
    """The committed fixtures (tests/golden/make_golden.py) are reproduced bit-for-bit in the
    active set and to 1e-9 in U by a fresh oracle build, and certified by the numpy twin."""
    gdir = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gdir) if f.endswith(".npz") and not f.startswith("golden_adapter"))  # (the adapter-tick fixture has its own test)
    assert len(files) >= 4 and "golden_refbuild_a1_h30.npz" in files
    for f in files:
        g = np.load(os.path.join(gdir, f))
        p = presets.preset(str(g["preset"]))
        batch = {k: g[k] for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
        out = oracle.solve_batch(p, batch, eliminate_swing=True)
        assert out["failures"] == 0
        if "source" in g.files:  # tests/golden/make_golden_ref.py: minimiser of the QP the REFERENCE'S code built
            assert str(g["source"]) == "oracle/_ref"
            assert np.abs(out["U"] - g["U"]).max() <= 1e-9 * max(1.0, np.abs(g["U"]).max())
            for i in range(len(g["x0"])):  # ... and the oracle's sparse QP reproduces the captured matrices' fingerprint
                Pd, q, A, l, u = oracle.sparse_qp(p, batch, i)
                fp = np.array([np.asarray(A).sum(), np.abs(np.asarray(A)).sum(), q.sum(), u[:12 * p.horizon].sum()])
                assert np.allclose(fp, g["qp_fingerprint"][i], rtol=1e-13, atol=0), (f, i)
        else:
            assert np.abs(out["U"] - g["U"]).max() < 1e-9
        assert np.array_equal(out["active_set"], g["active_set"])
        x0, rot, foot, cmd, contact = oracle._one(batch, 0, p.horizon)
        Hn, gn, _, _ = on.condensed_qp(p, x0, rot, foot, cmd)
        Cm, bb = on.condensed_rows(p, contact)
        cert = on.kkt_certificate(Hn, gn, Cm, bb, g["U"][0])
        assert cert["stationarity"] <= 1e-9 and cert["primal_violation"] <= 1e-9


def test_oracle_nan_rule(oracle):
    """Non-finite input -> zeros and a non-zero status (ConvexQPSolver.cpp:321-326)."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(2, p.horizon, p.dt, seed=1)
    b["x0"][0, 0] = np.nan
    out = oracle.solve_batch(p, b, eliminate_swing=True)
    assert out["status"][0] != 0 and np.all(out["grf"][0] == 0) and out["status"][1] == 0


def test_rollout_twin_closed_loop_is_stable(oracle):
    """numpy rollout twin (oracle/rollout_np.py): 8 trotting A1s for 60 control steps keep their
    commanded height and attitude; swing legs carry no force; phases advance by speed*dt."""
    from oracle import rollout_np
    p = presets.preset("northstar_a1")
    st = synth.make_rollout_state(8, seed=5)
    fr = st.pop("foot_rel")
    ph0 = st["phase"].copy()
    x_log, g_log = rollout_np.rollout(p, st, 60, fr)
    assert np.abs(x_log[-1][:, 5] - st["cmd"][:, 2]).max() < 0.03
    assert np.abs(x_log[:, :, 0:2]).max() < 0.2
    adv = (st["phase"] - ph0 - 60 * st["gait_speed"][:, None] * p.dt) % 1.0
    assert np.minimum(adv, 1.0 - adv).max() < 1e-9
    # trot: the diagonal pair in swing has exactly zero force in every step
    f = g_log.reshape(60, 8, 4, 3)
    assert ((np.abs(f[:, :, [0, 3]]).sum(axis=(2, 3)) == 0) | (np.abs(f[:, :, [1, 2]]).sum(axis=(2, 3)) == 0)).all()


def test_rollout_twin_plant_matches_step_matrices(oracle):
    """prepare() hands the solver exactly the state it integrates: foot_pos_abs = foot_world - pos,
    rot = Rz Ry Rx(euler) and the Raibert target of BaseInterface.cpp:358-399 for swing legs."""
    from oracle import rollout_np
    p = presets.preset("reference_a1")
    st = synth.make_rollout_state(4, seed=9)
    fr = st.pop("foot_rel")
    b = rollout_np.prepare(p, st, fr)
    assert np.allclose(b["foot_pos_abs"].reshape(4, 4, 3), st["foot_world"].reshape(4, 4, 3) - st["x"][:, None, 3:6])
    R = b["rot"].reshape(4, 3, 3).transpose(0, 2, 1)
    assert np.allclose(R @ R.transpose(0, 2, 1), np.eye(3)[None], atol=1e-14)
    i = 0
    c0 = b["contact"].reshape(4, p.horizon, 4)[i, 0]
    leg = int(np.nonzero(c0 == 0)[0][0])
    x = st["x"][i]
    cy, sy = np.cos(x[2]), np.sin(x[2])
    vd = np.array([cy * st["cmd"][i, 3] - sy * st["cmd"][i, 4], sy * st["cmd"][i, 3] + cy * st["cmd"][i, 4]])
    k = np.sqrt(abs(x[5]) / 9.8)
    delta = k * (x[9:11] - vd) + (1.0 / st["gait_speed"][i] / 2.0) / 2.0 * vd
    hip = np.array([cy * fr[3 * leg] - sy * fr[3 * leg + 1], sy * fr[3 * leg] + cy * fr[3 * leg + 1]])
    assert np.allclose(st["foot_world"][i, 3 * leg:3 * leg + 2], hip + delta + x[3:5])
