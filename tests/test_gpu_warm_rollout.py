"""GPU tests of the warm-started solve (lmpc_set_warm_start; OSQP warm-start analogue,
ConvexQPSolver.cpp:185) and the closed-loop batched rollout (lmpc_rollout; BASELINE config 5),
through the C ABI, against the CPU oracle / its numpy rollout twin."""
import numpy as np
import pytest

from legged_mpc_control_b200 import abi, presets, synth

pytestmark = pytest.mark.gpu

REL_TOL = 1e-6  # north_star: FP64 GRFs within 1e-6 relative of the exact solution


def _engine(params, max_batch):
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    return ConvexMpcEngine(params, max_batch, device=0)


def _rel_err(a, b):
    scale = np.maximum(1.0, np.abs(b).max(axis=1))
    return np.abs(a - b).max(axis=1) / scale


@pytest.mark.parametrize("preset,gaits", [("northstar_a1", (abi.GAIT_TROT,)),
                                          ("northstar_go1_h20", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK)),
                                          ("reference_a1", (abi.GAIT_TROT, abi.GAIT_CRAWL, abi.GAIT_STAND))])
def test_warm_start_gives_the_same_exact_solution(oracle, preset, gaits):
    """Second solve of a slightly moved state, warm-started from the first one's faces: the answer
    is still the oracle's exact minimiser (1e-6 rel., same active set) and most problems are solved
    without a single interior-point iteration."""
    p = presets.preset(preset)
    B = 512 if p.horizon <= 20 else 128
    b0 = synth.make_batch(B, p.horizon, p.dt, seed=11, gaits=gaits, robot="go1" if "go1" in preset else "a1")
    eng = _engine(p, B)
    eng.set_warm_start(True, horizon_shift=0)
    cold = eng.compute_grfs(b0)
    assert eng.last_warm_hits == 0
    cold_fact = cold["factorizations"].mean()
    b1 = dict(b0)
    rng = np.random.default_rng(0)
    b1["x0"] = b0["x0"] + 1e-3 * rng.standard_normal(b0["x0"].shape)
    warm = eng.compute_grfs(b1)
    ref = oracle.solve_batch(p, b1, eliminate_swing=True)
    assert ref["failures"] == 0
    assert (warm["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert _rel_err(warm["grf"], ref["grf"]).max() <= REL_TOL
    assert _rel_err(warm["U"], ref["U"]).max() <= REL_TOL
    hits = eng.last_warm_hits
    nz = int((b1["contact"].reshape(B, -1).sum(axis=1) > 0).sum())
    assert hits >= 0.6 * nz, (hits, nz)
    assert warm["factorizations"].mean() < 0.6 * cold_fact
    # the same problem again: every face verifies at once
    again = eng.compute_grfs(b1)
    assert np.abs(again["grf"] - warm["grf"]).max() <= 1e-9 * max(1.0, np.abs(warm["grf"]).max())
    # reset forgets the faces
    eng.reset_warm_start()
    eng.compute_grfs(b1)
    assert eng.last_warm_hits == 0
    # a wrong guess must not hurt: faces of a different batch
    b2 = synth.make_batch(B, p.horizon, p.dt, seed=12, gaits=gaits, robot="go1" if "go1" in preset else "a1")
    got2 = eng.compute_grfs(b2)
    ref2 = oracle.solve_batch(p, b2, eliminate_swing=True)
    assert (got2["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert _rel_err(got2["grf"], ref2["grf"]).max() <= REL_TOL
    eng.close()


def test_rollout_matches_cpu_twin(oracle):
    """16 robots x 30 control steps, trot/stand/bound: device rollout == numpy twin driven by the
    oracle's exact QP solver (states to 1e-7, GRFs to 1e-6 relative)."""
    from oracle import rollout_np
    p = presets.preset("northstar_a1")
    B, steps = 16, 30
    st = synth.make_rollout_state(B, seed=5, gaits=(abi.GAIT_TROT, abi.GAIT_STAND, abi.GAIT_BOUND))
    fr = st.pop("foot_rel")
    ref_state = {k: v.copy() for k, v in st.items()}
    x_ref, g_ref = rollout_np.rollout(p, ref_state, steps, fr)
    eng = _engine(p, B)
    out = eng.rollout(st, steps, fr, warm_start=1, log=True)
    assert out["rc"] == abi.OK and out["not_converged"] == 0 and out["solves"] == B * steps
    gscale = np.maximum(1.0, np.abs(g_ref).max(axis=2, keepdims=True))
    assert (np.abs(out["grf_log"] - g_ref) / gscale).max() <= REL_TOL
    assert np.abs(out["x_log"] - x_ref).max() <= 1e-7
    for k in ("x", "foot_world", "phase"):
        assert np.abs(st[k] - ref_state[k]).max() <= 1e-7, k
    assert np.abs(out["grf"] - g_ref[-1]).max() <= REL_TOL * gscale.max()
    assert out["warm_hits"] > 0.5 * out["solves"]
    eng.close()


def test_rollout_warm_equals_cold_and_reference_preset():
    """Warm start changes the work, not the trajectory (reference preset, H=30, dt=0.01)."""
    p = presets.preset("reference_a1")
    B, steps = 64, 20
    runs = []
    for warm in (0, 1):
        st = synth.make_rollout_state(B, seed=6, gaits=(abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL))
        fr = st.pop("foot_rel")
        eng = _engine(p, B)
        out = eng.rollout(st, steps, fr, warm_start=warm, log=True)
        assert out["rc"] == abi.OK and out["not_converged"] == 0
        runs.append((st, out))
        eng.close()
    (s0, o0), (s1, o1) = runs
    assert o0["warm_hits"] == 0 and o1["warm_hits"] > 0.5 * o1["solves"]
    assert o1["factorizations"] < 0.5 * o0["factorizations"]
    assert np.abs(o0["x_log"] - o1["x_log"]).max() <= 1e-8
    assert np.abs(o0["grf_log"] - o1["grf_log"]).max() <= 1e-6 * max(1.0, np.abs(o0["grf_log"]).max())
    # the robots keep standing/trotting: height stays near the commanded one
    assert np.abs(o1["x_log"][-1][:, 5] - s1["cmd"][:, 2]).max() < 0.05


def test_rollout_device_pointers_and_errors():
    import torch
    p = presets.preset("northstar_a1")
    B, steps = 256, 8
    st = synth.make_rollout_state(B, seed=7)
    fr = st.pop("foot_rel")
    host = {k: v.copy() for k, v in st.items()}
    eng = _engine(p, B)
    h = eng.rollout(host, steps, fr, warm_start=1)
    dev = {k: torch.from_numpy(v).cuda() for k, v in st.items()}
    grf = torch.zeros(B, 12, dtype=torch.float64, device="cuda")
    ptrs = {k: v.data_ptr() for k, v in dev.items()}
    ptrs.update(batch=B, grf=grf.data_ptr())
    n0 = eng.launch_count
    d = eng.rollout(ptrs, steps, fr, warm_start=1, on_device=True)
    # prepare + classify + solve_warp + solve_tile<15> + step per tick (a warm-started engine solves in ONE pass from the
    # classical interior-point start; the cold closed loop enqueues the retry pass too: 7 launches per tick)
    assert eng.launch_count - n0 == steps * 5
    assert d["solves"] == h["solves"] == B * steps
    assert np.abs(dev["x"].cpu().numpy() - host["x"]).max() <= 1e-12
    assert np.abs(grf.cpu().numpy() - h["grf"]).max() <= 1e-9
    from legged_mpc_control_b200.engine import LmpcError
    with pytest.raises(LmpcError):
        eng.rollout(host, 0, fr)
    n1 = eng.launch_count
    eng.rollout(ptrs, 2, fr, warm_start=0, on_device=True)
    assert eng.launch_count - n1 == 2 * 7
    eng.close()


def test_warm_equals_cold_at_headline_batch():
    """Full 65 536-problem headline batch: a warm-started solve of a moved state returns the same
    minimiser as a cold solve of it (size-independent property; both KKT-verified on device)."""
    p = presets.preset("northstar_a1")
    B = 65536
    b0 = synth.make_batch(B, p.horizon, p.dt, seed=3)
    b1 = dict(b0)
    b1["x0"] = b0["x0"] + 1e-3 * np.random.default_rng(1).standard_normal(b0["x0"].shape)
    eng = _engine(p, B)
    cold = eng.compute_grfs(b1, want_U=False)
    eng.set_warm_start(True, horizon_shift=0)
    eng.compute_grfs(b0, want_U=False)
    warm = eng.compute_grfs(b1, want_U=False)
    assert eng.last_warm_hits > 0.99 * B
    assert (warm["status"] <= abi.PROB_OK_IPM_ONLY).all() and (cold["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert _rel_err(warm["grf"], cold["grf"]).max() <= REL_TOL
    exact = (warm["status"] == 0) & (cold["status"] == 0)
    # same primal-active rows, up to rows whose residual sits within rounding of active_tol (the
    # two paths reach the minimiser from different sides; the oracle parity tests treat that band
    # the same way): at most a handful of problems per 65 536
    differ = int((warm["active_set"][exact] != cold["active_set"][exact]).any(axis=1).sum())
    assert differ <= 1e-3 * B, differ
    assert warm["factorizations"].mean() < 0.15 * cold["factorizations"].mean()
    eng.close()


def test_active_set_solver_enum(oracle):
    """LMPC_SOLVER_ACTIVE_SET: exact active-set iteration first (interior faces on a cold problem,
    stored faces afterwards), interior point as fallback — same minimiser as the oracle."""
    p = presets.preset("northstar_a1")
    p.solver = abi.SOLVER_ACTIVE_SET
    B = 1024
    b = synth.make_batch(B, p.horizon, p.dt, seed=21)
    eng = _engine(p, B)
    got = eng.compute_grfs(b)
    ref = oracle.solve_batch(p, b, eliminate_swing=True)
    assert ref["failures"] == 0 and (got["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert _rel_err(got["grf"], ref["grf"]).max() <= REL_TOL
    assert _rel_err(got["U"], ref["U"]).max() <= REL_TOL
    cold_hits = eng.last_warm_hits  # problems the active-set iteration solved from the interior faces alone
    assert 0 < cold_hits < B
    eng.set_warm_start(True, horizon_shift=0)  # same problems again (creation set the closed-loop shift of 1)
    again = eng.compute_grfs(b)
    assert eng.last_warm_hits > 0.99 * B and again["factorizations"].mean() < 1.2
    assert _rel_err(again["grf"], ref["grf"]).max() <= REL_TOL
    eng.close()
