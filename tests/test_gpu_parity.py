"""GPU parity tests: liblmpc.so (through the C ABI) against the CPU oracle on identical seeded
inputs.  Tolerances (BASELINE.json north_star): FP64 GRFs within 1e-6 relative, same active set
(rules in tests/parity_util.py)."""

import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from legged_mpc_control_b200 import abi, presets, synth

pytestmark = pytest.mark.gpu

from parity_util import REL_TOL, compare, norm_err as _rel_err  # noqa: E402


def _engine(params, max_batch):
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    return ConvexMpcEngine(params, max_batch, device=0)


def _check(params, batch, oracle, max_rel=REL_TOL, label=None):
    """CUDA path (through the C ABI) vs the oracle on the same inputs: element-wise 1e-6 on every
    component of grf and U, 1e-9 norm-wise regression guard on exactly polished answers, identical
    primal-active bitmask (ties within a decade of active_tol counted, reported and bounded)."""
    eng = _engine(params, len(batch["x0"]))
    got = eng.compute_grfs(batch)
    ref = oracle.solve_batch(params, batch, eliminate_swing=True)
    assert ref["failures"] == 0
    rep = compare(params, got, ref, batch, rel_tol=max_rel)
    print(f"[parity] {label or ''} {rep}")
    eng.close()
    return got, ref


def test_c1_single_solve_both_presets(oracle):
    """configs[0]: A1 trot single solve, H=10 (north-star) and the reference preset H=30."""
    for name in ("northstar_a1", "reference_a1"):
        p = presets.preset(name)
        H = p.horizon
        b = synth.standing_problem(H)
        b["cmd"][0, 3] = 0.5  # v_d = (0.5, 0, 0)
        from legged_mpc_control_b200 import gait
        b["contact"] = gait.contact_table(H, p.dt, [abi.GAIT_TROT], np.zeros((1, 4)), [presets.A1_GAIT_SPEED]).reshape(1, -1)
        got, ref = _check(p, b, oracle)
        # swing legs (FR, RL at phase 0) get exact zeros (SURVEY KAT K2)
        assert np.all(got["grf"][0, 3:9] == 0.0)


def test_standing_known_answer(oracle):
    """KAT K1: symmetric stand at the reference height: fx = fy = 0 and four equal fz near m g / 4."""
    p = presets.preset("northstar_a1")
    b = synth.standing_problem(p.horizon)
    got, _ = _check(p, b, oracle)
    f = got["grf"][0].reshape(4, 3)
    assert np.abs(f[:, :2]).max() < 1e-8
    assert np.ptp(f[:, 2]) < 1e-8
    assert abs(f[0, 2] - p.mass * p.gravity / 4) < 1.0


def test_c2_batch_4096_a1_trot(oracle):
    """configs[1]: A1 trot batch 4096 randomized states, H=10."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(4096, p.horizon, p.dt, seed=1)
    got, _ = _check(p, b, oracle)
    assert (got["status"] == abi.PROB_OK).mean() > 0.99


def test_c3_go1_mixed_gaits_h20(oracle):
    """configs[2] at oracle-sized batch: Go1, trot/bound/pronk, H=20."""
    p = presets.preset("northstar_go1_h20")
    b = synth.make_batch(512, p.horizon, p.dt, seed=2, robot="go1",
                         gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK))
    _check(p, b, oracle)


def test_reference_preset_h30_all_gaits(oracle):
    """The reference's own formulation (Euler, H=30, dt=0.01) over its four gait tables,
    including stand (n = 360, matrices spill to the global workspace)."""
    p = presets.preset("reference_a1")
    b = synth.make_batch(96, p.horizon, p.dt, seed=4,
                         gaits=(abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND))
    _check(p, b, oracle)


def test_all_swing_and_ragged_batches(oracle):
    """Edge cases: pronk flight (every leg swings for the whole horizon -> U = 0, KAT K3), batch
    of one, and a batch that is not a multiple of anything."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(37, p.horizon, p.dt, seed=9, gaits=(abi.GAIT_TROT, abi.GAIT_PRONK))
    b["contact"][0, :] = 0
    b["contact"][5, :] = 0
    got, _ = _check(p, b, oracle)
    assert np.all(got["U"][0] == 0) and np.all(got["U"][5] == 0)
    assert got["status"][0] == 0


def test_per_problem_parameters(oracle):
    """configs[3] shape: per-problem q, r, mu sweeps (C4)."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(256, p.horizon, p.dt, seed=4, gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK),
                         per_problem_params=True, base_params=p, gait_speed=(2.0, 4.5))
    _check(p, b, oracle)


def test_golden_fixtures(oracle):
    """Committed oracle outputs (tests/golden/make_golden.py) reproduced by the CUDA path."""
    gdir = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gdir) if f.endswith(".npz") and not f.startswith("golden_adapter"))  # (the adapter-tick fixture has its own test)
    assert files
    for f in files:
        g = np.load(os.path.join(gdir, f))
        p = presets.preset(str(g["preset"]))
        batch = {k: g[k] for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
        eng = _engine(p, len(batch["x0"]))
        got = eng.compute_grfs(batch)
        assert _rel_err(got["grf"], g["grf"]).max() <= REL_TOL
        assert _rel_err(got["U"], g["U"]).max() <= REL_TOL
        assert np.array_equal(got["active_set"], g["active_set"])
        eng.close()


def test_headline_batch_properties():
    """BASELINE.json's full size (batch 65536, H=10) through size-independent properties:
    (a) every GRF is primal feasible, (b) swing legs are exact zeros, (c) cost-scaling invariance
    (REF_HALF vs NS_TWO give the same U, KAT K10), (d) yaw equivariance (KAT K5): rotating state,
    feet and commands about z rotates the GRFs."""
    p = presets.preset("northstar_a1")
    B = 65536
    b = synth.make_batch(B, p.horizon, p.dt, seed=3)
    eng = _engine(p, B)
    got = eng.compute_grfs(b)
    assert (got["status"] <= 1).all()
    f = got["U"].reshape(B, p.horizon, 4, 3)
    c = b["contact"].reshape(B, p.horizon, 4).astype(bool)
    assert np.all(f[~c] == 0.0)
    tol = 1e-7
    assert (np.abs(f[..., 0]) <= p.mu * f[..., 2] + tol).all() and (np.abs(f[..., 1]) <= p.mu * f[..., 2] + tol).all()
    assert (f[..., 2] >= -tol).all() and (f[..., 2] <= p.fz_max + tol).all()
    # (c) scaling invariance on a slice
    p2 = p.copy(); p2.cost_scaling = abi.COST_REF_HALF
    sl = {k: v[:4096] for k, v in b.items()}
    eng2 = _engine(p2, 4096)
    got2 = eng2.compute_grfs(sl)
    assert _rel_err(got2["U"], got["U"][:4096]).max() <= REL_TOL
    # (d) yaw equivariance on a slice
    psi = 0.7
    cz, sz = np.cos(psi), np.sin(psi)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    r = {k: np.array(v[:4096], copy=True) for k, v in b.items()}
    x0 = r["x0"]
    x0[:, 2] += psi
    x0[:, 3:6] = x0[:, 3:6] @ Rz.T
    x0[:, 6:9] = x0[:, 6:9] @ Rz.T
    x0[:, 9:12] = x0[:, 9:12] @ Rz.T
    R = r["rot"].reshape(-1, 3, 3).transpose(0, 2, 1)
    r["rot"] = np.ascontiguousarray((Rz[None] @ R).transpose(0, 2, 1)).reshape(-1, 9)
    r["foot_pos_abs"] = np.ascontiguousarray(r["foot_pos_abs"].reshape(-1, 4, 3) @ Rz.T).reshape(-1, 12)
    got3 = eng2.compute_grfs(r)  # body-frame commands are unchanged by a world yaw rotation
    f3 = got3["U"].reshape(4096, p.horizon, 4, 3)
    f2 = got2["U"].reshape(4096, p.horizon, 4, 3) @ Rz.T
    # the friction pyramid is not rotation invariant: compare only problems with no friction row active
    H = p.horizon
    fric_words = got2["active_set"][:, : (16 * H) // 32]
    stance_fric = np.zeros(4096, bool)
    for i in range(4096):
        bits = np.unpackbits(fric_words[i].view(np.uint8), bitorder="little")[: 16 * H].reshape(H, 4, 4)
        stance_fric[i] = bits[c[i]].any()
    free = ~stance_fric
    bits3 = got3["active_set"][:, : (16 * H) // 32]
    for i in np.nonzero(free)[0]:
        b3 = np.unpackbits(bits3[i].view(np.uint8), bitorder="little")[: 16 * H].reshape(H, 4, 4)
        if b3[c[i]].any():
            free[i] = False
    if free.any():
        err = np.abs(f3[free] - f2[free]).reshape(int(free.sum()), -1).max(axis=1)
        assert (err / np.maximum(1.0, np.abs(f2[free]).reshape(int(free.sum()), -1).max(axis=1))).max() <= 1e-6
    eng.close(); eng2.close()


def test_error_behaviour():
    """ABI error convention: invalid args are reported, never thrown across; failed problems
    return zeros + status (reference NaN rule, ConvexQPSolver.cpp:321-326)."""
    from legged_mpc_control_b200.engine import ConvexMpcEngine, LmpcError
    p = presets.preset("northstar_a1")
    eng = ConvexMpcEngine(p, 8)
    b = synth.make_batch(16, p.horizon, p.dt, seed=1)
    with pytest.raises(LmpcError):
        eng.compute_grfs(b)  # batch > max_batch
    b = synth.make_batch(4, p.horizon, p.dt, seed=1)
    b["x0"][1, 0] = np.nan
    got = eng.compute_grfs(b)
    assert got["rc"] == abi.ERR_NOT_CONVERGED
    assert got["status"][1] >= abi.PROB_MAX_ITER and np.all(got["grf"][1] == 0)
    assert (got["status"][[0, 2, 3]] == 0).all()
    bad = p.copy(); bad.horizon = 99
    with pytest.raises(LmpcError):
        ConvexMpcEngine(bad, 8)
    eng.close()


def test_device_pointer_path_and_launch_count():
    """Device-resident path (torch only for memory): same answer as the host-pointer path."""
    import torch
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    p = presets.preset("northstar_a1")
    B = 2048
    b = synth.make_batch(B, p.horizon, p.dt, seed=11)
    eng = ConvexMpcEngine(p, B)
    host = eng.compute_grfs(b)
    dev = torch.device("cuda:0")
    tin = {k: torch.from_numpy(b[k]).to(dev) for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
    grf = torch.zeros(B, 12, dtype=torch.float64, device=dev)
    status = torch.full((B,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    n0 = eng.launch_count
    eng.solve_device({k: v.data_ptr() for k, v in tin.items()}, {"grf": grf.data_ptr(), "status": status.data_ptr()}, B)
    assert eng.launch_count - n0 >= 2
    assert np.array_equal(grf.cpu().numpy(), host["grf"])
    assert (status.cpu().numpy() == host["status"]).all()
    st = eng.last_stats()
    assert abs(st["mean_n_eff"] - 60.0) < 1e-9 and st["mean_factorizations"] > 5
    assert eng.last_kernel_ms > 0
    eng.close()
