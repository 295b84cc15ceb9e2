"""Full-size GPU parity: every BASELINE.json config at (or near) its quoted batch, CUDA path through
the C ABI against the CPU oracle on the same seeded inputs — not through properties.  The oracle
solves these batches in seconds on the box's host cores (one problem per thread).  Rules and
tolerances: tests/parity_util.py (element-wise 1e-6 on every component, 1e-9 regression guard on
exactly polished answers, identical active set with counted ties, every status-1 answer checked
individually)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from legged_mpc_control_b200 import abi, presets, synth  # noqa: E402
from parity_util import compare  # noqa: E402

pytestmark = pytest.mark.gpu


def _run(params, batch, oracle, label, elem_floor_U=1e-3):
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    B = len(batch["x0"])
    eng = ConvexMpcEngine(params, B, device=0)
    got = eng.compute_grfs(batch)
    stats = eng.last_stats()
    eng.close()
    ref = oracle.solve_batch(params, batch, eliminate_swing=True, want_kkt=False)
    assert ref["failures"] == 0
    rep = compare(params, got, ref, batch, elem_floor_U=elem_floor_U)
    rep["mean_n_eff"] = stats["mean_n_eff"]
    print(f"[parity-full] {label}: {rep}")
    return got, ref, rep


def test_headline_65536_vs_oracle(oracle):
    """The metric's own batch (A1 trot, H=10, 65 536 problems, seed 3 — bench.py's workload)."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(65536, p.horizon, p.dt, seed=3)
    got, _, rep = _run(p, b, oracle, "headline 65536")
    assert (got["status"] == abi.PROB_OK).mean() > 0.999


def test_c3_go1_h20_16384_vs_oracle(oracle):
    """config 3: Go1 trot / bound / pronk, H=20 (tile classes 15 and 24), 16 384 problems."""
    p = presets.preset("northstar_go1_h20")
    b = synth.make_batch(16384, p.horizon, p.dt, seed=2, robot="go1",
                         gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK))
    _run(p, b, oracle, "c3 16384")


@pytest.mark.parametrize("gait,name", [(abi.GAIT_TROT, "trot"), (abi.GAIT_TROT_WITH_STAND, "trot_with_stand"),
                                        (abi.GAIT_CRAWL, "crawl"), (abi.GAIT_STAND, "stand")])
def test_reference_preset_h30_1024_per_gait(oracle, gait, name):
    """The reference's own operating point (Euler, H=30, dt=0.01): 1024 problems of EACH of its four
    gait tables, so size classes 2 (n=180), 4 (n<=288) and 3 (n up to 360) each see hundreds."""
    p = presets.preset("reference_a1")
    b = synth.make_batch(1024, p.horizon, p.dt, seed=40 + gait, gaits=(gait,))
    # stand at H=30 is n = 360 with the yaml weights (no position weight): cond(H) ~ 1e8, so two exact
    # FP64 solvers agree to ~1e-9 of the force scale on late-horizon forces, not better: the element
    # floor of the full-horizon U is 1e-2 ||ref||_inf there.  The first-step GRFs keep 1e-3 everywhere.
    _, _, rep = _run(p, b, oracle, f"ref_h30 {name} 1024", elem_floor_U=1e-2 if gait == abi.GAIT_STAND else 1e-3)


def test_c4_sweep_65536_vs_oracle(oracle):
    """config 4 shape at 65 536: per-problem q, r x 10^U(-1,1), mu in [0.2, 0.8], three gaits."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(65536, p.horizon, p.dt, seed=4, gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK),
                         per_problem_params=True, base_params=p, gait_speed=(2.0, 4.5))
    # weights swept over two decades push cond(H) to 2e8 on some problems: U floor 1e-2 (measured worst
    # U component at the 1e-3 floor 2.5e-6, an interior-point-only answer); GRFs keep 1e-3
    _run(p, b, oracle, "c4 sweep 65536", elem_floor_U=1e-2)
