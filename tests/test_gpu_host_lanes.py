"""Large host-pointer batches are solved in slices that overlap their copies with the neighbouring slice's
solve (lmpc_api.cu solve_host_lanes).  Every problem is independent and goes through the same kernels, so
the sliced call must return exactly what the one-shot call returns."""

import os

import numpy as np
import pytest

from legged_mpc_control_b200 import abi, presets, synth

pytestmark = pytest.mark.gpu


def _solve(params, batch, lanes):
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    old = os.environ.get("LMPC_HOST_LANES")
    os.environ["LMPC_HOST_LANES"] = str(lanes)  # read once, at lmpc_create
    try:
        eng = ConvexMpcEngine(params, len(batch["x0"]), device=0)
    finally:
        if old is None:
            del os.environ["LMPC_HOST_LANES"]
        else:
            os.environ["LMPC_HOST_LANES"] = old
    n0 = eng.launch_count
    out = eng.compute_grfs(batch, want_U=True)
    st = eng.last_stats()
    launches = eng.launch_count - n0
    ms = eng.last_kernel_ms
    eng.close()
    return out, st, launches, ms


@pytest.mark.parametrize("name,robot,gaits,nb", [
    ("northstar_a1", "a1", (abi.GAIT_TROT,), 16384 + 37),
    ("northstar_go1_h20", "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), 8192 + 5),
])
def test_sliced_host_batch_equals_one_shot(name, robot, gaits, nb):
    p = presets.preset(name)
    b = synth.make_batch(nb, p.horizon, p.dt, 11, robot=robot, gaits=gaits)
    one, st1, l1, ms1 = _solve(p, b, 1)
    for lanes in (4, 3):
        got, st, ln, ms = _solve(p, b, lanes)
        for k in ("grf", "U", "status", "iters", "active_set"):
            assert np.array_equal(np.asarray(got[k]), np.asarray(one[k])), (lanes, k)
        for k in ("mean_n_eff", "mean_factorizations", "flops_executed"):  # statistics are summed over the slices
            assert abs(st[k] - st1[k]) <= 1e-9 * abs(st1[k]), (lanes, k, st[k], st1[k])
        assert ln >= lanes * 2 and ln > l1, (ln, l1)   # classify + at least one solve kernel per slice
        assert ms > 0.0


def test_small_and_warm_batches_stay_one_shot():
    """Below the slice threshold, and with warm start on (the face records are keyed by batch index inside ONE
    engine), the call takes the one-shot path: the launch count is that of a single engine."""
    p = presets.preset("northstar_a1")
    b = synth.make_batch(4096, p.horizon, p.dt, 5)
    _, _, l_small, _ = _solve(p, b, 4)
    _, _, l_one, _ = _solve(p, b, 1)
    assert l_small == l_one
