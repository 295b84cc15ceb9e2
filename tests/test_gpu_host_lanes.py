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


def test_retry_pass_from_the_conservative_start():
    """The interior point starts from a small complementarity (KParams::z0_scale, LMPC_Z0_SCALE); what that pass
    cannot solve is handed to a second launch with the classical start.  A start far too small to work at all must
    therefore still return every minimiser — from the second pass."""
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    p = presets.preset("northstar_a1")
    b = synth.make_batch(2048, p.horizon, p.dt, 21)

    def run(scale):
        old = os.environ.get("LMPC_Z0_SCALE")
        os.environ["LMPC_Z0_SCALE"] = scale
        try:
            eng = ConvexMpcEngine(p, 2048, device=0)
        finally:
            if old is None:
                del os.environ["LMPC_Z0_SCALE"]
            else:
                os.environ["LMPC_Z0_SCALE"] = old
        out = eng.compute_grfs(b, want_U=False)
        st = eng.last_stats()
        eng.close()
        return out, st

    ref, st_ref = run("1")
    bad, st_bad = run("1e-300")
    assert (ref["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert (bad["status"] <= abi.PROB_OK_IPM_ONLY).all()
    scale = np.maximum(np.abs(ref["grf"]).max(axis=1, keepdims=True), 1.0)
    assert (np.abs(bad["grf"] - ref["grf"]) / scale).max() <= 1e-9
    assert st_bad["mean_factorizations"] > st_ref["mean_factorizations"]   # the wasted first attempts are counted
    dflt, _ = run("1e-4")
    assert (dflt["status"] <= abi.PROB_OK_IPM_ONLY).all()
    assert (np.abs(dflt["grf"] - ref["grf"]) / scale).max() <= 1e-9
