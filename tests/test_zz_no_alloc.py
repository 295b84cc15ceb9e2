"""lmpc_solve_noalloc (include/lmpc.h): the allocation-free entry point, called through the C ABI and
checked against the existing entry point (lmpc_solve) on the same inputs — two calls in a row on the
shared static scratch, different batches and sizes, so the second call runs on the first call's
leftovers.  Tolerance: the parity tests' own (tests/parity_util.py: 1e-6 element-wise, same status,
same active set).  Collected last on purpose."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from legged_mpc_control_b200 import abi, presets, synth  # noqa: E402
from parity_util import REL_TOL, elem_err  # noqa: E402

pytestmark = pytest.mark.gpu

IN_KEYS = ("x0", "rot", "foot_pos_abs", "cmd", "contact")


def _device_io(torch, b, H, dev):
    B = len(b["x0"])
    tin = {k: torch.from_numpy(np.ascontiguousarray(b[k])).to(dev) for k in IN_KEYS}
    # outputs start as garbage so that every element must be written by the call
    tout = {
        "grf": torch.full((B, 12), 7.0, dtype=torch.float64, device=dev),
        "status": torch.full((B,), -1, dtype=torch.int32, device=dev),
        "iters": torch.full((B,), -1, dtype=torch.int32, device=dev),
        "active_set": torch.full((B, abi.active_words(H)), -1, dtype=torch.int32, device=dev),
        "U": torch.full((B, 12 * H), 7.0, dtype=torch.float64, device=dev),
    }
    return tin, tout


def _call(eng, tin, tout, B, **kw):
    from legged_mpc_control_b200 import _lib
    st = eng.solve_device_noalloc({k: v.data_ptr() for k, v in tin.items()}, {k: v.data_ptr() for k, v in tout.items()}, B, **kw)
    if st != abi.OK:
        print(f"[noalloc] status {st}: {_lib.last_error()}")
    return st


def _check(tout, ref, label):
    grf = tout["grf"].cpu().numpy()
    U = tout["U"].cpu().numpy()
    st = tout["status"].cpu().numpy()
    act = tout["active_set"].cpu().numpy().view(np.uint32)
    # lmpc_solve starts its interior point from a small complementarity and re-solves what that leaves in a second
    # launch; the allocation-free entry point is ONE launch from the conservative start (no retry list to allocate).
    # Both end in the same exact polish: same answers, possibly different iteration counts / exact-vs-interior status.
    assert np.array_equal(st <= abi.PROB_OK_IPM_ONLY, ref["status"] <= abi.PROB_OK_IPM_ONLY), f"{label}: status differs"
    assert (st <= abi.PROB_OK_IPM_ONLY).all()
    eg, eu = elem_err(grf, ref["grf"]).max(), elem_err(U, ref["U"]).max()
    print(f"[noalloc] {label}: grf err {eg:.2e} U err {eu:.2e} bit-identical {np.array_equal(grf, ref['grf'])}")
    assert eg <= REL_TOL and eu <= REL_TOL, f"{label}: grf {eg:.3e} U {eu:.3e} > {REL_TOL}"
    both = (st == 0) & (ref["status"] == 0)  # exactly polished by both: the same vertex of the same face
    assert both.mean() > 0.99
    assert np.array_equal(act[both], np.asarray(ref["active_set"]).view(np.uint32)[both]), f"{label}: active set differs"
    assert (tout["iters"].cpu().numpy() > 0).all()


def test_noalloc_matches_lmpc_solve_on_two_calls_in_a_row():
    import torch
    from legged_mpc_control_b200.engine import ConvexMpcEngine
    dev = torch.device("cuda:0")
    p = presets.preset("northstar_a1")
    H = p.horizon
    b1 = synth.make_batch(3000, H, p.dt, seed=21)  # trot at H = 10: n = 60 for every phase, the one-warp class
    b2 = synth.make_batch(1111, H, p.dt, seed=22)
    for b in (b1, b2):
        assert (3 * b["contact"].reshape(len(b["x0"]), -1).astype(np.int64).sum(axis=1) <= 63).all()
    eng = ConvexMpcEngine(p, 4096)
    ref1 = eng.compute_grfs(b1)  # existing entry point: lmpc_solve, host pointers
    ref2 = eng.compute_grfs(b2)
    io1, io2 = _device_io(torch, b1, H, dev), _device_io(torch, b2, H, dev)
    warm = _device_io(torch, b2, H, dev)
    assert _call(eng, *warm, 1111) == abi.OK  # set-up call: symbol address, module load
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    n0 = eng.launch_count
    assert _call(eng, *io1, 3000) == abi.OK   # call 1
    assert _call(eng, *io2, 1111) == abi.OK   # call 2, on call 1's leftovers in the scratch
    assert eng.launch_count - n0 == 2          # one kernel per call
    assert torch.cuda.mem_get_info(dev)[0] == free0, "device memory in use changed across the allocation-free calls"
    _check(io1[1], ref1, "call 1 (3000)")
    _check(io2[1], ref2, "call 2 (1111)")
    # third call: non-blocking on a caller's stream, the caller synchronises
    io3 = _device_io(torch, b1, H, dev)
    s = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    assert _call(eng, *io3, 3000, stream=s.cuda_stream, blocking=False) == abi.OK
    s.synchronize()
    _check(io3[1], ref1, "call 3 (async)")
    eng.close()


def test_noalloc_limits():
    import torch
    from legged_mpc_control_b200 import _lib
    from legged_mpc_control_b200.engine import ConvexMpcEngine, LmpcError
    dev = torch.device("cuda:0")
    p = presets.preset("northstar_a1")
    H = p.horizon
    eng = ConvexMpcEngine(p, 8192)
    # beyond the stated maximum: refused with a clear message, nothing launched
    big = synth.make_batch(abi.NOALLOC_MAX_BATCH + 1, H, p.dt, seed=23)
    tin, tout = _device_io(torch, big, H, dev)
    n0 = eng.launch_count
    with pytest.raises(LmpcError) as ei:
        _call(eng, tin, tout, abi.NOALLOC_MAX_BATCH + 1)
    assert ei.value.code == abi.ERR_INVALID_ARG and "LMPC_NOALLOC_MAX_BATCH" in str(ei.value)
    assert eng.launch_count == n0
    assert (tout["status"].cpu().numpy() == -1).all()
    # the maximum itself is accepted
    assert _call(eng, tin, tout, abi.NOALLOC_MAX_BATCH) == abi.OK
    # host pointers are refused (staging them would allocate)
    hb = synth.make_batch(4, H, p.dt, seed=24)
    sin, _keep = abi.host_batch_in(hb, H)[:2]
    sout = abi.LmpcBatchOut()
    grf = np.zeros((4, 12))
    sout.grf = grf.ctypes.data
    sout.on_device = 0
    import ctypes as C
    st = _lib.lib().lmpc_solve_noalloc(eng._h, 4, C.byref(sin), C.byref(sout), None, 1)
    assert st == abi.ERR_INVALID_ARG and "device pointers" in _lib.last_error()
    # a problem above the one-warp limit (stand at H = 10: n = 120) is refused per problem, the others are solved
    mix = synth.make_batch(64, H, p.dt, seed=25, gaits=(abi.GAIT_TROT, abi.GAIT_STAND))
    n = 3 * mix["contact"].reshape(64, -1).astype(np.int64).sum(axis=1)
    assert (n > 63).any() and (n <= 63).any()
    ref = eng.compute_grfs(mix)
    tin, tout = _device_io(torch, mix, H, dev)
    assert _call(eng, tin, tout, 64) == abi.ERR_NOT_CONVERGED
    st = tout["status"].cpu().numpy()
    grf = tout["grf"].cpu().numpy()
    assert (st[n > 63] == abi.PROB_TOO_LARGE).all() and (grf[n > 63] == 0).all()
    assert (tout["U"].cpu().numpy()[n > 63] == 0).all() and (tout["active_set"].cpu().numpy()[n > 63] == 0).all()
    assert (st[n <= 63] <= abi.PROB_OK_IPM_ONLY).all() and (ref["status"][n <= 63] <= abi.PROB_OK_IPM_ONLY).all()
    assert elem_err(grf[n <= 63], ref["grf"][n <= 63]).max() <= REL_TOL
    eng.close()
