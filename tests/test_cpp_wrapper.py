"""include/lmpc.hpp (C++ RAII face of the C ABI) and include/lmpc_legged_adapter.hpp
(ConvexMpcB200 : ConvexMpc, compiled against the field-name mock in tests/cpp/mock_reference/):
the C++ host side builds with g++ alone, refuses to run without a GPU, and on a GPU returns the
oracle's GRFs for the standing-trot tick it drives through LeggedMPC::update."""
import os
import subprocess

import numpy as np
import pytest

from legged_mpc_control_b200 import abi, gait, presets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def wrapper_exe(lmpc_lib, tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("cpp") / "test_wrapper")
    libdir = os.path.join(ROOT, "legged_mpc_control_b200")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", f"-I{ROOT}/include", f"-I{ROOT}/tests/cpp/mock_reference",
           f"{ROOT}/tests/cpp/test_wrapper.cpp", "-o", exe, f"-L{libdir}", "-llmpc", f"-Wl,-rpath,{libdir}",
           "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64", "-pthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_cpp_wrapper_builds_and_has_no_cpu_fallback(wrapper_exe):
    if _has_gpu():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([wrapper_exe], capture_output=True, text=True)
    assert r.returncode == 3, (r.returncode, r.stdout, r.stderr)  # lmpc::Error(LMPC_ERR_NO_DEVICE)
    assert "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_adapter_update_matches_oracle(wrapper_exe, oracle):
    r = subprocess.run([wrapper_exe], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("grf")][0]
    got = np.array([float(v) for v in line.split()[1:]])
    p = presets.preset("reference_a1")
    H = p.horizon
    x0 = np.zeros((1, 12)); x0[0, 5] = 0.3
    cmd = np.zeros((1, 9)); cmd[0, 2] = 0.3
    batch = {"x0": x0, "rot": np.eye(3).reshape(1, 9), "foot_pos_abs": np.asarray(presets.A1_FOOTHOLDS).reshape(1, 12),
             "cmd": cmd, "contact": gait.contact_table(H, p.dt, [abi.GAIT_TROT], np.zeros((1, 4)), [3.5]).reshape(1, 4 * H)}
    ref = oracle.solve_batch(p, batch, eliminate_swing=True)
    assert ref["failures"] == 0
    assert np.abs(got - ref["grf"][0]).max() <= 1e-6 * max(1.0, np.abs(ref["grf"]).max())
    assert got[3:9].tolist() == [0.0] * 6  # swing legs


@pytest.mark.gpu
def test_sharded_engine_matches_oracle(wrapper_exe, oracle, tmp_path):
    """lmpc::ShardedEngine (one engine + host thread per shard, contiguous index ranges, results in
    disjoint slices of one output array): 3 shards over a ragged batch of 1001 problems."""
    from legged_mpc_control_b200 import synth
    p = presets.preset("northstar_a1")
    B, H = 1001, p.horizon
    b = synth.make_batch(B, H, p.dt, seed=31, gaits=(abi.GAIT_TROT, abi.GAIT_PRONK, abi.GAIT_STAND))
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        f.write(np.array([B, H], np.int32).tobytes())
        for k in ("x0", "rot", "foot_pos_abs", "cmd"):
            f.write(np.ascontiguousarray(b[k], np.float64).tobytes())
        f.write(np.ascontiguousarray(b["contact"], np.uint8).tobytes())
    r = subprocess.run([wrapper_exe, "shard", str(fin), str(fout), "northstar_a1", "3"], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "shards 3" in r.stdout and "overflow_rejected 1" in r.stdout and "unconverged 0" in r.stdout
    raw = open(fout, "rb").read()
    grf = np.frombuffer(raw[:B * 96], np.float64).reshape(B, 12)
    status = np.frombuffer(raw[B * 96:], np.int32)
    ref = oracle.solve_batch(p, b, eliminate_swing=True)
    assert ref["failures"] == 0 and (status >= 0).all() and (status <= abi.PROB_OK_IPM_ONLY).all()
    scale = np.maximum(1.0, np.abs(ref["grf"]).max(axis=1))
    assert (np.abs(grf - ref["grf"]).max(axis=1) / scale).max() <= 1e-6
