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
           "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"]
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
