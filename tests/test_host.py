"""CPU tests of the host-side pieces: the C ABI surface of liblmpc.so (loads, exports every
declared symbol, struct layouts, presets, error convention without a GPU), the gait schedule in
its three restatements, the synthetic generator and the batch sharding (gloo, world_size 2)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from legged_mpc_control_b200 import _lib, abi, gait, presets, shard, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lmpc_[a-z0-9_]+)\s*\(", src)) - {"lmpc_handle"})


def test_library_exports_every_declared_symbol(lmpc_lib):
    names = _declared_functions("lmpc.h") + _declared_functions("lmpc_bench.h")
    assert len(names) >= 15
    for n in names:
        assert hasattr(lmpc_lib, n), f"liblmpc.so does not export {n}"
    assert lmpc_lib.lmpc_abi_version() == abi.ABI_VERSION


def test_struct_layout_and_presets(lmpc_lib):
    from legged_mpc_control_b200 import engine
    assert C.sizeof(abi.LmpcParams) == 8 * 4 + 6 * 8 + (9 + 13 + 12) * 8
    assert C.sizeof(abi.LmpcBatchIn) == 11 * 8 + 8 and C.sizeof(abi.LmpcBatchOut) == 5 * 8 + 8
    for name in presets.preset_names():
        assert bytes(engine.c_preset(name)) == bytes(presets.preset(name)), name
    with pytest.raises(engine.LmpcError):
        engine.c_preset("nope")
    ref = presets.preset("reference_a1")  # the reference's compile-time constants
    assert (ref.horizon, ref.dt, ref.mu, ref.fz_max, ref.discretisation) == (30, 0.01, 0.3, 180.0, abi.DISC_EULER)


def test_no_cpu_fallback_without_device(lmpc_lib):
    """On a box without a GPU the engine must fail loudly (no silent CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from legged_mpc_control_b200 import engine
    with pytest.raises(engine.LmpcError) as ei:
        engine.ConvexMpcEngine(presets.preset("northstar_a1"), 8)
    assert ei.value.code == abi.ERR_NO_DEVICE and "no CPU fallback" in str(ei.value)


def test_invalid_params_rejected(lmpc_lib):
    from legged_mpc_control_b200 import engine
    for field, val in (("horizon", 0), ("horizon", 33), ("dt", 0.0), ("mu", -1.0), ("precision", 7)):
        p = presets.preset("northstar_a1")
        setattr(p, field, val)
        with pytest.raises(engine.LmpcError) as ei:
            engine.ConvexMpcEngine(p, 8)
        assert ei.value.code == abi.ERR_INVALID_ARG
    p = presets.preset("northstar_a1"); p.r[3] = 0.0
    with pytest.raises(engine.LmpcError):
        engine.ConvexMpcEngine(p, 8)


def test_package_does_not_import_oracle():
    """The product package must not reach into oracle/ (checked in a fresh interpreter)."""
    code = "import sys; import legged_mpc_control_b200, legged_mpc_control_b200.engine; print(any(m.startswith('oracle') for m in sys.modules))"
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True)
    assert r.stdout.strip() == "False", r.stdout + r.stderr
    for dirpath, _, files in os.walk(os.path.join(ROOT, "legged_mpc_control_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read().replace("test oracle", ""), f


# ------------------------------------------------------------------ gait schedule
def test_gait_tables_three_restatements(oracle, lmpc_lib):
    """numpy (gait.py), C oracle and liblmpc's lmpc_gait_predict_contact agree on every gait, leg
    and a dense phase sweep incl. the wrap (LeggedContactFSM.cpp:280-294)."""
    rng = np.random.default_rng(1)
    for g in range(6):
        for leg in range(4):
            ph = np.concatenate([rng.random(200), [0.0, 0.1, 0.25, 0.5, 0.6, 0.75, 1.0]])
            dts = np.concatenate([rng.random(200) * 0.6, np.zeros(7)])
            a = gait.predict_contact(g, leg, ph, 3.5, dts)
            b = [oracle.gait_predict_contact(g, leg, x, 3.5, d) for x, d in zip(ph, dts)]
            c = [lmpc_lib.lmpc_gait_predict_contact(g, leg, x, 3.5, d) for x, d in zip(ph, dts)]
            assert list(a) == b == c


def test_trot_table_matches_reference_defaults():
    """Default gait = trot: legs {0,3} stance on [0,0.5], {1,2} on (0.5,1] (LeggedContactFSM.cpp:93-114)."""
    t = gait.contact_table(10, 0.025, [abi.GAIT_TROT], np.zeros((1, 4)), [3.5])[0]
    assert t[0].tolist() == [1, 0, 0, 1]
    # phase advances 3.5 * 0.025 = 0.0875 per step: switch after phase 0.5 -> step 6
    assert t[5].tolist() == [1, 0, 0, 1] and t[6].tolist() == [0, 1, 1, 0]
    s = gait.contact_table(30, 0.01, [abi.GAIT_STAND], np.zeros((1, 4)), [3.5])[0]
    assert s.all()


def test_contact_table_c_twin(lmpc_lib):
    from legged_mpc_control_b200 import engine
    rng = np.random.default_rng(2)
    B = 257
    g = rng.integers(0, 6, B); ph = rng.random((B, 4)); sp = 2 + 3 * rng.random(B)
    for H, dt in ((10, 0.025), (30, 0.01)):
        assert np.array_equal(gait.contact_table(H, dt, g, ph, sp), engine.gait_contact_table(H, dt, g, ph, sp))


# ------------------------------------------------------------------ synthetic inputs
def test_synth_is_counter_based():
    a = synth.make_batch(64, 10, 0.025, seed=3)
    b = synth.make_batch(32, 10, 0.025, seed=3, start=32)
    for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact"):
        assert np.array_equal(a[k][32:], b[k]), k
    c = synth.make_batch(64, 10, 0.025, seed=4)
    assert not np.array_equal(a["x0"], c["x0"])
    R = a["rot"].reshape(-1, 3, 3)
    assert np.abs(np.einsum("bij,bkj->bik", R, R) - np.eye(3)).max() < 1e-12  # orthonormal
    assert a["contact"].reshape(64, 10, 4).sum(axis=2).tolist() == [[2] * 10] * 64  # trot: two stance legs
    u = synth.u01(1, np.arange(200000), 7)
    assert 0 <= u.min() and u.max() < 1 and abs(u.mean() - 0.5) < 5e-3 and abs(u.var() - 1 / 12) < 2e-3


# ------------------------------------------------------------------ sharding
def test_shard_range_partitions():
    for total in (0, 1, 7, 65536, 1 << 20):
        for world in (1, 2, 3, 8):
            r = [shard.shard_range(total, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1


_GLOO_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from legged_mpc_control_b200 import shard, synth, presets
from oracle import binding as ob
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
p = presets.preset("northstar_a1")
B = 37  # uneven on purpose
full = synth.make_batch(B, p.horizon, p.dt, seed=5)
mine = shard.shard_batch(full, rank, world)
lo, hi = shard.shard_range(B, rank, world)
# each rank's shard equals regenerating exactly its index range from the counter-based stream
regen = synth.make_batch(hi - lo, p.horizon, p.dt, seed=5, start=lo)
assert all(np.array_equal(mine[k], regen[k]) for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact"))
out = ob.solve_batch(p, mine, eliminate_swing=True, nthreads=1)   # stand-in for the per-GPU engine
g = shard.gather_results(torch.from_numpy(out["grf"]), B, rank, world)
if rank == 0:
    ref = ob.solve_batch(p, full, eliminate_swing=True, nthreads=2)
    assert g.shape == (B, 12) and np.array_equal(g.numpy(), ref["grf"])
    print("GATHER_OK")
else:
    assert g is None
dist.destroy_process_group()
'''


def test_shard_and_gather_world2_gloo(tmp_path, oracle):
    """world_size-2 run of the N>1 path on CPU: shard by batch index, solve locally (the oracle
    stands in for the per-GPU engine — this tests the host logic only), gather to rank 0."""
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", str(script), ROOT],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GATHER_OK" in r.stdout


def test_bench_reference_arm_contract(oracle):
    """`bench.py --impl reference` (the oracle port on the host cores; the reference itself cannot be
    built here) prints ONE JSON line with the contract's keys; non-zero ranks print nothing."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-seconds", "0.5"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "convex_mpc_qp_solves_per_sec" and d["unit"] == "solves/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"] == {"value": d["value"], "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("a1_trot_h10")
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r1 = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r1.returncode == 0 and r1.stdout.strip() == ""


# ------------------------------------------------------------------ ABI layout
def test_ctypes_mirror_matches_library_layout(lmpc_lib):
    """sizeof and every offsetof of the ctypes structs in abi.py against what the built library
    reports for include/lmpc.h (lmpc_abi_sizeof): a drifted field would silently shift pointers."""
    import ctypes as C
    structs = [abi.LmpcParams, abi.LmpcBatchIn, abi.LmpcBatchOut, abi.LmpcRolloutState, abi.LmpcRolloutCfg, abi.LmpcRolloutResult,
               abi.LmpcFsmIo, abi.LmpcLegIo, abi.LmpcLegGeometry]
    for which, S in enumerate(structs):
        assert lmpc_lib.lmpc_abi_sizeof(which, -1) == C.sizeof(S), S.__name__
        for i, (name, *_rest) in enumerate(S._fields_):
            assert lmpc_lib.lmpc_abi_sizeof(which, i) == getattr(S, name).offset, (S.__name__, name)
        assert lmpc_lib.lmpc_abi_sizeof(which, len(S._fields_)) == -1  # no field the mirror does not know about
    assert lmpc_lib.lmpc_abi_sizeof(9, -1) == -1


def test_gait_predict_never_hangs():
    """the reference's `while (phase > 1) phase -= 1` never ends on inf / huge phase; the
    O(1) wrap returns the reference's own fall-through (STANCE) there and is bit-identical otherwise."""
    L = _lib.lib()
    for ph, sp in ((np.inf, 1.0), (np.nan, 1.0), (0.3, np.inf), (1e300, 1.0), (0.5, 1e300)):
        assert L.lmpc_gait_predict_contact(abi.GAIT_TROT, 1, float(ph), float(sp), 0.01) == 1
    rng = np.random.default_rng(0)
    for _ in range(2000):
        ph, sp, dt = rng.uniform(0, 1), rng.uniform(0.5, 400.0), rng.uniform(0, 0.5)
        x = ph + sp * dt
        while x > 1.0:
            x -= 1.0
        g, leg = int(rng.integers(0, 6)), int(rng.integers(0, 4))
        got = L.lmpc_gait_predict_contact(g, leg, ph, sp, dt)
        assert got == int(gait.predict_contact(g, leg, ph, sp, dt))           # python twin (loops like the reference)
        assert got == L.lmpc_gait_predict_contact(g, leg, x, 0.0, 0.0)        # the wrapped phase itself
