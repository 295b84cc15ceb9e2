"""Generates the committed golden fixtures from the CPU oracle (run here, in the build
container: `python tests/golden/make_golden.py`).  Each .npz holds the inputs of a small batch
and the oracle's exact outputs (U, grf, active set).  The reference has no fixtures for this
path, so these pin the oracle-vs-CUDA comparison and guard against regressions of
the oracle itself; they are cross-checked against the independent numpy twin (KKT certificates)
in tests/test_oracle.py."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from legged_mpc_control_b200 import abi, presets, synth  # noqa: E402
from oracle import binding as ob  # noqa: E402

CASES = [
    # file, preset, batch, seed, robot, gaits
    ("golden_ns_a1_trot.npz", "northstar_a1", 32, 101, "a1", (abi.GAIT_TROT,)),
    ("golden_ns_go1_h20_mixed.npz", "northstar_go1_h20", 16, 102, "go1", (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK)),
    ("golden_ref_a1_h30.npz", "reference_a1", 12, 103, "a1", (abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND)),
]


def main():
    for fname, preset, B, seed, robot, gaits in CASES:
        p = presets.preset(preset)
        b = synth.make_batch(B, p.horizon, p.dt, seed=seed, robot=robot, gaits=gaits)
        full = ob.solve_batch(p, b, eliminate_swing=False)   # the reference's full row set
        red = ob.solve_batch(p, b, eliminate_swing=True)
        assert full["failures"] == 0 and red["failures"] == 0
        assert np.abs(full["U"] - red["U"]).max() < 1e-7
        assert np.array_equal(full["active_set"], red["active_set"])
        assert full["kkt"][:, 0].max() < 1e-9 and full["kkt"][:, 1].max() < 1e-9
        np.savez_compressed(os.path.join(HERE, fname), preset=preset, seed=seed,
                            x0=b["x0"], rot=b["rot"], foot_pos_abs=b["foot_pos_abs"], cmd=b["cmd"],
                            contact=b["contact"], U=red["U"], grf=red["grf"], active_set=red["active_set"])
        print(fname, "ok", "max|U|", np.abs(red["U"]).max())


if __name__ == "__main__":
    main()
