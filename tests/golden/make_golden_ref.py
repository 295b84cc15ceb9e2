"""Generates tests/golden/golden_refbuild_a1_h30.npz FROM THE REFERENCE'S OWN CODE (run here, where
/root/reference is mounted: `python tests/golden/make_golden_ref.py`).

For 48 seeded A1 states over the reference's four gait tables, oracle/_ref (the unmodified
ConvexQPSolver.cpp / LeggedContactFSM.cpp compiled against oracle/shim) builds the QP it would hand to
OSQP; that captured (P, q, A, l, u) is condensed with its own equality rows and minimised exactly.  The
file holds the inputs, that minimiser (U, grf), the primal-active bitmask and a fingerprint of the captured
matrices, so that on a box without /root/reference both the oracle (CPU suite) and the CUDA path (GPU
suite) are still checked against numbers the reference's code produced."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from legged_mpc_control_b200 import abi, presets, synth  # noqa: E402
from oracle import binding as ob, ref_binding as rb  # noqa: E402

H = 30
GAITS = (abi.GAIT_TROT, abi.GAIT_TROT_WITH_STAND, abi.GAIT_CRAWL, abi.GAIT_STAND)


def solve_captured(cap):
    Hm, g, Cm, lo, up, X0, S = rb.condense_captured(cap["Pdiag"], cap["q"], cap["A"], cap["l"], cap["u"])
    rows, rhs = [], []
    for k in range(Cm.shape[0]):
        if lo[k] > -1e29:
            rows.append(Cm[k]); rhs.append(lo[k])
        if up[k] < 1e29:
            rows.append(-Cm[k]); rhs.append(-up[k])
    rc, U, lam, it = ob.gi_dense(Hm, g, np.array(rows), np.array(rhs))
    assert rc == 0
    return U


def main():
    p = presets.preset("reference_a1")
    B = 48
    b = synth.make_batch(B, H, p.dt, seed=104, gaits=GAITS)
    U = np.zeros((B, 12 * H)); fp = np.zeros((B, 4))
    for i in range(B):
        cap = rb.qp_build(list(p.q)[:12], list(p.r), p.mass, list(p.inertia), b["x0"][i], b["rot"][i], b["foot_pos_abs"][i],
                          b["cmd"][i], b["contact"][i].reshape(H, 4)[0], b["gait"][i], b["phase"][i], b["gait_speed"][i])
        U[i] = solve_captured(cap)
        fp[i] = [cap["A"].sum(), np.abs(cap["A"]).sum(), cap["q"].sum(), cap["u"][:12 * H].sum()]
    orc = ob.solve_batch(p, b, eliminate_swing=True)
    assert orc["failures"] == 0
    err = np.abs(U - orc["U"]).max()
    assert err < 1e-8, err
    np.savez_compressed(os.path.join(HERE, "golden_refbuild_a1_h30.npz"), preset="reference_a1", seed=104,
                        x0=b["x0"], rot=b["rot"], foot_pos_abs=b["foot_pos_abs"], cmd=b["cmd"], contact=b["contact"],
                        U=U, grf=U[:, :12].copy(), active_set=orc["active_set"], qp_fingerprint=fp, source="oracle/_ref")
    print("golden_refbuild_a1_h30.npz ok: max|U_ref - U_oracle| =", err)


if __name__ == "__main__":
    main()
