"""Generates tests/golden/golden_adapter_tick.npz FROM THE REFERENCE'S OWN CONTROLLER TICK (run here,
where /root/reference is mounted: `python tests/golden/make_golden_adapter.py`).

A scripted sequence of 50 ticks (joystick command, a slowly drifting state, footholds, early-contact
flags, one movement_mode-0 reset, one estimation-not-inited tick) is fed to the reference's unmodified
ConvexMpc::update (oracle/_ref; ConvexMpc.cpp:24-62 with LeggedContactFSM::update, the command
filters and ConvexQPSolver) with the exact minimiser of the QP IT built supplied where OSQP would
answer.  The file holds the scenario (the bytes oracle/_ref/adapter_ref reads) and everything the tick
wrote into LeggedState, so that on the GPU box include/lmpc_legged_adapter.hpp — compiled against the
same reference headers — can be checked tick by tick against numbers the reference's code produced."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from legged_mpc_control_b200 import presets, synth  # noqa: E402
from oracle import binding as ob, ref_binding as rb  # noqa: E402

H, T = 30, 50


def solve_captured(Pd, q, A, l, u):
    Hm, g, Cm, lo, up, X0, S = rb.condense_captured(Pd, q, A, l, u)
    rows, rhs = [], []
    for k in range(Cm.shape[0]):
        if lo[k] > -1e29:
            rows.append(Cm[k]); rhs.append(lo[k])
        if up[k] < 1e29:
            rows.append(-Cm[k]); rhs.append(-up[k])
    rc, U, lam, it = ob.gi_dense(Hm, g, np.array(rows), np.array(rhs))
    assert rc == 0
    X = X0 + S @ U
    z = np.zeros(24 * H)
    for k in range(H):
        z[24 * k:24 * k + 12] = U[12 * k:12 * k + 12]
        z[24 * k + 12:24 * k + 24] = X[12 * k:12 * k + 12]
    return z


def main():
    assert rb.build() and rb.horizon() == H
    yaml = rb.yaml_path("gazebo_a1_convex.yaml")
    prm = rb.load_params(yaml)
    q, r, mass, inertia, gspeed = prm["q"], prm["r"], prm["mass"], prm["inertia"], prm["gait_counter_speed"]
    p = presets.preset("reference_a1")
    b = synth.make_batch(1, H, p.dt, seed=77)
    x0, rot, foot_abs = b["x0"][0].copy(), b["rot"][0].copy(), b["foot_pos_abs"][0].copy()
    rng = np.random.default_rng(5)
    foot_world = (foot_abs.reshape(4, 3) + x0[3:6]).reshape(12)
    tick = rb.Tick(yaml)
    rows_in, rows_out = [], []
    for k in range(T):
        t, dt = 0.01 * (k + 1), 0.01
        movement_mode = 0.0 if k == 30 else 1.0      # one FSM reset mid-run (ConvexMpc.cpp:86-93)
        inited = 0.0 if k == 0 else 1.0              # the estimation guard (ConvexMpc.cpp:27-31)
        joy = np.array([0.6 + 0.2 * np.sin(0.3 * k), 0.1, 0.2, 0.27])
        rp = np.array([0.01, -0.02])
        xk = x0.copy(); xk[3] += 0.005 * k; xk[9] = 0.5; xk[8] = 0.2; xk[2] += 0.002 * k  # drifting state (rot kept: R need not match yaw for the MPC)
        tgt = foot_world + np.array([0.05, 0.0, 0.0] * 4) + 0.002 * rng.standard_normal(12)
        cf = (rng.uniform(size=4) < 0.15).astype(np.float64)  # occasional early contact (LeggedContactFSM.cpp:58-63)
        out = tick.update(t, dt, xk, rot, foot_abs, foot_world, tgt, joy=joy, contact_flag=cf, roll_pitch_d=rp,
                          movement_mode=movement_mode, estimation_inited=bool(inited), solver=solve_captured)
        assert out["rc"] == 0
        rows_in.append(np.concatenate([[t, dt, movement_mode, inited], joy, rp, xk, rot, foot_abs, foot_world, tgt, cf]))
        rows_out.append(np.concatenate([out["optimized_state"], out["optimized_input"], out["plan_contacts"].astype(np.float64),
                                        out["v_d_world"], out["euler_d"], out["lin_vel_d_rel"], [float(out["mpc_solver_inited"])]]))
    hdr = np.concatenate([q, r, [mass], np.asarray(inertia).reshape(-1), [gspeed]])
    np.savez_compressed(os.path.join(HERE, "golden_adapter_tick.npz"), T=np.int32(T), header=hdr, ticks_in=np.array(rows_in),
                        ticks_out=np.array(rows_out))
    print("wrote golden_adapter_tick.npz:", np.array(rows_in).shape, np.array(rows_out).shape,
          "stance pattern changes:", int(np.abs(np.diff(np.array(rows_out)[:, 42:46], axis=0)).sum()))


if __name__ == "__main__":
    main()
