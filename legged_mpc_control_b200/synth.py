"""Synthetic robot-state batches, keyed by a counter-based splitmix64 stream.

Every value is a pure function of (seed, problem_index, field_id): batches can be generated in
any order, on any rank, in any chunking, and shards never overlap.  No dataset is read.
"""
import numpy as np

from . import abi, gait as gaitmod, presets

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
_C_SEED = np.uint64(0x9E3779B97F4A7C15)
_C_IDX = np.uint64(0xD1B54A32D192ED03)
_C_FLD = np.uint64(0x8CB92BA72F3D8DD7)


def _mix(z):
    z = np.asarray(z, np.uint64)
    with np.errstate(over="ignore"):
        z = (z + _C_SEED) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def u01(seed: int, idx, field: int):
    """Uniform double in [0,1) from the 53 high bits of splitmix64(key(seed, idx, field))."""
    idx = np.asarray(idx, np.uint64)
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) * _C_SEED + idx * _C_IDX + np.uint64(field) * _C_FLD) & _M64
    z = _mix(_mix(key))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def uniform(seed, idx, field, lo, hi):
    return lo + (hi - lo) * u01(seed, idx, field)


def normal(seed, idx, field, sigma):
    """Box-Muller on fields (field, field+1000)."""
    u1 = 1.0 - u01(seed, idx, field)  # (0,1]
    u2 = u01(seed, idx, field + 1000)
    return sigma * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


def _rot_zyx(roll, pitch, yaw):
    """R = Rz(yaw) Ry(pitch) Rx(roll); returns [B][3][3] (row, col)."""
    cr, sr, cp, sp, cy, sy = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    R = np.empty(roll.shape + (3, 3))
    R[..., 0, 0] = cy * cp
    R[..., 0, 1] = cy * sp * sr - sy * cr
    R[..., 0, 2] = cy * sp * cr + sy * sr
    R[..., 1, 0] = sy * cp
    R[..., 1, 1] = sy * sp * sr + cy * cr
    R[..., 1, 2] = sy * sp * cr - cy * sr
    R[..., 2, 0] = -sp
    R[..., 2, 1] = cp * sr
    R[..., 2, 2] = cp * cr
    return R


def make_batch(n: int, horizon: int, dt: float, seed: int, robot: str = "a1", gaits=(abi.GAIT_TROT,),
               start: int = 0, gait_speed=None, per_problem_params: bool = False, base_params=None):
    """Problems [start, start+n) of the stream `seed`.  Returns a dict of numpy arrays in the
    ABI's SoA layout (see include/lmpc.h lmpc_batch_in) plus 'gait' and 'phase' for reference.

    Distributions: roll,pitch~U(-.2,.2), yaw~U(-pi,pi), p_xy~U(-1,1),
    p_z~U(.22,.34), omega~N(0,.5^2), v_xy~U(-1,1), v_z~N(0,.1^2), feet = R (foothold + U(-.05,.05)^3),
    v_d,x~U(-2.5,2.5), v_d,y~U(-.4,.4), w_z,d~U(-.8,.8), z_d~U(.22,.30), gait phase~U(0,1).
    """
    idx = np.arange(start, start + n, dtype=np.uint64)
    f = lambda fld, lo, hi: uniform(seed, idx, fld, lo, hi)  # noqa: E731
    roll, pitch, yaw = f(1, -0.2, 0.2), f(2, -0.2, 0.2), f(3, -np.pi, np.pi)
    x0 = np.stack([
        roll, pitch, yaw,
        f(4, -1, 1), f(5, -1, 1), f(6, 0.22, 0.34),
        normal(seed, idx, 7, 0.5), normal(seed, idx, 8, 0.5), normal(seed, idx, 9, 0.5),
        f(10, -1, 1), f(11, -1, 1), normal(seed, idx, 12, 0.1),
    ], axis=1)
    R = _rot_zyx(roll, pitch, yaw)
    rot = np.ascontiguousarray(np.transpose(R, (0, 2, 1))).reshape(n, 9)  # column-major
    fh = np.asarray(presets.footholds(robot), np.float64)  # [4][3]
    jitter = np.stack([f(20 + k, -0.05, 0.05) for k in range(12)], axis=1).reshape(n, 4, 3)
    foot_rel = fh[None] + jitter
    foot_abs = np.einsum("bij,blj->bli", R, foot_rel)  # [B][leg][xyz] == 3x4 column-major
    cmd = np.stack([
        np.zeros(n), np.zeros(n), f(40, 0.22, 0.30),
        f(41, -2.5, 2.5), f(42, -0.4, 0.4), np.zeros(n),
        np.zeros(n), np.zeros(n), f(43, -0.8, 0.8),
    ], axis=1)
    gaits = list(gaits)
    gsel = np.minimum((u01(seed, idx, 50) * len(gaits)).astype(np.int64), len(gaits) - 1)
    gait = np.asarray(gaits, np.int32)[gsel]
    phase1 = u01(seed, idx, 51)
    phase = np.repeat(phase1[:, None], 4, axis=1)
    if gait_speed is None:
        gait_speed = presets.A1_GAIT_SPEED if robot == "a1" else presets.GO1_GAIT_SPEED
    if np.isscalar(gait_speed):
        gspeed = np.full(n, float(gait_speed))
    else:  # (lo, hi) range
        gspeed = f(52, gait_speed[0], gait_speed[1])
    contact = gaitmod.contact_table(horizon, dt, gait, phase, gspeed)
    out = {
        "x0": np.ascontiguousarray(x0), "rot": rot, "foot_pos_abs": np.ascontiguousarray(foot_abs.reshape(n, 12)),
        "cmd": np.ascontiguousarray(cmd), "contact": np.ascontiguousarray(contact.reshape(n, 4 * horizon)),
        "gait": gait, "phase": phase, "gait_speed": gspeed,
    }
    if per_problem_params:
        # C4 sweep: q,r x logU(0.1,10) per entry, mu~U(.2,.8)
        assert base_params is not None
        q0 = np.asarray(list(base_params.q)[:12])
        r0 = np.asarray(list(base_params.r))
        lq = np.stack([f(60 + k, -1.0, 1.0) for k in range(12)], axis=1)
        lr = np.stack([f(80 + k, -1.0, 1.0) for k in range(12)], axis=1)
        out["q"] = np.ascontiguousarray(q0[None] * 10.0 ** lq)
        out["r"] = np.ascontiguousarray(r0[None] * 10.0 ** lr)
        out["mu"] = np.ascontiguousarray(f(100, 0.2, 0.8))
    return out


def make_rollout_state(n: int, seed: int, robot: str = "a1", gaits=(abi.GAIT_TROT,), start: int = 0, gait_speed=None):
    """Initial closed-loop state of robots [start, start+n) for lmpc_rollout (config 5): level trunk
    at standing height with small random attitude/velocity, feet on the ground under the hips,
    random gait phase and joystick command.  Returns the dict lmpc_rollout_state expects."""
    idx = np.arange(start, start + n, dtype=np.uint64)
    f = lambda fld, lo, hi: uniform(seed, idx, fld, lo, hi)  # noqa: E731
    fh = np.asarray(presets.footholds(robot), np.float64)  # [4][3]
    z0 = -fh[0, 2]
    yaw = f(3, -np.pi, np.pi)
    x = np.stack([f(1, -0.05, 0.05), f(2, -0.05, 0.05), yaw, f(4, -1, 1), f(5, -1, 1), z0 + f(6, -0.02, 0.02),
                  normal(seed, idx, 7, 0.1), normal(seed, idx, 8, 0.1), normal(seed, idx, 9, 0.1),
                  f(10, -0.3, 0.3), f(11, -0.3, 0.3), normal(seed, idx, 12, 0.05)], axis=1)
    cy, sy = np.cos(yaw), np.sin(yaw)
    fw = np.zeros((n, 4, 3))
    for leg in range(4):
        fw[:, leg, 0] = x[:, 3] + cy * fh[leg, 0] - sy * fh[leg, 1]
        fw[:, leg, 1] = x[:, 4] + sy * fh[leg, 0] + cy * fh[leg, 1]
    cmd = np.stack([np.zeros(n), np.zeros(n), np.full(n, z0), f(41, -1.0, 1.0), f(42, -0.3, 0.3), np.zeros(n),
                    np.zeros(n), np.zeros(n), f(43, -0.8, 0.8)], axis=1)
    gaits = list(gaits)
    gsel = np.minimum((u01(seed, idx, 50) * len(gaits)).astype(np.int64), len(gaits) - 1)
    gait_arr = np.asarray(gaits, np.int32)[gsel]
    standing = gait_arr == abi.GAIT_STAND  # feet never relocate: a standing robot is commanded to stay
    cmd[standing, 3:6] = 0.0
    cmd[standing, 8] = 0.0
    x[standing, 9:11] *= 0.1
    if gait_speed is None:
        gait_speed = presets.A1_GAIT_SPEED if robot == "a1" else presets.GO1_GAIT_SPEED
    return {"x": np.ascontiguousarray(x), "foot_world": np.ascontiguousarray(fw.reshape(n, 12)),
            "phase": np.ascontiguousarray(np.repeat(u01(seed, idx, 51)[:, None], 4, axis=1)),
            "gait": np.ascontiguousarray(gait_arr),
            "gait_speed": np.full(n, float(gait_speed)), "cmd": np.ascontiguousarray(cmd),
            "foot_rel": np.ascontiguousarray(fh.reshape(12))}


def standing_problem(horizon: int, robot: str = "a1", z: float = 0.28, contact_all: bool = True):
    """Deterministic C1 state: euler=0, pos=(0,0,z), zero velocities, feet at the yaml footholds,
    z_d=z (fixture idea from the reference's test_wbc.cpp:126-129: f_z = m g / 4 per leg)."""
    x0 = np.zeros((1, 12)); x0[0, 5] = z
    rot = np.eye(3).reshape(1, 9)
    foot = np.asarray(presets.footholds(robot), np.float64).reshape(1, 12)
    cmd = np.zeros((1, 9)); cmd[0, 2] = z
    contact = np.ones((1, 4 * horizon), np.uint8) if contact_all else None
    return {"x0": x0, "rot": rot, "foot_pos_abs": foot, "cmd": cmd, "contact": contact}
