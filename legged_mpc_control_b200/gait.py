"""Gait tables and the horizon contact schedule (numpy, vectorised over robots).

Mirrors LeggedContactFSM::predict_contact_state (/root/reference/src/legged_ctrl/src/utils/
LeggedContactFSM.cpp:280-294) and the gait tables of :93-212.  BOUND and PRONK are additions
(the reference has trot / trot-with-stand / crawl / stand only).  The C twin is
lmpc_gait_contact_table in csrc/lmpc_host.cpp; tests compare the two.
"""
import numpy as np

from . import abi

# per gait, per leg: list of (state, switch_time); state 1 = STANCE, 0 = SWING
_TROT = {0: [(1, 0.5), (0, 1.0)], 3: [(1, 0.5), (0, 1.0)], 1: [(0, 0.5), (1, 1.0)], 2: [(0, 0.5), (1, 1.0)]}
_TROT_STAND = {0: [(1, 0.6), (0, 1.0)], 3: [(1, 0.6), (0, 1.0)],
               1: [(1, 0.1), (0, 0.5), (1, 1.0)], 2: [(1, 0.1), (0, 0.5), (1, 1.0)]}
_CRAWL = {0: [(0, 0.25), (1, 1.0)], 1: [(1, 0.25), (0, 0.5), (1, 1.0)],
          2: [(1, 0.5), (0, 0.75), (1, 1.0)], 3: [(1, 0.75), (0, 1.0)]}
_STAND = {leg: [(1, 1.0)] for leg in range(4)}
_BOUND = {0: [(1, 0.5), (0, 1.0)], 1: [(1, 0.5), (0, 1.0)], 2: [(0, 0.5), (1, 1.0)], 3: [(0, 0.5), (1, 1.0)]}
_PRONK = {leg: [(1, 0.5), (0, 1.0)] for leg in range(4)}
GAIT_TABLES = {abi.GAIT_TROT: _TROT, abi.GAIT_TROT_WITH_STAND: _TROT_STAND, abi.GAIT_CRAWL: _CRAWL,
               abi.GAIT_STAND: _STAND, abi.GAIT_BOUND: _BOUND, abi.GAIT_PRONK: _PRONK}


def predict_contact(gait: int, leg: int, phase, gait_speed, delta_t):
    """Vectorised predict_contact_state: arrays broadcast; returns uint8 (1 = stance)."""
    ph = np.asarray(phase, np.float64) + np.asarray(gait_speed, np.float64) * np.asarray(delta_t, np.float64)
    ph = np.array(ph, np.float64, copy=True)
    # `while (p > 1.0) p -= 1.0`  (:283-285) — repeated subtraction, not fmod, to match bit for bit
    while True:
        m = ph > 1.0
        if not m.any():
            break
        ph[m] -= 1.0
    out = np.ones(ph.shape, np.uint8)  # "should not reach here" -> STANCE (:293)
    done = np.zeros(ph.shape, bool)
    for state, sw in GAIT_TABLES[gait][leg]:
        hit = (~done) & (ph <= sw)
        out[hit] = state
        done |= hit
    return out


def contact_table(horizon: int, dt: float, gait, phase, gait_speed):
    """contact[B][H][4]: row i = predict_contact_state(i*dt) per leg (row 0: delta_t = 0).

    gait: [B] int, phase: [B][4] (per-leg gait phase in [0,1]), gait_speed: [B].
    Follows ConvexQPSolver::update_bound_constraints (ConvexQPSolver.cpp:329-346)."""
    gait = np.asarray(gait, np.int32).reshape(-1)
    B = gait.shape[0]
    phase = np.asarray(phase, np.float64).reshape(B, 4)
    gait_speed = np.asarray(gait_speed, np.float64).reshape(B)
    out = np.zeros((B, horizon, 4), np.uint8)
    steps = np.arange(horizon, dtype=np.float64) * dt  # i * mpc_dt_  (:342)
    for g in np.unique(gait):
        sel = np.nonzero(gait == g)[0]
        for leg in range(4):
            out[sel, :, leg] = predict_contact(int(g), leg, phase[sel, leg][:, None],
                                               gait_speed[sel][:, None], steps[None, :])
    return out


class ContactFsm:
    """Host mirror of the LeggedContactFSM state machine that drives the contact schedule
    (LeggedContactFSM.cpp:16-36 reset, :38-84 update, :214-229 common_enter, :269-278 percent_in_state);
    the swing-foot Bezier targets are not part of it.  Kept quirk for quirk: `update` TOGGLES stance/swing
    at every transition instead of reading the pattern entry, so the three-entry legs of trot-with-stand /
    crawl run their (STANCE, SWING, STANCE) table with alternating meaning, and the one-entry `stand`
    pattern — whose index never wraps, so its phase is never brought back (:218-221) — leaves stance at
    phase 1 for good.  The reference only ever activates the trot pattern (LeggedContactFSM.cpp:10).  The
    device twin is fsm_update in csrc/lmpc_gait.h; tests compare both with the compiled reference."""

    def __init__(self, gait: int, leg: int, gait_speed: float):
        self.table = GAIT_TABLES[gait][leg]
        self.speed = float(gait_speed)
        self.reset()

    def reset(self):
        self.phase = 0.0
        self.index = 0
        self.start = 0.0
        self.end = self.table[0][1]
        self.state = self.table[0][0]

    def percent(self):
        with np.errstate(divide="ignore", invalid="ignore"):
            p = np.float64(self.phase - self.start) / np.float64(self.end - self.start)
        if p < 0.0:
            p = np.float64(0.0)
        elif p > 1.0:
            p = np.float64(1.0)
        return float(p)

    def _enter(self):
        prev = self.index
        self.index = (self.index + 1) % len(self.table)
        if self.index < prev:
            self.phase -= 1.0
        self.start = self.phase
        self.end = self.table[self.index][1]

    def update(self, dt: float, force_flag: bool = False):
        self.phase += self.speed * dt
        if self.state == 1:
            if self.phase >= self.end:
                self._enter()
                self.state = 0
        else:
            pct = self.percent()
            if (pct > 0.9 and force_flag) or pct >= 1.0:
                self.state = 1
                self._enter()
        return self.phase

    def predict(self, delta_t: float) -> int:
        ph = self.phase + self.speed * delta_t
        while ph > 1.0:
            ph -= 1.0
        for state, sw in self.table:
            if ph <= sw:
                return state
        return 1
