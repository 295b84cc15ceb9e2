// lmpc_gait.h — gait tables and the contact prediction rule, shared by the host ABI
// (lmpc_host.cpp) and the device gait scheduler (lmpc_rollout.cuh).  Reference citations are
// relative to /root/reference/src/legged_ctrl/.
#pragma once
#include <math.h>

#include "../../include/lmpc.h"

#ifdef __CUDACC__
#define LMPC_HD __host__ __device__
#else
#define LMPC_HD
#endif

namespace lmpc {

// Gait tables: (state, switch_time) per leg.  1 = STANCE, 0 = SWING (LeggedContactFSM.h:11-15).
struct GaitLeg { int n; int state[3]; double sw[3]; };

LMPC_HD inline GaitLeg gait_table(int gait, int leg) {
  GaitLeg t = {1, {1, 1, 1}, {1.0, 1.0, 1.0}};
  const bool diagA = (leg == 0 || leg == 3);
  switch (gait) {
    case LMPC_GAIT_TROT:  // set_default_gait_pattern, LeggedContactFSM.cpp:93-114
      t.n = 2; t.sw[0] = 0.5; t.sw[1] = 1.0;
      t.state[0] = diagA ? 1 : 0; t.state[1] = diagA ? 0 : 1;
      break;
    case LMPC_GAIT_TROT_WITH_STAND:  // :116-157
      if (diagA) { t.n = 2; t.state[0] = 1; t.state[1] = 0; t.sw[0] = 0.6; t.sw[1] = 1.0; }
      else { t.n = 3; t.state[0] = 1; t.state[1] = 0; t.state[2] = 1; t.sw[0] = 0.1; t.sw[1] = 0.5; t.sw[2] = 1.0; }
      break;
    case LMPC_GAIT_CRAWL:  // :158-199
      if (leg == 0) { t.n = 2; t.state[0] = 0; t.state[1] = 1; t.sw[0] = 0.25; t.sw[1] = 1.0; }
      else if (leg == 1) { t.n = 3; t.state[0] = 1; t.state[1] = 0; t.state[2] = 1; t.sw[0] = 0.25; t.sw[1] = 0.5; t.sw[2] = 1.0; }
      else if (leg == 2) { t.n = 3; t.state[0] = 1; t.state[1] = 0; t.state[2] = 1; t.sw[0] = 0.5; t.sw[1] = 0.75; t.sw[2] = 1.0; }
      else { t.n = 2; t.state[0] = 1; t.state[1] = 0; t.sw[0] = 0.75; t.sw[1] = 1.0; }
      break;
    case LMPC_GAIT_STAND:  // :201-212
      break;
    case LMPC_GAIT_BOUND: {
      const bool front = (leg == 0 || leg == 1);
      t.n = 2; t.sw[0] = 0.5; t.sw[1] = 1.0;
      t.state[0] = front ? 1 : 0; t.state[1] = front ? 0 : 1;
      break;
    }
    default:  // LMPC_GAIT_PRONK
      t.n = 2; t.sw[0] = 0.5; t.sw[1] = 1.0; t.state[0] = 1; t.state[1] = 0;
      break;
  }
  return t;
}


// LeggedContactFSM::predict_contact_state, src/utils/LeggedContactFSM.cpp:280-294.
LMPC_HD inline int gait_predict(int gait, int leg, double phase, double gait_speed, double delta_t) {
  const GaitLeg t = gait_table(gait, leg);
#ifdef __CUDA_ARCH__
  double ph = __dadd_rn(phase, __dmul_rn(gait_speed, delta_t));  // :281, unfused like the host build
#else
  double ph = phase + gait_speed * delta_t;  // :281
#endif
  // :283-285 `while (ph > 1) ph -= 1`.  Every one of those subtractions is exact in binary floating
  // point (the difference lies on ph's own ulp grid), so the loop equals ph - (ceil(ph) - 1) bit for
  // bit — evaluated in O(1).  Where the reference's loop would never end (non-finite phase or speed;
  // |ph| beyond 2^52, where ph - 1 == ph) the row falls through to its `should not reach here`
  // STANCE instead of hanging the kernel.
  if (!(ph == ph) || ph > 4503599627370496.0) return 1;
  if (ph > 1.0) ph -= ceil(ph) - 1.0;
  for (int i = 0; i < t.n; ++i)              // :287-291
    if (ph <= t.sw[i]) return t.state[i];
  return 1;                                  // :293 "should not reach here" -> STANCE
}

// ---- the contact state machine itself: LeggedContactFSM::reset (:16-36), update (:38-84),
// common_enter (:214-229), percent_in_state (:269-278); the swing-foot Bezier targets are not part of
// it.  Kept quirk for quirk (host mirror: legged_mpc_control_b200/gait.py ContactFsm, both checked
// against the compiled reference): `update` TOGGLES stance/swing at every transition instead of
// reading the pattern entry; the phase is brought back by 1 only when the pattern index wraps, so it
// is NOT confined to [0, 1] between transitions (predict_contact_state wraps its own copy).
struct FsmLeg { double phase, start, end; int index, state; };

LMPC_HD inline void fsm_reset(int gait, int leg, FsmLeg& f) {
  const GaitLeg t = gait_table(gait, leg);
  f.phase = 0.0; f.index = 0; f.start = 0.0; f.end = t.sw[0]; f.state = t.state[0];
}
LMPC_HD inline double fsm_percent(const FsmLeg& f) {
  double p = (f.phase - f.start) / (f.end - f.start);  // :270 (0/0 -> NaN falls through both clamps, as in the reference)
  if (p < 0.0) p = 0.0;
  else if (p > 1.0) p = 1.0;
  return p;
}
LMPC_HD inline void fsm_enter(const GaitLeg& t, FsmLeg& f) {  // common_enter
  const int prev = f.index;
  f.index = (f.index + 1) % t.n;
  if (f.index < prev) f.phase -= 1.0;
  f.start = f.phase;
  f.end = t.sw[f.index];
}
// one LeggedContactFSM::update(dt, ., ., foot_force_flag); returns the new contact state (1 = STANCE)
LMPC_HD inline int fsm_update(int gait, int leg, double gait_speed, double dt, bool force_flag, FsmLeg& f) {
  const GaitLeg t = gait_table(gait, leg);
#ifdef __CUDA_ARCH__
  f.phase = __dadd_rn(f.phase, __dmul_rn(gait_speed, dt));  // :50, unfused like the host build
#else
  f.phase += gait_speed * dt;
#endif
  if (f.state == 1) {
    if (f.phase >= f.end) { fsm_enter(t, f); f.state = 0; }  // :53-58
  } else {
    const double pct = fsm_percent(f);
    if ((pct > 0.9 && force_flag) || pct >= 1.0) { f.state = 1; fsm_enter(t, f); }  // :59-71
  }
  return f.state;
}

}  // namespace lmpc
