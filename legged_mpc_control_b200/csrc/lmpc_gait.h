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

}  // namespace lmpc
