// lmpc_host.cpp — host-only parts of the C ABI: parameter presets and the gait schedule.
// No CUDA here.  Reference citations are relative to /root/reference/src/legged_ctrl/.
#include <string.h>

#include "../../include/lmpc.h"

namespace {

// config/gazebo_a1_convex.yaml:40-56 and config/gazebo_go1_convex.yaml:39-55
const double kA1Q[13] = {60.0, 100.0, 0.0, 0.0, 0.0, 450.0, 0.15, 0.15, 100.0, 3.0, 3.0, 5.0, 0.0};
const double kGo1Q[13] = {50.0, 100.0, 0.0, 0.0, 0.0, 3500.0, 0.01, 0.01, 10.0, 15.0, 15.0, 20.0, 0.0};
// a1_trunk_inertia, config/gazebo_a1_convex.yaml:135-140 (Go1 falls back to the same loader
// defaults, src/LeggedState.cpp:146-160)
const double kA1Inertia[9] = {0.0158533, 0.0, 0.0, 0.0, 0.0377999, 0.0, 0.0, 0.0, 0.0456542};

void base(lmpc_params* p, int H, double dt, int disc, int scal, int off, const double* q) {
  memset(p, 0, sizeof(*p));
  p->horizon = H;
  p->discretisation = disc;
  p->cost_scaling = scal;
  p->xref_index_offset = off;
  p->precision = LMPC_PREC_F64;
  p->solver = LMPC_SOLVER_IPM;
  p->max_iter = 0;
  p->dt = dt;
  p->mu = 0.3;        // ConvexQPSolver.cpp:25
  p->fz_max = 180.0;  // ConvexQPSolver.cpp:171
  p->mass = 13.0;     // a1_robot_mass, config/gazebo_a1_convex.yaml:133
  p->gravity = 9.8;   // ConvexQPSolver.cpp:175
  p->active_tol = 1e-6;
  memcpy(p->inertia, kA1Inertia, sizeof(kA1Inertia));
  memcpy(p->q, q, sizeof(double) * 13);
  for (int i = 0; i < 12; ++i) p->r[i] = 1e-4;  // r_weights_*, yaml :58-72
}

// Gait tables: (state, switch_time) per leg.  1 = STANCE, 0 = SWING (LeggedContactFSM.h:11-15).
struct GaitLeg { int n; int state[3]; double sw[3]; };

GaitLeg gait_table(int gait, int leg) {
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

}  // namespace

extern "C" lmpc_status lmpc_params_preset(const char* name, lmpc_params* p) {
  if (!name || !p) return LMPC_ERR_INVALID_ARG;
  if (!strcmp(name, "reference_a1")) base(p, 30, 0.01, LMPC_DISC_EULER, LMPC_COST_REF_HALF, 0, kA1Q);
  else if (!strcmp(name, "reference_go1")) base(p, 30, 0.01, LMPC_DISC_EULER, LMPC_COST_REF_HALF, 0, kGo1Q);
  else if (!strcmp(name, "northstar_a1")) base(p, 10, 0.025, LMPC_DISC_ZOH, LMPC_COST_NS_TWO, 1, kA1Q);
  else if (!strcmp(name, "northstar_go1_h20")) base(p, 20, 0.025, LMPC_DISC_ZOH, LMPC_COST_NS_TWO, 1, kGo1Q);
  else return LMPC_ERR_INVALID_ARG;
  return LMPC_OK;
}

extern "C" int32_t lmpc_gait_predict_contact(int32_t gait, int32_t leg, double phase,
                                             double gait_speed, double delta_t) {
  const GaitLeg t = gait_table(gait, leg);
  double ph = phase + gait_speed * delta_t;  // LeggedContactFSM.cpp:281
  while (ph > 1.0) ph -= 1.0;                // :283-285
  for (int i = 0; i < t.n; ++i)              // :287-291
    if (ph <= t.sw[i]) return t.state[i];
  return 1;                                  // :293 "should not reach here" -> STANCE
}

extern "C" lmpc_status lmpc_gait_contact_table(int32_t batch, int32_t horizon, double dt,
                                               const int32_t* gait, const double* phase,
                                               const double* gait_speed, uint8_t* contact) {
  if (batch < 0 || horizon < 1 || horizon > LMPC_MAX_HORIZON || !gait || !phase || !gait_speed || !contact)
    return LMPC_ERR_INVALID_ARG;
  for (int b = 0; b < batch; ++b)
    for (int i = 0; i < horizon; ++i)  // ConvexQPSolver.cpp:329-346: row i uses predict(i * mpc_dt_)
      for (int leg = 0; leg < LMPC_NUM_LEG; ++leg)
        contact[((size_t)b * horizon + i) * 4 + leg] =
            (uint8_t)lmpc_gait_predict_contact(gait[b], leg, phase[4 * (size_t)b + leg], gait_speed[b], (double)i * dt);
  return LMPC_OK;
}
