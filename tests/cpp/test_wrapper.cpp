// Drives include/lmpc.hpp and include/lmpc_legged_adapter.hpp (against the field-name mock of the
// reference types) through liblmpc.so.  Prints one line per check; exit code 0 = all good,
// 3 = no usable GPU (the engine refused to construct: there is no CPU fallback).
//   argv[1] = "adapter": standing A1 through ConvexMpcB200::update -> prints the 12 GRFs
#include <cmath>
#include <cstdio>
#include <cstring>

#include "lmpc_legged_adapter.hpp"

int main(int argc, char** argv) {
  legged::LeggedState s;
  const double q[12] = {60.0, 100.0, 0.0, 0.0, 0.0, 450.0, 0.15, 0.15, 100.0, 3.0, 3.0, 5.0};
  for (int i = 0; i < 12; ++i) { s.param.q_weights[i] = q[i]; s.param.r_weights[i] = 1e-4; }
  s.param.a1_trunk_inertia(0, 0) = 0.0158533; s.param.a1_trunk_inertia(1, 1) = 0.0377999; s.param.a1_trunk_inertia(2, 2) = 0.0456542;
  s.fbk.root_pos[2] = 0.3; s.ctrl.root_pos_d[2] = 0.3;
  for (int i = 0; i < 3; ++i) s.fbk.root_rot_mat(i, i) = 1.0;
  const double fh[4][3] = {{0.17, 0.17, -0.3}, {0.17, -0.17, -0.3}, {-0.17, 0.17, -0.3}, {-0.17, -0.17, -0.3}};
  for (int l = 0; l < 4; ++l)
    for (int a = 0; a < 3; ++a) s.fbk.foot_pos_abs(a, l) = fh[l][a];
  s.ctrl.plan_contacts[1] = s.ctrl.plan_contacts[2] = false;  // trot phase 0: legs 0 and 3 in stance
  try {
    legged::ConvexMpcB200 mpc(s);
    for (int tick = 0; tick < 3; ++tick) mpc.update(s, 0.01 * tick, 0.01);
    std::printf("grf");
    for (int i = 0; i < 12; ++i) std::printf(" %.17g", s.ctrl.optimized_input[i]);
    std::printf("\n");
    // direct use of the wrapper with a RobotTick and the gait helper
    lmpc::ConvexMpcEngine eng(q, s.param.r_weights, "northstar_a1", 4);
    lmpc::RobotTick t;
    std::memcpy(t.x0.data() + 3, s.fbk.root_pos.data(), 24);
    std::memcpy(t.rot.data(), s.fbk.root_rot_mat.data(), 72);
    std::memcpy(t.foot_pos_abs.data(), s.fbk.foot_pos_abs.data(), 96);
    t.cmd[2] = 0.3;
    const double ph[4] = {0.1, 0.1, 0.1, 0.1};
    const auto c = eng.contact_table(LMPC_GAIT_STAND, ph, 3.5);
    std::memcpy(t.contact.data(), c.data(), c.size());
    int st = -1;
    const auto g = eng.compute_grfs(t, &st);
    std::printf("stand status %d fz %.17g %.17g %.17g %.17g\n", st, g[2], g[5], g[8], g[11]);
    return (st == 0 && std::fabs(g[2] - g[11]) < 1e-6) ? 0 : 1;
  } catch (const lmpc::Error& e) {
    std::printf("lmpc::Error %d: %s\n", (int)e.code(), e.what());
    return e.code() == LMPC_ERR_NO_DEVICE ? 3 : 2;
  }
}
