// lmpc_legged_adapter.hpp — the class a legged_ctrl maintainer adds to run the convex MPC tick on
// liblmpc.so: ConvexMpcB200 keeps LeggedMPC::update(LeggedState&, double t, double dt)
// (/root/reference/src/legged_ctrl/include/mpc_ctrl/LeggedMPC.h:13) and everything
// ConvexMpc::update does around the solver (src/mpc_ctrl/convex_mpc/ConvexMpc.cpp:24-62: joystick
// -> references, foot_update, output packing); only grf_update (ConvexMpc.cpp:64-78) changes.
//
// Needs the reference's headers (Eigen + ROS) on the include path and two one-word edits to
// include/mpc_ctrl/convex_mpc/ConvexMpc.h: `grf_update` becomes virtual (:34) and `private:`
// becomes `protected:` (:39, for leg_FSM and foot_forces_grf_world).  This repository compiles it
// only against the field-name mock in tests/cpp/mock_reference/ (Eigen and ROS are absent here).
#ifndef LMPC_LEGGED_ADAPTER_HPP_
#define LMPC_LEGGED_ADAPTER_HPP_

#include <cstdint>
#include <memory>

#include "lmpc.hpp"
#include "mpc_ctrl/convex_mpc/ConvexMpc.h"

namespace legged {

class ConvexMpcB200 : public ConvexMpc {
 public:
  explicit ConvexMpcB200(LeggedState& state, int device = 0) : ConvexMpc(state) {
    lmpc_params p = lmpc::preset("reference_a1");  // H=30, dt=0.01, Euler, mu=0.3, fz<=180 (ConvexQPSolver.cpp:25-26,171)
    for (int i = 0; i < 12; ++i) {                  // ConvexQPSolver(q_weights, r_weights), ConvexMpc.cpp:15
      p.q[i] = state.param.q_weights[i];
      p.r[i] = state.param.r_weights[i];
    }
    p.mass = state.param.robot_mass;                                                 // LeggedState.h:164
    for (int i = 0; i < 9; ++i) p.inertia[i] = state.param.a1_trunk_inertia.data()[i];  // :165, column-major
    engine_ = std::make_unique<lmpc::ConvexMpcEngine>(p, /*max_batch=*/1, device);   // throws: no CPU fallback
    engine_->set_warm_start(true, /*horizon_shift=*/1);  // OSQP warm start, ConvexQPSolver.cpp:185
  }

  bool grf_update(LeggedState& state, double /*t*/, double /*dt*/) override {  // ConvexMpc.cpp:64-78
    const int H = engine_->horizon();
    const double mpc_dt = engine_->params().dt;
    double x0[12], cmd[9];
    std::uint8_t contact[LMPC_MAX_HORIZON * LMPC_NUM_LEG];
    for (int a = 0; a < 3; ++a) {  // ConvexQPSolver.cpp:256-259
      x0[a] = state.fbk.root_euler[a];
      x0[3 + a] = state.fbk.root_pos[a];
      x0[6 + a] = state.fbk.root_ang_vel[a];
      x0[9 + a] = state.fbk.root_lin_vel[a];
      cmd[3 + a] = state.ctrl.root_lin_vel_d_rel[a];  // :260
      cmd[6 + a] = state.ctrl.root_ang_vel_d_rel[a];  // :264-276
    }
    cmd[0] = state.ctrl.root_euler_d[0];
    cmd[1] = state.ctrl.root_euler_d[1];
    cmd[2] = state.ctrl.root_pos_d[2];
    for (int j = 0; j < NUM_LEG; ++j) contact[j] = state.ctrl.plan_contacts[j] ? 1 : 0;  // :331-337
    for (int i = 1; i < H; ++i)
      for (int j = 0; j < NUM_LEG; ++j)
        contact[4 * i + j] = leg_FSM[j].predict_contact_state(i * mpc_dt) == STANCE ? 1 : 0;  // :338-344
    const auto grf = engine_->compute_grfs(x0, state.fbk.root_rot_mat, state.fbk.foot_pos_abs, cmd, contact);
    for (int i = 0; i < NUM_LEG; ++i)
      for (int a = 0; a < 3; ++a) foot_forces_grf_world(a, i) = grf[3 * i + a];  // ConvexMpc.cpp:73-75
    return true;  // the reference never reports failure (:77); a failed solve already returned zeros
  }

 private:
  std::unique_ptr<lmpc::ConvexMpcEngine> engine_;
};

}  // namespace legged
#endif  // LMPC_LEGGED_ADAPTER_HPP_
