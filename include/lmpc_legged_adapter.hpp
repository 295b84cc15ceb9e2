// lmpc_legged_adapter.hpp — the class a legged_ctrl maintainer adds to run the convex-MPC tick on
// liblmpc.so.  ConvexMpcB200 implements LeggedMPC::update(LeggedState&, double t, double dt)
// (/root/reference/src/legged_ctrl/include/mpc_ctrl/LeggedMPC.h:13) and is a sibling of the
// reference's ConvexMpc, NOT a subclass: ConvexMpc holds a `ConvexQPSolver fastConvex` member
// (include/mpc_ctrl/convex_mpc/ConvexMpc.h:46) that sets up an OSQP problem in its constructor
// (ConvexQPSolver.cpp:181-195), so anything derived from it still links OsqpEigen and pays that
// setup.  This class owns the four LeggedContactFSMs and the two command filters itself and runs the
// same tick around a different solver:
//
//   ConvexMpc::ConvexMpc      src/mpc_ctrl/convex_mpc/ConvexMpc.cpp:6-22     constructor below
//   ConvexMpc::update         ConvexMpc.cpp:24-62                            update()
//   ConvexMpc::grf_update     ConvexMpc.cpp:64-78                            grf_update()  -> lmpc_solve
//   ConvexMpc::foot_update    ConvexMpc.cpp:80-108                           foot_update()
//
// It needs only the reference's own headers (LeggedState.h, utils/LeggedContactFSM.h,
// utils/MovingWindowFilter.hpp, mpc_ctrl/LeggedMPC.h) — no edit to any reference file, no
// ConvexQPSolver.h, no OsqpEigen.  In this repository it is compiled against those real headers
// over the Eigen/ros shims of oracle/shim (oracle/Makefile, target _ref/adapter_ref) and linked with
// the reference's unmodified LeggedContactFSM.cpp / LeggedState.cpp / Utils.cpp and liblmpc.so.
#ifndef LMPC_LEGGED_ADAPTER_HPP_
#define LMPC_LEGGED_ADAPTER_HPP_

#include <cstdint>
#include <iostream>
#include <memory>

#include "LeggedParams.h"
#include "LeggedState.h"
#include "lmpc.hpp"
#include "mpc_ctrl/LeggedMPC.h"
#include "utils/LeggedContactFSM.h"
#include "utils/MovingWindowFilter.hpp"

namespace legged {

class ConvexMpcB200 : public LeggedMPC {
 public:
  explicit ConvexMpcB200(LeggedState& state, int device = 0)
      : total_run_time_(0.0),
        vel_filter_x_(int(1000.0 / MPC_UPDATE_FREQUENCY * 0.3)),  // ConvexMpc.cpp:19-20
        vel_filter_y_(int(1000.0 / MPC_UPDATE_FREQUENCY * 0.3)) {
    for (int i = 0; i < NUM_LEG; ++i) leg_FSM_[i].reset_params(state, i);  // ConvexMpc.cpp:8-11
    foot_forces_grf_world_.setZero();
    lmpc_params p = lmpc::preset("reference_a1");  // H=30, dt=0.01, Euler, mu=0.3, fz<=180 (ConvexQPSolver.cpp:25-26,171)
    for (int i = 0; i < 12; ++i) {                  // ConvexQPSolver(q_weights, r_weights), ConvexMpc.cpp:15-16
      p.q[i] = state.param.q_weights[i];
      p.r[i] = state.param.r_weights[i];
    }
    p.mass = state.param.robot_mass;                                                     // ConvexQPSolver.cpp:204
    for (int i = 0; i < 9; ++i) p.inertia[i] = state.param.a1_trunk_inertia.data()[i];  // :201, column-major
    engine_ = std::make_unique<lmpc::ConvexMpcEngine>(p, /*max_batch=*/1, device);      // throws: no CPU fallback
    engine_->set_warm_start(true, /*horizon_shift=*/1);  // OSQP warm start, ConvexQPSolver.cpp:185
  }

  bool update(LeggedState& state, double t, double dt) override {  // ConvexMpc.cpp:24-62
    total_run_time_ += dt;
    if (!state.estimation_inited) {
      std::cout << "estimation not inited" << std::endl;
      return true;
    }
    state.ctrl.root_pos_d[2] = state.joy.body_height;
    state.ctrl.root_lin_vel_d_rel[0] = vel_filter_x_.CalculateAverage(state.joy.velx);
    state.ctrl.root_lin_vel_d_rel[1] = vel_filter_y_.CalculateAverage(state.joy.vely);
    state.ctrl.root_ang_vel_d_rel[2] = state.joy.yaw_rate;
    state.ctrl.root_euler_d[2] += state.joy.yaw_rate * dt;
    foot_update(state, t, dt);
    grf_update(state, t, dt);
    for (int a = 0; a < 3; ++a) {  // ConvexMpc.cpp:48-49
      state.ctrl.optimized_state[a] = state.ctrl.root_pos_d[a];
      state.ctrl.optimized_state[3 + a] = state.ctrl.root_euler_d[a];
    }
    for (int i = 0; i < NUM_LEG; ++i)  // :51-56
      for (int a = 0; a < 3; ++a) {
        state.ctrl.optimized_input[3 * i + a] = foot_forces_grf_world_(a, i);
        state.ctrl.optimized_state[6 + 3 * i + a] = leg_FSM_[i].FSM_foot_pos_target_world[a];
        state.ctrl.optimized_input[12 + 3 * i + a] = leg_FSM_[i].FSM_foot_vel_target_world[a];
      }
    state.mpc_solver_inited = true;  // :59
    return true;
  }

  // calc_mpc_reference + update_cons_matrix + compute_grfs (ConvexMpc.cpp:70-72) on the GPU
  bool grf_update(LeggedState& state, double /*t*/, double /*dt*/) {
    const int H = engine_->horizon();
    const double mpc_dt = engine_->params().dt;
    double x0[12], cmd[9];
    std::uint8_t contact[LMPC_MAX_HORIZON * LMPC_NUM_LEG];
    for (int a = 0; a < 3; ++a) {  // ConvexQPSolver.cpp:256-259
      x0[a] = state.fbk.root_euler[a];
      x0[3 + a] = state.fbk.root_pos[a];
      x0[6 + a] = state.fbk.root_ang_vel[a];
      x0[9 + a] = state.fbk.root_lin_vel[a];
      cmd[3 + a] = state.ctrl.root_lin_vel_d_rel[a];
      cmd[6 + a] = state.ctrl.root_ang_vel_d_rel[a];  // :264-276
    }
    cmd[0] = state.ctrl.root_euler_d[0];
    cmd[1] = state.ctrl.root_euler_d[1];
    cmd[2] = state.ctrl.root_pos_d[2];
    // side effect of calc_mpc_reference the low-level controller reads: v_d,world = R v_d,rel (:260)
    state.ctrl.root_lin_vel_d_world = state.fbk.root_rot_mat * state.ctrl.root_lin_vel_d_rel;
    for (int j = 0; j < NUM_LEG; ++j) contact[j] = state.ctrl.plan_contacts[j] ? 1 : 0;  // :331-337
    for (int i = 1; i < H; ++i)
      for (int j = 0; j < NUM_LEG; ++j)
        contact[4 * i + j] = leg_FSM_[j].predict_contact_state(i * mpc_dt) == STANCE ? 1 : 0;  // :338-344
    const auto grf = engine_->compute_grfs(x0, state.fbk.root_rot_mat, state.fbk.foot_pos_abs, cmd, contact);
    for (int i = 0; i < NUM_LEG; ++i)
      for (int a = 0; a < 3; ++a) foot_forces_grf_world_(a, i) = grf[3 * i + a];  // ConvexMpc.cpp:73-75
    return true;  // the reference never reports failure (:77); a failed solve already returned zeros
  }

  bool foot_update(LeggedState& state, double /*t*/, double dt) {  // ConvexMpc.cpp:80-108
    if (state.ctrl.movement_mode == 0) {
      for (int i = 0; i < NUM_LEG; ++i) {
        leg_FSM_[i].reset();
        state.ctrl.plan_contacts[i] = true;
      }
    } else {
      for (int i = 0; i < NUM_LEG; ++i) {
        const Eigen::Vector3d cur = state.fbk.foot_pos_world.col(i), tgt = state.ctrl.foot_pos_target_world.col(i);
        state.ctrl.gait_counter[i] = leg_FSM_[i].update(dt, cur, tgt, state.fbk.foot_contact_flag[i]);
      }
      for (int i = 0; i < NUM_LEG; ++i) state.ctrl.plan_contacts[i] = leg_FSM_[i].get_contact_state();
    }
    return true;
  }

  LeggedContactFSM& leg_fsm(int i) { return leg_FSM_[i]; }
  lmpc::ConvexMpcEngine& engine() { return *engine_; }

 private:
  double total_run_time_;
  Eigen::Matrix<double, 3, NUM_LEG> foot_forces_grf_world_;
  LeggedContactFSM leg_FSM_[NUM_LEG];
  MovingWindowFilter vel_filter_x_, vel_filter_y_;
  std::unique_ptr<lmpc::ConvexMpcEngine> engine_;
};

}  // namespace legged
#endif  // LMPC_LEGGED_ADAPTER_HPP_
