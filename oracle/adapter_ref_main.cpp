/*
 * adapter_ref_main.cpp — drives include/lmpc_legged_adapter.hpp (ConvexMpcB200 : LeggedMPC) compiled
 * against the reference's REAL headers (LeggedState.h, utils/LeggedContactFSM.h, MovingWindowFilter.hpp,
 * mpc_ctrl/LeggedMPC.h, over the Eigen/ros shims of oracle/shim) and linked with the reference's
 * unmodified LeggedContactFSM.cpp / LeggedState.cpp / LeggedMPC.cpp / Utils.cpp and liblmpc.so —
 * and with NEITHER ConvexQPSolver.cpp NOR ConvexMpc.cpp NOR anything of OsqpEigen: that the link
 * succeeds is the proof that the adapter removes the OSQP dependency of the convex-MPC tick.
 * TEST INFRASTRUCTURE (built by oracle/Makefile into oracle/_ref/adapter_ref, run by
 * tests/test_cpp_wrapper.py on the GPU box against tests/golden/golden_adapter_tick.npz, which the
 * reference's own ConvexMpc::update produced here for the same scripted ticks).
 *
 *   adapter_ref <scenario.bin> <out.bin>
 * scenario: int32 T, then doubles q[12] r[12] mass inertia[9] gait_counter_speed, then per tick
 *   t dt movement_mode estimation_inited joy[4] roll_pitch_d[2] x0[12] rot[9] foot_abs[12]
 *   foot_world[12] foot_target_world[12] contact_flag[4]
 * out per tick (doubles): optimized_state[18] optimized_input[24] plan_contacts[4] v_d_world[3]
 *   euler_d[3] lin_vel_d_rel[2] mpc_solver_inited
 */
#include <cstdio>
#include <cstdint>
#include <fstream>
#include <sstream>
#include <vector>

#include "shim/mini_eigen.h"
#include "shim/ros/ros.h"
#include "lmpc_legged_adapter.hpp"

int main(int argc, char** argv) {
  if (argc < 3) { std::printf("usage: adapter_ref <scenario.bin> <out.bin>\n"); return 2; }
  std::ifstream f(argv[1], std::ios::binary);
  std::int32_t T = 0;
  f.read(reinterpret_cast<char*>(&T), 4);
  double hdr[12 + 12 + 1 + 9 + 1];
  f.read(reinterpret_cast<char*>(hdr), sizeof(hdr));
  if (!f || T < 1) { std::printf("bad scenario\n"); return 2; }
  legged::LeggedState st;
  Eigen::VectorXd qw(12), rw(12);
  for (int i = 0; i < 12; ++i) { qw[i] = hdr[i]; rw[i] = hdr[12 + i]; }
  st.param.q_weights = qw; st.param.r_weights = rw; st.param.robot_mass = hdr[24];
  for (int k = 0; k < 9; ++k) st.param.a1_trunk_inertia.data()[k] = hdr[25 + k];
  st.param.gait_counter_speed = hdr[34];
  st.ctrl.root_pos_d.setZero(); st.ctrl.root_euler_d.setZero();
  st.ctrl.root_lin_vel_d_rel.setZero(); st.ctrl.root_lin_vel_d_world.setZero();
  st.ctrl.root_ang_vel_d_rel.setZero(); st.ctrl.root_ang_vel_d_world.setZero();
  st.ctrl.gait_counter.setZero(); st.ctrl.foot_pos_target_world.setZero();
  st.ctrl.optimized_state.setZero(); st.ctrl.optimized_input.setZero();
  st.fbk.foot_pos_world.setZero(); st.fbk.foot_contact_flag.setZero();
  for (int j = 0; j < NUM_LEG; ++j) st.ctrl.plan_contacts[j] = true;
  try {
    legged::ConvexMpcB200 mpc(st);
    for (int j = 0; j < NUM_LEG; ++j) mpc.leg_fsm(j).reset();  // as oracle/ref_harness.cpp does for the reference's ConvexMpc
    std::ofstream o(argv[2], std::ios::binary);
    std::ostringstream sink;  // the reference's FSM prints from common_enter()
    std::streambuf* old = std::cout.rdbuf(sink.rdbuf());
    for (int k = 0; k < T; ++k) {
      double in[4 + 4 + 2 + 12 + 9 + 12 + 12 + 12 + 4];
      f.read(reinterpret_cast<char*>(in), sizeof(in));
      if (!f) { std::cout.rdbuf(old); std::printf("short scenario at tick %d\n", k); return 2; }
      const double *joy = in + 4, *rp = in + 8, *x0 = in + 10, *rot = in + 22, *fa = in + 31, *fw = in + 43, *ft = in + 55, *cf = in + 67;
      st.estimation_inited = in[3] != 0.0;
      st.ctrl.movement_mode = in[2];
      st.joy.velx = joy[0]; st.joy.vely = joy[1]; st.joy.yaw_rate = joy[2]; st.joy.body_height = joy[3];
      st.ctrl.root_euler_d[0] = rp[0]; st.ctrl.root_euler_d[1] = rp[1];
      for (int i = 0; i < 3; ++i) { st.fbk.root_euler[i] = x0[i]; st.fbk.root_pos[i] = x0[3 + i]; st.fbk.root_ang_vel[i] = x0[6 + i]; st.fbk.root_lin_vel[i] = x0[9 + i]; }
      for (int i = 0; i < 9; ++i) st.fbk.root_rot_mat.data()[i] = rot[i];
      for (int i = 0; i < 12; ++i) { st.fbk.foot_pos_abs.data()[i] = fa[i]; st.fbk.foot_pos_world.data()[i] = fw[i]; st.ctrl.foot_pos_target_world.data()[i] = ft[i]; }
      for (int j = 0; j < 4; ++j) st.fbk.foot_contact_flag[j] = cf[j];
      mpc.update(st, in[0], in[1]);
      double out[18 + 24 + 4 + 3 + 3 + 2 + 1];
      for (int i = 0; i < 18; ++i) out[i] = st.ctrl.optimized_state[i];
      for (int i = 0; i < 24; ++i) out[18 + i] = st.ctrl.optimized_input[i];
      for (int j = 0; j < 4; ++j) out[42 + j] = st.ctrl.plan_contacts[j] ? 1.0 : 0.0;
      for (int i = 0; i < 3; ++i) { out[46 + i] = st.ctrl.root_lin_vel_d_world[i]; out[49 + i] = st.ctrl.root_euler_d[i]; }
      out[52] = st.ctrl.root_lin_vel_d_rel[0]; out[53] = st.ctrl.root_lin_vel_d_rel[1];
      out[54] = st.mpc_solver_inited ? 1.0 : 0.0;
      o.write(reinterpret_cast<const char*>(out), sizeof(out));
    }
    std::cout.rdbuf(old);
    std::printf("adapter_ref: %d ticks, launches %lld\n", T, (long long)mpc.engine().launch_count());
    return 0;
  } catch (const lmpc::Error& e) {
    std::printf("lmpc::Error %d: %s\n", (int)e.code(), e.what());
    return e.code() == LMPC_ERR_NO_DEVICE ? 3 : 2;
  }
}
