// Field-name mock of the reference types lmpc_legged_adapter.hpp touches (test infrastructure).
// NOT the reference's code: only the member names and shapes the adapter reads or writes, with
// the two edits INTEGRATION.md asks for already applied (virtual grf_update, protected members).
// A fixed-size column-major matrix stands in for Eigen::Matrix<double,R,C>.
#pragma once
#include <cstddef>

#define NUM_LEG 4

namespace mock {
template <int R, int C>
struct Mat {
  double v[R * C] = {};
  double* data() { return v; }
  const double* data() const { return v; }
  double& operator[](int i) { return v[i]; }
  double operator[](int i) const { return v[i]; }
  double& operator()(int r, int c) { return v[r + R * c]; }
  double operator()(int r, int c) const { return v[r + R * c]; }
};
}  // namespace mock

namespace legged {

enum LeggedContactState { SWING, STANCE };

struct LeggedState {
  struct Fbk {
    mock::Mat<3, 1> root_euler, root_pos, root_ang_vel, root_lin_vel;
    mock::Mat<3, 3> root_rot_mat;
    mock::Mat<3, 4> foot_pos_abs;
  } fbk;
  struct Ctrl {
    mock::Mat<3, 1> root_euler_d, root_pos_d, root_lin_vel_d_rel, root_ang_vel_d_rel;
    bool plan_contacts[NUM_LEG] = {true, true, true, true};
    mock::Mat<24, 1> optimized_input;
  } ctrl;
  struct Param {
    mock::Mat<12, 1> q_weights, r_weights;
    double robot_mass = 13.0;
    mock::Mat<3, 3> a1_trunk_inertia;
  } param;
};

struct LeggedContactFSM {
  double gait_phase = 0.0, gait_speed = 3.5;
  bool diag_a = true;
  LeggedContactState predict_contact_state(double dt) {
    double p = gait_phase + gait_speed * dt;
    while (p > 1.0) p -= 1.0;
    return ((p <= 0.5) == diag_a) ? STANCE : SWING;
  }
};

class LeggedMPC {
 public:
  virtual ~LeggedMPC() = default;
  virtual bool update(LeggedState& state, double t, double dt) = 0;
};

class ConvexMpc : public LeggedMPC {
 public:
  explicit ConvexMpc(LeggedState&) {
    for (int j = 0; j < NUM_LEG; ++j) leg_FSM[j].diag_a = (j == 0 || j == 3);
  }
  bool update(LeggedState& state, double t, double dt) override {
    grf_update(state, t, dt);
    for (int i = 0; i < NUM_LEG; ++i)
      for (int a = 0; a < 3; ++a) state.ctrl.optimized_input[3 * i + a] = foot_forces_grf_world(a, i);
    return true;
  }
  virtual bool grf_update(LeggedState& state, double t, double dt) = 0;

 protected:
  mock::Mat<3, NUM_LEG> foot_forces_grf_world;
  LeggedContactFSM leg_FSM[NUM_LEG];
};

}  // namespace legged
