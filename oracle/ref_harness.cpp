/*
 * ref_harness.cpp — C entry points around the UNMODIFIED reference sources, for oracle/_ref.
 * TEST INFRASTRUCTURE ONLY (tests/ and bench.py's cpu_baseline leg load the resulting library).
 *
 * oracle/Makefile (target `ref`) compiles, from where they lie under /root/reference,
 *   src/legged_ctrl/src/mpc_ctrl/convex_mpc/ConvexQPSolver.cpp   (the hot path)
 *   src/legged_ctrl/src/mpc_ctrl/convex_mpc/ConvexMpc.cpp        (the controller tick)
 *   src/legged_ctrl/src/mpc_ctrl/LeggedMPC.cpp, src/legged_ctrl/src/LeggedState.cpp
 *   src/legged_ctrl/src/utils/{LeggedContactFSM,Utils,A1Kinematics}.cpp
 * against oracle/shim (an Eigen-subset, a capturing OsqpEigen stub, a ros::NodeHandle stub) and
 * links them with this file into oracle/_ref/libref_convex_mpc.so.  This file contains no
 * restatement of the reference's arithmetic: it fills a LeggedState, calls the reference's own
 * functions and copies out what they produced — in particular the exact (P, q, A, l, u) that
 * ConvexQPSolver::compute_grfs hands to OSQP (ConvexQPSolver.cpp:315-318).
 *
 * The reference keeps the FSM phase and the solver private; this translation unit (only) sees
 * them through `#define private public` so that a test can place a leg at a given gait phase.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <iostream>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "shim/mini_eigen.h"
#include "shim/mini_eigen_sparse.h"
#include "shim/OsqpEigen/OsqpEigen.h"
#include "shim/ros/ros.h"

#define private public
#include "mpc_ctrl/convex_mpc/ConvexMpc.h"
#include "utils/A1Kinematics.h"
#undef private

using legged::ConvexMpc;
using legged::ConvexQPSolver;
using legged::LeggedContactFSM;
using legged::LeggedState;

namespace {

const int H = PLAN_HORIZON;
const int NZ = (MPC_STATE_DIM_SPARSE + DIM_GRF) * PLAN_HORIZON;
const int NROW = (MPC_STATE_DIM_SPARSE + 4 * NUM_LEG + 4) * PLAN_HORIZON;

/* solution source for OsqpEigen::Solver::solve() */
typedef int (*ref_solve_cb)(int n, int m, const double* Pdiag, const double* q, const double* A,
                            const double* l, const double* u, double* z);
struct HookState {
  const double* fixed_z = nullptr; /* if set: copy (size n) */
  ref_solve_cb cb = nullptr;       /* else if set: ask the callback */
} g_hook;

void dense_from_captured(const OsqpEigen::CapturedQP& qp, double* Pdiag, double* q, double* A, double* l, double* u) {
  const int n = qp.n, m = qp.m;
  if (Pdiag) {
    for (int j = 0; j < n; ++j) Pdiag[j] = 0.0;
    for (int j = 0; j < n; ++j)
      for (Eigen::SparseMatrix<double>::InnerIterator it(qp.P, j); it; ++it)
        if (it.row() == j) Pdiag[j] = it.value(); else Pdiag[j] = std::numeric_limits<double>::quiet_NaN(); /* reference P is diagonal */
  }
  if (q) for (int j = 0; j < n; ++j) q[j] = qp.q(j);
  if (A) {
    std::memset(A, 0, sizeof(double) * (size_t)n * (size_t)m);
    for (int j = 0; j < n; ++j)
      for (Eigen::SparseMatrix<double>::InnerIterator it(qp.A, j); it; ++it) A[(size_t)it.row() + (size_t)j * (size_t)m] = it.value();
  }
  if (l) for (int i = 0; i < m; ++i) l[i] = qp.l(i);
  if (u) for (int i = 0; i < m; ++i) u[i] = qp.u(i);
}

bool hook(const OsqpEigen::CapturedQP& qp, Eigen::VectorXd& z, void*) {
  if (g_hook.fixed_z) { for (int j = 0; j < qp.n; ++j) z(j) = g_hook.fixed_z[j]; return true; }
  if (g_hook.cb) {
    std::vector<double> P((size_t)qp.n), q((size_t)qp.n), A((size_t)qp.n * (size_t)qp.m), l((size_t)qp.m), u((size_t)qp.m), zz((size_t)qp.n, 0.0);
    dense_from_captured(qp, P.data(), q.data(), A.data(), l.data(), u.data());
    if (g_hook.cb(qp.n, qp.m, P.data(), q.data(), A.data(), l.data(), u.data(), zz.data()) != 0) return false;
    for (int j = 0; j < qp.n; ++j) z(j) = zz[(size_t)j];
    return true;
  }
  return false;
}

void set_gait(LeggedContactFSM& f, int leg, int gait, double speed) {
  f.leg_id = leg;
  f.gait_speed = speed;
  f.s = legged::STANCE;
  switch (gait) {
    case 0: f.set_default_gait_pattern(); break;
    case 1: f.set_trot_with_stand_gait_pattern(); break;
    case 2: f.set_crawl_gait_pattern(); break;
    default: f.set_default_stand_pattern(); break;
  }
}

void fill_state(LeggedState& st, const double x0[12], const double rot[9], const double foot_abs[12], const double cmd[9]) {
  for (int i = 0; i < 3; ++i) {
    st.fbk.root_euler[i] = x0[i];
    st.fbk.root_pos[i] = x0[3 + i];
    st.fbk.root_ang_vel[i] = x0[6 + i];
    st.fbk.root_lin_vel[i] = x0[9 + i];
  }
  for (int k = 0; k < 9; ++k) st.fbk.root_rot_mat.data()[k] = rot[k];
  for (int k = 0; k < 12; ++k) st.fbk.foot_pos_abs.data()[k] = foot_abs[k];
  st.ctrl.root_euler_d.setZero(); st.ctrl.root_pos_d.setZero();
  st.ctrl.root_euler_d[0] = cmd[0]; st.ctrl.root_euler_d[1] = cmd[1]; st.ctrl.root_pos_d[2] = cmd[2];
  for (int i = 0; i < 3; ++i) { st.ctrl.root_lin_vel_d_rel[i] = cmd[3 + i]; st.ctrl.root_ang_vel_d_rel[i] = cmd[6 + i]; }
  st.ctrl.root_lin_vel_d_world.setZero(); st.ctrl.root_ang_vel_d_world.setZero();
}

struct StreamSilencer { /* the reference prints from common_enter() and update() */
  std::ostringstream sink; /* declared (hence constructed) before `old` */
  std::streambuf* old;
  StreamSilencer() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~StreamSilencer() { std::cout.rdbuf(old); }
};

struct Tick {
  LeggedState state;
  std::unique_ptr<ConvexMpc> mpc;
};

}  // namespace

extern "C" {

int ref_horizon(void) { return H; }
double ref_mpc_dt(void) { return MPC_UPDATE_FREQUENCY / 1000.0; }
int ref_nz(void) { return NZ; }
int ref_nrow(void) { return NROW; }
double ref_osqp_infty(void) { return OsqpEigen::INFTY; }

/* ---- LeggedParam::load on a reference yaml (LeggedState.cpp:20-209) ---- */
int ref_load_params(const char* yaml_path, double q[12], double r[12], double* mass, double inertia[9],
                    double* gait_counter_speed, double default_foot_pos_rel[12], double kp_kd_km[27]) {
  if (ros::shim::load_yaml(yaml_path) < 0) return -1;
  legged::LeggedParam p; ros::NodeHandle nh;
  if (!p.load(nh)) return -2;
  for (int i = 0; i < 12; ++i) { q[i] = p.q_weights[i]; r[i] = p.r_weights[i]; default_foot_pos_rel[i] = p.default_foot_pos_rel.data()[i]; }
  *mass = p.robot_mass; *gait_counter_speed = p.gait_counter_speed;
  for (int k = 0; k < 9; ++k) inertia[k] = p.a1_trunk_inertia.data()[k];
  if (kp_kd_km) {
    for (int k = 0; k < 12; ++k) { kp_kd_km[k] = p.kp_foot.data()[k]; kp_kd_km[12 + k] = p.kd_foot.data()[k]; }
    for (int k = 0; k < 3; ++k) kp_kd_km[24 + k] = p.km_foot[k];
  }
  return 0;
}

/* ---- the QP the reference hands to OSQP, for one robot ----
 * Runs ConvexQPSolver(q,r) (ConvexQPSolver.cpp:16-196), calc_mpc_reference (:254-313),
 * update_cons_matrix (:230-239) and compute_grfs (:314-327) exactly as ConvexMpc::grf_update does
 * (ConvexMpc.cpp:70-72), with leg FSMs placed at (gait, phase[leg], gait_speed).
 * Outputs: Pdiag[24H], qv[24H], A[32H x 24H col-major], l[32H], u[32H], v_d_world[3]
 * (ctrl.root_lin_vel_d_world, :260), grf[12] (what compute_grfs returned given z_solution, or the
 * NaN->0 rule when z_solution == NULL). */
int ref_qp_build(const double q_w[12], const double r_w[12], double mass, const double inertia[9],
                 const double x0[12], const double rot[9], const double foot_abs[12], const double cmd[9],
                 const uint8_t plan_contacts[4], int gait, const double phase[4], double gait_speed,
                 const double* z_solution, double* Pdiag, double* qv, double* A, double* l, double* u,
                 double* v_d_world, double* grf) {
  Eigen::VectorXd qw(12), rw(12);
  for (int i = 0; i < 12; ++i) { qw[i] = q_w[i]; rw[i] = r_w[i]; }
  LeggedState st;
  st.param.q_weights = qw; st.param.r_weights = rw; st.param.robot_mass = mass;
  for (int k = 0; k < 9; ++k) st.param.a1_trunk_inertia.data()[k] = inertia[k];
  st.param.gait_counter_speed = gait_speed;
  fill_state(st, x0, rot, foot_abs, cmd);
  for (int j = 0; j < NUM_LEG; ++j) st.ctrl.plan_contacts[j] = plan_contacts[j] != 0;
  LeggedContactFSM fsm[NUM_LEG];
  for (int j = 0; j < NUM_LEG; ++j) { set_gait(fsm[j], j, gait, gait_speed); fsm[j].gait_phase = phase[j]; }
  ConvexQPSolver solver(qw, rw);
  OsqpEigen::solve_hook() = hook; g_hook.fixed_z = z_solution; g_hook.cb = nullptr;
  solver.calc_mpc_reference(st, fsm);
  solver.update_cons_matrix();
  Eigen::Matrix<double, DIM_GRF, 1> f = solver.compute_grfs(st);
  g_hook.fixed_z = nullptr;
  const OsqpEigen::CapturedQP& qp = solver.solver.captured();
  if (qp.n != NZ || qp.m != NROW) return -1;
  dense_from_captured(qp, Pdiag, qv, A, l, u);
  if (v_d_world) for (int i = 0; i < 3; ++i) v_d_world[i] = st.ctrl.root_lin_vel_d_world[i];
  if (grf) for (int i = 0; i < 12; ++i) grf[i] = f[i];
  return 0;
}

/* OSQP settings the reference selects (ConvexQPSolver.cpp:182-185) */
int ref_osqp_settings(double out[4]) {
  Eigen::VectorXd qw(12), rw(12); qw.setConstant(1.0); rw.setConstant(1.0);
  ConvexQPSolver solver(qw, rw);
  const OsqpEigen::CapturedQP& qp = solver.solver.captured();
  out[0] = qp.eps_abs; out[1] = qp.eps_rel; out[2] = qp.warm_start; out[3] = qp.verbose;
  return 0;
}

/* ---- LeggedContactFSM, one leg (LeggedContactFSM.cpp) ---- */
void* ref_fsm_new(int leg, int gait, double gait_speed) {
  LeggedContactFSM* f = new LeggedContactFSM();
  set_gait(*f, leg, gait, gait_speed);
  f->reset(); /* phase 0, pattern index 0, s = pattern[0] (LeggedContactFSM.cpp:16-36) */
  return f;
}
void ref_fsm_free(void* h) { delete (LeggedContactFSM*)h; }
void ref_fsm_reset(void* h) { ((LeggedContactFSM*)h)->reset(); }
void ref_fsm_set_phase(void* h, double phase) { ((LeggedContactFSM*)h)->gait_phase = phase; }
int ref_fsm_predict(void* h, double delta_t) { return (int)((LeggedContactFSM*)h)->predict_contact_state(delta_t); }
/* one update() call; info = {phase, state, pattern_index, cur_state_start_time, cur_state_end_time, percent_in_state} */
void ref_fsm_update(void* h, double dt, const double foot_cur[3], const double foot_target[3], int force_flag,
                    double info[6], double pos_target[3], double vel_target[3]) {
  LeggedContactFSM* f = (LeggedContactFSM*)h;
  StreamSilencer quiet;
  Eigen::Vector3d c(foot_cur[0], foot_cur[1], foot_cur[2]), t(foot_target[0], foot_target[1], foot_target[2]);
  double ph = f->update(dt, c, t, force_flag != 0);
  info[0] = ph; info[1] = (double)f->get_contact_state(); info[2] = f->gait_pattern_index;
  info[3] = f->cur_state_start_time; info[4] = f->cur_state_end_time; info[5] = f->percent_in_state();
  for (int i = 0; i < 3; ++i) { if (pos_target) pos_target[i] = f->FSM_foot_pos_target_world[i]; if (vel_target) vel_target[i] = f->FSM_foot_vel_target_world[i]; }
}

/* ---- the whole controller tick: ConvexMpc(state) + ConvexMpc::update (ConvexMpc.cpp:6-62) ---- */
void* ref_tick_new(const char* yaml_path) {
  if (ros::shim::load_yaml(yaml_path) < 0) return nullptr;
  Tick* t = new Tick();
  ros::NodeHandle nh;
  if (!t->state.param.load(nh)) { delete t; return nullptr; }
  /* LeggedCtrl::reset()/LeggedFeedback::reset() are empty in the reference (LeggedState.cpp:14-18):
   * give every field the tick reads a defined value */
  t->state.ctrl.root_pos_d.setZero(); t->state.ctrl.root_euler_d.setZero();
  t->state.ctrl.root_lin_vel_d_rel.setZero(); t->state.ctrl.root_lin_vel_d_world.setZero();
  t->state.ctrl.root_ang_vel_d_rel.setZero(); t->state.ctrl.root_ang_vel_d_world.setZero();
  t->state.ctrl.gait_counter.setZero(); t->state.ctrl.foot_pos_target_world.setZero();
  t->state.ctrl.optimized_state.setZero(); t->state.ctrl.optimized_input.setZero();
  t->state.fbk.foot_pos_world.setZero(); t->state.fbk.foot_contact_flag.setZero();
  for (int j = 0; j < NUM_LEG; ++j) t->state.ctrl.plan_contacts[j] = true;
  t->mpc.reset(new ConvexMpc(t->state));
  for (int j = 0; j < NUM_LEG; ++j) t->mpc->leg_FSM[j].reset();
  return t;
}
void ref_tick_free(void* h) { delete (Tick*)h; }
void ref_tick_set_gait(void* h, int gait) {
  Tick* t = (Tick*)h;
  for (int j = 0; j < NUM_LEG; ++j) { set_gait(t->mpc->leg_FSM[j], j, gait, t->state.param.gait_counter_speed); t->mpc->leg_FSM[j].reset(); }
}
/* joy = {velx, vely, yaw_rate, body_height}; roll_pitch_d = ctrl.root_euler_d[0..1] (never written by the
 * tick).  foot_world / foot_target_world / contact_flag feed foot_update (ConvexMpc.cpp:80-108).
 * Outputs: optimized_state[18], optimized_input[24], plan_contacts[4], gait_phase[4],
 * v_d_world[3], euler_d[3], lin_vel_d_rel[2], and the captured QP (any pointer may be NULL). */
int ref_tick_update(void* h, double t_now, double dt, int estimation_inited, double movement_mode,
                    const double joy[4], const double roll_pitch_d[2], const double x0[12], const double rot[9],
                    const double foot_abs[12], const double foot_world[12], const double foot_target_world[12],
                    const double contact_flag[4], ref_solve_cb cb, double* optimized_state, double* optimized_input,
                    uint8_t* plan_contacts, double* gait_phase, double* v_d_world, double* euler_d, double* lin_vel_d_rel,
                    double* Pdiag, double* qv, double* A, double* l, double* u, int* mpc_solver_inited) {
  Tick* t = (Tick*)h;
  LeggedState& st = t->state;
  StreamSilencer quiet;
  st.estimation_inited = estimation_inited != 0;
  st.ctrl.movement_mode = movement_mode;
  st.joy.velx = joy[0]; st.joy.vely = joy[1]; st.joy.yaw_rate = joy[2]; st.joy.body_height = joy[3];
  st.ctrl.root_euler_d[0] = roll_pitch_d[0]; st.ctrl.root_euler_d[1] = roll_pitch_d[1];
  for (int i = 0; i < 3; ++i) { st.fbk.root_euler[i] = x0[i]; st.fbk.root_pos[i] = x0[3 + i]; st.fbk.root_ang_vel[i] = x0[6 + i]; st.fbk.root_lin_vel[i] = x0[9 + i]; }
  for (int k = 0; k < 9; ++k) st.fbk.root_rot_mat.data()[k] = rot[k];
  for (int k = 0; k < 12; ++k) { st.fbk.foot_pos_abs.data()[k] = foot_abs[k]; st.fbk.foot_pos_world.data()[k] = foot_world[k]; st.ctrl.foot_pos_target_world.data()[k] = foot_target_world[k]; }
  for (int j = 0; j < 4; ++j) st.fbk.foot_contact_flag[j] = contact_flag[j];
  OsqpEigen::solve_hook() = hook; g_hook.fixed_z = nullptr; g_hook.cb = cb;
  bool ok = t->mpc->update(st, t_now, dt);
  g_hook.cb = nullptr;
  if (optimized_state) for (int i = 0; i < 18; ++i) optimized_state[i] = st.ctrl.optimized_state[i];
  if (optimized_input) for (int i = 0; i < 24; ++i) optimized_input[i] = st.ctrl.optimized_input[i];
  if (plan_contacts) for (int j = 0; j < 4; ++j) plan_contacts[j] = st.ctrl.plan_contacts[j] ? 1 : 0;
  if (gait_phase) for (int j = 0; j < 4; ++j) gait_phase[j] = t->mpc->leg_FSM[j].gait_phase;
  if (v_d_world) for (int i = 0; i < 3; ++i) v_d_world[i] = st.ctrl.root_lin_vel_d_world[i];
  if (euler_d) for (int i = 0; i < 3; ++i) euler_d[i] = st.ctrl.root_euler_d[i];
  if (lin_vel_d_rel) { lin_vel_d_rel[0] = st.ctrl.root_lin_vel_d_rel[0]; lin_vel_d_rel[1] = st.ctrl.root_lin_vel_d_rel[1]; }
  if (mpc_solver_inited) *mpc_solver_inited = st.mpc_solver_inited ? 1 : 0;
  if (estimation_inited && (Pdiag || qv || A || l || u)) dense_from_captured(t->mpc->fastConvex.solver.captured(), Pdiag, qv, A, l, u);
  return ok ? 0 : 1;
}

/* ---- Utils / A1Kinematics (downstream rows of SURVEY §8f-3) ---- */
void ref_skew(const double v[3], double out9[9]) {
  Eigen::Matrix3d m = Utils::skew(Eigen::Vector3d(v[0], v[1], v[2]));
  for (int k = 0; k < 9; ++k) out9[k] = m.data()[k];
}
void ref_a1_fk(const double q[3], const double rho_opt[3], const double rho_fix[5], double p[3]) {
  legged::A1Kinematics kin; Eigen::VectorXd ro(3), rf(5);
  for (int i = 0; i < 3; ++i) ro[i] = rho_opt[i];
  for (int i = 0; i < 5; ++i) rf[i] = rho_fix[i];
  Eigen::Vector3d r = kin.fk(Eigen::Vector3d(q[0], q[1], q[2]), ro, rf);
  for (int i = 0; i < 3; ++i) p[i] = r[i];
}
void ref_a1_jac(const double q[3], const double rho_opt[3], const double rho_fix[5], double J9[9]) {
  legged::A1Kinematics kin; Eigen::VectorXd ro(3), rf(5);
  for (int i = 0; i < 3; ++i) ro[i] = rho_opt[i];
  for (int i = 0; i < 5; ++i) rf[i] = rho_fix[i];
  Eigen::Matrix3d J = kin.jac(Eigen::Vector3d(q[0], q[1], q[2]), ro, rf);
  for (int k = 0; k < 9; ++k) J9[k] = J.data()[k];
}
void ref_a1_inv_kin(const double p[3], const double cur_q[3], const double rho_opt[3], const double rho_fix[5], double q[3]) {
  legged::A1Kinematics kin; Eigen::VectorXd ro(3), rf(5);
  for (int i = 0; i < 3; ++i) ro[i] = rho_opt[i];
  for (int i = 0; i < 5; ++i) rf[i] = rho_fix[i];
  Eigen::Vector3d r = kin.inv_kin(Eigen::Vector3d(p[0], p[1], p[2]), Eigen::Vector3d(cur_q[0], cur_q[1], cur_q[2]), ro, rf);
  for (int i = 0; i < 3; ++i) q[i] = r[i];
}
void ref_bezier_foot_curve(double t, const double start[3], const double final_[3], double terrain_pitch, double out6[6]) {
  BezierUtils b;
  Eigen::Matrix<double, 6, 1> r = b.get_foot_pos_curve((float)t, Eigen::Vector3d(start[0], start[1], start[2]),
                                                       Eigen::Vector3d(final_[0], final_[1], final_[2]), terrain_pitch);
  for (int i = 0; i < 6; ++i) out6[i] = r[i];
}

}  /* extern "C" */
