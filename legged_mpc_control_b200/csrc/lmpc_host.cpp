// lmpc_host.cpp — host-only parts of the C ABI: parameter presets and the gait schedule.
// No CUDA here.  Reference citations are relative to /root/reference/src/legged_ctrl/.
#include <stddef.h>
#include <string.h>

#include "../../include/lmpc.h"
#include "lmpc_gait.h"

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
  return lmpc::gait_predict(gait, leg, phase, gait_speed, delta_t);
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

// Layout self-description for FFI bindings (tests/test_host.py checks the ctypes mirror against it).
extern "C" int64_t lmpc_abi_sizeof(int32_t which, int32_t field) {
#define OFF(T, m) (int64_t) offsetof(T, m)
  static const int64_t params[] = {OFF(lmpc_params, horizon), OFF(lmpc_params, discretisation), OFF(lmpc_params, cost_scaling),
                                   OFF(lmpc_params, xref_index_offset), OFF(lmpc_params, precision), OFF(lmpc_params, solver),
                                   OFF(lmpc_params, max_iter), OFF(lmpc_params, reserved0), OFF(lmpc_params, dt), OFF(lmpc_params, mu),
                                   OFF(lmpc_params, fz_max), OFF(lmpc_params, mass), OFF(lmpc_params, gravity), OFF(lmpc_params, active_tol),
                                   OFF(lmpc_params, inertia), OFF(lmpc_params, q), OFF(lmpc_params, r)};
  static const int64_t bin[] = {OFF(lmpc_batch_in, x0), OFF(lmpc_batch_in, rot), OFF(lmpc_batch_in, foot_pos_abs), OFF(lmpc_batch_in, cmd),
                                OFF(lmpc_batch_in, contact), OFF(lmpc_batch_in, q), OFF(lmpc_batch_in, r), OFF(lmpc_batch_in, mu),
                                OFF(lmpc_batch_in, fz_max), OFF(lmpc_batch_in, mass), OFF(lmpc_batch_in, inertia), OFF(lmpc_batch_in, on_device),
                                OFF(lmpc_batch_in, reserved0)};
  static const int64_t bout[] = {OFF(lmpc_batch_out, grf), OFF(lmpc_batch_out, status), OFF(lmpc_batch_out, iters),
                                 OFF(lmpc_batch_out, active_set), OFF(lmpc_batch_out, U), OFF(lmpc_batch_out, on_device),
                                 OFF(lmpc_batch_out, reserved0)};
  static const int64_t rst[] = {OFF(lmpc_rollout_state, x), OFF(lmpc_rollout_state, foot_world), OFF(lmpc_rollout_state, phase),
                                OFF(lmpc_rollout_state, gait), OFF(lmpc_rollout_state, gait_speed), OFF(lmpc_rollout_state, cmd),
                                OFF(lmpc_rollout_state, on_device), OFF(lmpc_rollout_state, reserved0),
                                OFF(lmpc_rollout_state, fsm_state), OFF(lmpc_rollout_state, fsm_index), OFF(lmpc_rollout_state, fsm_start),
                                OFF(lmpc_rollout_state, fsm_end)};
  static const int64_t rcfg[] = {OFF(lmpc_rollout_cfg, steps), OFF(lmpc_rollout_cfg, warm_start), OFF(lmpc_rollout_cfg, default_foot_pos_rel),
                                 OFF(lmpc_rollout_cfg, foot_delta_x_limit), OFF(lmpc_rollout_cfg, foot_delta_y_limit), OFF(lmpc_rollout_cfg, ground_z)};
  static const int64_t rres[] = {OFF(lmpc_rollout_result, grf), OFF(lmpc_rollout_result, x_log), OFF(lmpc_rollout_result, grf_log),
                                 OFF(lmpc_rollout_result, solves), OFF(lmpc_rollout_result, factorizations), OFF(lmpc_rollout_result, polish_rounds),
                                 OFF(lmpc_rollout_result, not_converged), OFF(lmpc_rollout_result, warm_hits), OFF(lmpc_rollout_result, flops_executed),
                                 OFF(lmpc_rollout_result, mean_n_eff), OFF(lmpc_rollout_result, device_ms), OFF(lmpc_rollout_result, reserved0)};
  static const int64_t fsm[] = {OFF(lmpc_fsm_io, gait_phase), OFF(lmpc_fsm_io, fsm_state), OFF(lmpc_fsm_io, fsm_index), OFF(lmpc_fsm_io, fsm_start),
                                OFF(lmpc_fsm_io, fsm_end), OFF(lmpc_fsm_io, gait), OFF(lmpc_fsm_io, gait_speed), OFF(lmpc_fsm_io, foot_contact_flag),
                                OFF(lmpc_fsm_io, plan_contacts), OFF(lmpc_fsm_io, contact_table), OFF(lmpc_fsm_io, on_device), OFF(lmpc_fsm_io, reserved0)};
  static const int64_t leg[] = {OFF(lmpc_leg_io, joint_pos), OFF(lmpc_leg_io, joint_vel), OFF(lmpc_leg_io, rot), OFF(lmpc_leg_io, root_pos),
                                OFF(lmpc_leg_io, root_lin_vel), OFF(lmpc_leg_io, imu_ang_vel), OFF(lmpc_leg_io, grf), OFF(lmpc_leg_io, foot_pos_target_world),
                                OFF(lmpc_leg_io, foot_vel_target_world), OFF(lmpc_leg_io, movement_mode), OFF(lmpc_leg_io, foot_pos_rel), OFF(lmpc_leg_io, j_foot),
                                OFF(lmpc_leg_io, foot_vel_rel), OFF(lmpc_leg_io, foot_pos_abs), OFF(lmpc_leg_io, foot_vel_abs), OFF(lmpc_leg_io, foot_pos_world),
                                OFF(lmpc_leg_io, foot_vel_world), OFF(lmpc_leg_io, joint_tau_tgt), OFF(lmpc_leg_io, joint_ang_tgt),
                                OFF(lmpc_leg_io, prev_joint_ang_tgt), OFF(lmpc_leg_io, joint_vel_tgt), OFF(lmpc_leg_io, on_device), OFF(lmpc_leg_io, reserved0)};
  static const int64_t geo[] = {OFF(lmpc_leg_geometry, rho_fix), OFF(lmpc_leg_geometry, rho_opt)};
#undef OFF
  static const struct { const int64_t* off; int n; int64_t size; } tab[] = {
      {params, (int)(sizeof(params) / 8), (int64_t)sizeof(lmpc_params)},         {bin, (int)(sizeof(bin) / 8), (int64_t)sizeof(lmpc_batch_in)},
      {bout, (int)(sizeof(bout) / 8), (int64_t)sizeof(lmpc_batch_out)},          {rst, (int)(sizeof(rst) / 8), (int64_t)sizeof(lmpc_rollout_state)},
      {rcfg, (int)(sizeof(rcfg) / 8), (int64_t)sizeof(lmpc_rollout_cfg)},        {rres, (int)(sizeof(rres) / 8), (int64_t)sizeof(lmpc_rollout_result)},
      {fsm, (int)(sizeof(fsm) / 8), (int64_t)sizeof(lmpc_fsm_io)},               {leg, (int)(sizeof(leg) / 8), (int64_t)sizeof(lmpc_leg_io)},
      {geo, (int)(sizeof(geo) / 8), (int64_t)sizeof(lmpc_leg_geometry)}};
  if (which < 0 || which >= 9) return -1;
  if (field < 0) return tab[which].size;
  return field < tab[which].n ? tab[which].off[field] : -1;
}
