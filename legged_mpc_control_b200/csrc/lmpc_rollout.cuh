// lmpc_rollout.cuh — device gait scheduler + closed-loop single-rigid-body rollout (BASELINE
// config 5).  Headless replacement of the loop the reference closes through
// Gazebo: every control step, for every robot,
//   rollout_prepare_kernel   contact table from the gait phase (LeggedContactFSM::predict_contact_state,
//                            src/utils/LeggedContactFSM.cpp:280-294 via ConvexQPSolver.cpp:329-346),
//                            Raibert footholds for swing legs (src/interfaces/BaseInterface.cpp:358-399),
//                            rotation matrix, foot_pos_abs -> the solver's input arrays
//   solve kernels            lmpc_api.cu (warm-started from the previous tick's faces)
//   rollout_step_kernel      x+ = A_d x + B_d u - g dt e_vz with the solver's own discretisation
//                            (ConvexQPSolver.cpp:198-228), gait phase advance (LeggedContactFSM.cpp:50)
// The robot state never leaves the GPU between steps.  Included by lmpc_api.cu only.
#pragma once
#include "lmpc_gait.h"

namespace lmpc {

struct RolloutDev {
  double* x;           // [B][12]
  double* foot_world;  // [B][12]
  double* phase;       // [B][4]
  const int* gait;     // [B]
  const double* gait_speed;  // [B]
  const double* cmd;   // [B][9]
  // solver inputs written by the prepare kernel
  double* x0; double* rot; double* foot_abs; uint8_t* contact;
  // solver output read by the step kernel
  const double* grf;
  double* x_log; double* grf_log;  // optional [steps][B][12]
  // optional contact state machine (LeggedContactFSM::update): all four or none, [B][4]
  int* fsm_state; int* fsm_index; double* fsm_start; double* fsm_end;
  int B;
};

struct RolloutCfgDev {
  double foot_rel[12];  // default_foot_pos_rel, 3x4 col-major (LeggedState.cpp:120-133)
  double dlim_x, dlim_y, ground_z;
};

// Rz(yaw) Ry(pitch) Rx(roll), column-major.
__device__ inline void euler_to_rot(double roll, double pitch, double yaw, double* R) {
  double sr, cr, sp, cp, sy, cy;
  sincos(roll, &sr, &cr); sincos(pitch, &sp, &cp); sincos(yaw, &sy, &cy);
  R[0] = cy * cp; R[1] = sy * cp; R[2] = -sp;
  R[3] = cy * sp * sr - sy * cr; R[4] = sy * sp * sr + cy * cr; R[5] = cp * sr;
  R[6] = cy * sp * cr + sy * sr; R[7] = sy * sp * cr - cy * sr; R[8] = cp * cr;
}

__global__ void rollout_prepare_kernel(const __grid_constant__ KParams P, const __grid_constant__ RolloutDev D,
                                       const __grid_constant__ RolloutCfgDev C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= D.B) return;
  const size_t p = (size_t)i;
  const int H = P.H;
  double x[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) x[k] = D.x[12 * p + k];
  const double* cmd = D.cmd + 9 * p;
  const int gait = D.gait[p];
  const double speed = D.gait_speed[p];
  double sy, cy;
  sincos(x[2], &sy, &cy);
  // Raibert heuristic, BaseInterface.cpp:360-397
  const double vdx = cy * cmd[3] - sy * cmd[4], vdy = sy * cmd[3] + cy * cmd[4];
  const double kk = sqrt(fabs(x[5]) / 9.8);
  const double half_stance = (1.0 / speed / 2.0) / 2.0;
  double dx = kk * (x[9] - vdx) + half_stance * vdx;
  double dy = kk * (x[10] - vdy) + half_stance * vdy;
  dx = fmin(fmax(dx, -C.dlim_x), C.dlim_x);
  dy = fmin(fmax(dy, -C.dlim_y), C.dlim_y);
  uint8_t* ct = D.contact + (size_t)(4 * H) * p;
  for (int leg = 0; leg < 4; ++leg) {
    const double ph = D.phase[4 * p + leg];
    for (int k = 0; k < H; ++k) ct[4 * k + leg] = (uint8_t)gait_predict(gait, leg, ph, speed, __dmul_rn((double)k, P.dt));
    if (D.fsm_state) ct[leg] = (uint8_t)D.fsm_state[4 * p + leg];  // row 0 = plan_contacts = the FSM's own state (ConvexMpc.cpp:101-104)
    double* fw = D.foot_world + 12 * p + 3 * leg;
    if (!ct[leg]) {  // swing: the foot tracks its foothold target (lands where the target is at touchdown)
      const double rx = C.foot_rel[3 * leg], ry = C.foot_rel[3 * leg + 1];
      fw[0] = cy * rx - sy * ry + dx + x[3];
      fw[1] = sy * rx + cy * ry + dy + x[4];
      fw[2] = C.ground_z;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) D.foot_abs[12 * p + 3 * leg + a] = fw[a] - x[3 + a];
  }
  double R[9];
  euler_to_rot(x[0], x[1], x[2], R);
#pragma unroll
  for (int k = 0; k < 9; ++k) D.rot[9 * p + k] = R[k];
#pragma unroll
  for (int k = 0; k < 12; ++k) D.x0[12 * p + k] = x[k];
}

struct RolloutTotals {
  unsigned long long solves, sum_n, sum_fact, sum_polish, not_ok, ipm_only, warm_hits;
  double flops;
};

__global__ void rollout_step_kernel(const __grid_constant__ KParams P, const __grid_constant__ RolloutDev D,
                                    const double* __restrict__ per_inertia, const double* __restrict__ per_mass,
                                    const KStats* __restrict__ stats, RolloutTotals* totals, int step) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {  // the solve kernels of this step are complete (same stream): fold their counters
    totals->solves += stats->problems; totals->sum_n += stats->sum_n; totals->sum_fact += stats->sum_fact;
    totals->sum_polish += stats->sum_polish; totals->not_ok += stats->not_ok; totals->ipm_only += stats->ipm_only;
    totals->warm_hits += stats->warm_hits; totals->flops += stats->flops;
  }
  if (i >= D.B) return;
  const size_t p = (size_t)i;
  double x[12], f[12], R[9], I[9], RI[9], Iw[9];
#pragma unroll
  for (int k = 0; k < 12; ++k) { x[k] = D.x[12 * p + k]; f[k] = D.grf[12 * p + k]; }
#pragma unroll
  for (int k = 0; k < 9; ++k) { R[k] = D.rot[9 * p + k]; I[k] = per_inertia ? per_inertia[9 * p + k] : P.inertia[k]; }
  const double mass = per_mass ? per_mass[p] : P.mass;
  // I_w = R I R', closed-form inverse (same arithmetic as the solver's SRBD build)
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) RI[r + 3 * c] = R[r] * I[3 * c] + R[r + 3] * I[1 + 3 * c] + R[r + 6] * I[2 + 3 * c];
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) Iw[r + 3 * c] = RI[r] * R[c] + RI[r + 3] * R[c + 3] + RI[r + 6] * R[c + 6];
  const double a = Iw[0], b = Iw[3], c = Iw[6], d = Iw[1], ee = Iw[4], ff = Iw[7], g = Iw[2], h = Iw[5], ii = Iw[8];
  const double c00 = ee * ii - ff * h, c01 = -(d * ii - ff * g), c02 = d * h - ee * g;
  const double c10 = -(b * ii - c * h), c11 = a * ii - c * g, c12 = -(a * h - b * g);
  const double c20 = b * ff - c * ee, c21 = -(a * ff - c * d), c22 = a * ee - b * d;
  const double id = 1.0 / (a * c00 + b * c01 + c * c02);
  const double Ii[9] = {c00 * id, c01 * id, c02 * id, c10 * id, c11 * id, c12 * id, c20 * id, c21 * id, c22 * id};
  double tau[3] = {0, 0, 0}, fs[3] = {0, 0, 0};
  for (int leg = 0; leg < 4; ++leg) {
    const double* r = D.foot_abs + 12 * p + 3 * leg;
    const double* fl = f + 3 * leg;
    tau[0] += r[1] * fl[2] - r[2] * fl[1];
    tau[1] += r[2] * fl[0] - r[0] * fl[2];
    tau[2] += r[0] * fl[1] - r[1] * fl[0];
    fs[0] += fl[0]; fs[1] += fl[1]; fs[2] += fl[2];
  }
  double al[3], ac[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) al[r] = Ii[r] * tau[0] + Ii[r + 3] * tau[1] + Ii[r + 6] * tau[2];
  ac[0] = fs[0] / mass; ac[1] = fs[1] / mass; ac[2] = fs[2] / mass - P.gravity;
  const double dt = P.dt, th = P.zoh ? 0.5 : 0.0;
  double sy, cy;
  sincos(x[2], &sy, &cy);
  // A_c = [0 0 T 0; 0 0 0 I; 0; 0] (T = Rz(yaw)' acting on omega), A_c^2 = 0:
  //   x+ = x + dt A_c x + (dt B_c + th dt^2 A_c B_c) u + gravity terms
  const double w0 = x[6] + th * dt * al[0], w1 = x[7] + th * dt * al[1], w2 = x[8] + th * dt * al[2];
  x[0] += dt * (cy * w0 + sy * w1);
  x[1] += dt * (-sy * w0 + cy * w1);
  x[2] += dt * w2;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    x[3 + r] += dt * (x[9 + r] + th * dt * ac[r]);
    x[6 + r] += dt * al[r];
    x[9 + r] += dt * ac[r];
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) D.x[12 * p + k] = x[k];
  const double speed = D.gait_speed[p];
  for (int leg = 0; leg < 4 && D.fsm_state; ++leg) {  // the reference's state machine (LeggedContactFSM.cpp:38-84); no contact sensing in this plant
    const size_t k = 4 * p + leg;
    FsmLeg f = {D.phase[k], D.fsm_start[k], D.fsm_end[k], D.fsm_index[k], D.fsm_state[k]};
    fsm_update(D.gait[p], leg, speed, dt, false, f);
    D.phase[k] = f.phase; D.fsm_start[k] = f.start; D.fsm_end[k] = f.end; D.fsm_index[k] = f.index; D.fsm_state[k] = f.state;
  }
  for (int leg = 0; leg < 4 && !D.fsm_state; ++leg) {  // stateless variant: wrapped phase only
    double ph = __dadd_rn(D.phase[4 * p + leg], __dmul_rn(speed, dt));  // LeggedContactFSM.cpp:50 (unfused)
    if (ph > 1.0 && ph <= 4503599627370496.0) ph -= ceil(ph) - 1.0;  // == `while (ph > 1) ph -= 1` bit for bit (lmpc_gait.h), O(1)
    D.phase[4 * p + leg] = ph;
  }
  if (D.x_log) {
#pragma unroll
    for (int k = 0; k < 12; ++k) D.x_log[((size_t)step * D.B + p) * 12 + k] = x[k];
  }
  if (D.grf_log) {
#pragma unroll
    for (int k = 0; k < 12; ++k) D.grf_log[((size_t)step * D.B + p) * 12 + k] = f[k];
  }
}

}  // namespace lmpc
