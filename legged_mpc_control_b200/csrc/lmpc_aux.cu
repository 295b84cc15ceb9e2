// lmpc_aux.cu — the batched stages either side of the MPC solve: the contact state
// machine (LeggedContactFSM::update) and the GRF -> joint command map (BaseInterface::tau_ctrl_update with
// the A1 leg kinematics).  One thread per robot-leg; both are HBM-bound streaming kernels (a few hundred
// bytes per robot), sized in whole waves of the SM count by the launch helper.
#include <string.h>

#include <vector>

#include "lmpc_internal.h"
#include "lmpc_gait.h"
#include "lmpc_kin.h"

namespace lmpc {

// ---------------------------------------------------------------- contact state machine
struct FsmDev {
  double* phase; int* state; int* index; double* start; double* end;
  const int* gait; const double* speed; const double* flag;
  uint8_t* plan; uint8_t* table;
  int B, H;
  double mpc_dt;
};

__global__ void fsm_kernel(const __grid_constant__ FsmDev D, double dt, int do_reset) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // robot-leg
  if (i >= 4 * D.B) return;
  const int p = i >> 2, leg = i & 3;
  const int gait = D.gait[p];
  const double speed = D.speed[p];
  FsmLeg f;
  if (do_reset) {
    fsm_reset(gait, leg, f);
  } else {
    f.phase = D.phase[i]; f.state = D.state[i]; f.index = D.index[i]; f.start = D.start[i]; f.end = D.end[i];
    fsm_update(gait, leg, speed, dt, D.flag ? D.flag[i] != 0.0 : false, f);
  }
  D.phase[i] = f.phase; D.state[i] = f.state; D.index[i] = f.index; D.start[i] = f.start; D.end[i] = f.end;
  if (D.plan) D.plan[i] = (uint8_t)f.state;
  if (D.table) {  // row 0 = plan_contacts, row k = predict_contact_state(k * mpc_dt) (ConvexQPSolver.cpp:329-346)
    uint8_t* t = D.table + (size_t)(4 * D.H) * p + leg;
    t[0] = (uint8_t)f.state;
    for (int k = 1; k < D.H; ++k) t[4 * k] = (uint8_t)gait_predict(gait, leg, f.phase, speed, __dmul_rn((double)k, D.mpc_dt));
  }
}

// ---------------------------------------------------------------- leg kinematics / torque map
struct LegDev {
  const double *joint_pos, *joint_vel, *rot, *root_pos, *root_lin_vel, *imu_ang_vel, *grf, *ptgt, *vtgt, *mode;
  double *foot_pos_rel, *j_foot, *foot_vel_rel, *foot_pos_abs, *foot_vel_abs, *foot_pos_world, *foot_vel_world;
  double *tau, *ang_tgt, *prev_ang_tgt, *vel_tgt;
  int B;
};
struct GeomDev { LegGeom leg[4]; };

__device__ __forceinline__ void rot_mul(const double* R, const double* v, double* o) {  // o = R v
  for (int r = 0; r < 3; ++r) o[r] = R[r] * v[0] + R[r + 3] * v[1] + R[r + 6] * v[2];
}
__device__ __forceinline__ void rot_tmul(const double* R, const double* v, double* o) {  // o = R' v
  for (int c = 0; c < 3; ++c) o[c] = R[3 * c] * v[0] + R[3 * c + 1] * v[1] + R[3 * c + 2] * v[2];
}

// BaseInterface.cpp:280-300
__global__ void leg_kinematics_kernel(const __grid_constant__ LegDev D, const __grid_constant__ GeomDev G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 4 * D.B) return;
  const int p = i >> 2, leg = i & 3;
  const size_t o3 = 3 * (size_t)i;
  double q[3], dq[3], R[9], pr[3], J[9], vr[3], pa[3], va[3];
  for (int a = 0; a < 3; ++a) { q[a] = D.joint_pos[o3 + a]; dq[a] = D.joint_vel[o3 + a]; }
  for (int k = 0; k < 9; ++k) R[k] = D.rot[9 * (size_t)p + k];
  a1_fk(q, G.leg[leg], pr);
  a1_jac(q, G.leg[leg], J);
  for (int r = 0; r < 3; ++r) vr[r] = J[r] * dq[0] + J[r + 3] * dq[1] + J[r + 6] * dq[2];
  rot_mul(R, pr, pa);
  rot_mul(R, vr, va);
  for (int a = 0; a < 3; ++a) {
    if (D.foot_pos_rel) D.foot_pos_rel[o3 + a] = pr[a];
    if (D.foot_vel_rel) D.foot_vel_rel[o3 + a] = vr[a];
    if (D.foot_pos_abs) D.foot_pos_abs[o3 + a] = pa[a];
    if (D.foot_vel_abs) D.foot_vel_abs[o3 + a] = va[a];
    if (D.foot_pos_world) D.foot_pos_world[o3 + a] = pa[a] + D.root_pos[3 * (size_t)p + a];
  }
  if (D.j_foot)
    for (int k = 0; k < 9; ++k) D.j_foot[9 * (size_t)i + k] = J[k];
  if (D.foot_vel_world) {  // v_abs + v_root + R (omega x p_rel)   (:299-300)
    double w[3] = {0, 0, 0}, wxp[3], rw[3];
    if (D.imu_ang_vel) for (int a = 0; a < 3; ++a) w[a] = D.imu_ang_vel[3 * (size_t)p + a];
    wxp[0] = w[1] * pr[2] - w[2] * pr[1]; wxp[1] = w[2] * pr[0] - w[0] * pr[2]; wxp[2] = w[0] * pr[1] - w[1] * pr[0];
    rot_mul(R, wxp, rw);
    for (int a = 0; a < 3; ++a) D.foot_vel_world[o3 + a] = va[a] + D.root_lin_vel[3 * (size_t)p + a] + rw[a];
  }
}

// BaseInterface::tau_ctrl_update, BaseInterface.cpp:451-500
__global__ void tau_ctrl_kernel(const __grid_constant__ LegDev D, const __grid_constant__ GeomDev G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 4 * D.B) return;
  const int p = i >> 2, leg = i & 3;
  const size_t o3 = 3 * (size_t)i, p3 = 3 * (size_t)p;
  double q[3], dq[3], R[9], f[3], fr[3], J[9];
  for (int a = 0; a < 3; ++a) { q[a] = D.joint_pos[o3 + a]; dq[a] = D.joint_vel[o3 + a]; f[a] = D.grf[o3 + a]; }
  for (int k = 0; k < 9; ++k) R[k] = D.rot[9 * (size_t)p + k];
  a1_jac(q, G.leg[leg], J);  // fbk.j_foot block of this leg (:286-290)
  rot_tmul(R, f, fr);        // world -> body (:455-456)
  for (int c = 0; c < 3; ++c) D.tau[o3 + c] = -(J[3 * c] * fr[0] + J[3 * c + 1] * fr[1] + J[3 * c + 2] * fr[2]);  // -J' f (:459)
  double ang[3], vel[3];
  if (D.mode[p] > 0) {
    double d[3], pt[3], vt[3], qi[3], qd[3];
    for (int a = 0; a < 3; ++a) d[a] = D.ptgt[o3 + a] - D.root_pos[p3 + a];
    rot_tmul(R, d, pt);  // :464-465
    for (int a = 0; a < 3; ++a) d[a] = D.vtgt[o3 + a] - D.root_lin_vel[p3 + a];
    rot_tmul(R, d, vt);  // :468-469
    a1_inv_kin(pt, q, G.leg[leg], qi);
    const bool bad_q = isnan(qi[0]) || isnan(qi[1]) || isnan(qi[2]);  // :472-478
    solve3_pivoted(J, vt, qd);
    const bool bad_v = isnan(qd[0]) || isnan(qd[1]) || isnan(qd[2]);  // :480-485
    for (int a = 0; a < 3; ++a) { ang[a] = bad_q ? q[a] : qi[a]; vel[a] = bad_v ? dq[a] : qd[a]; }
    if (D.prev_ang_tgt) for (int a = 0; a < 3; ++a) D.prev_ang_tgt[o3 + a] = D.ang_tgt[o3 + a];
  } else {  // :486-490
    for (int a = 0; a < 3; ++a) { ang[a] = q[a]; vel[a] = dq[a]; }
    if (D.prev_ang_tgt) for (int a = 0; a < 3; ++a) D.prev_ang_tgt[o3 + a] = q[a];
  }
  for (int a = 0; a < 3; ++a) { D.ang_tgt[o3 + a] = ang[a]; D.vel_tgt[o3 + a] = vel[a]; }
}

// Host-pointer calls: every non-NULL array is staged through ONE handle-owned device buffer
// (h->d_aux, grown on demand and kept), carved into 256-byte aligned slices; no cudaMalloc per call
// once the buffer has reached the batch's size.
struct Stager {
  lmpc_handle* h;
  struct Out { void* host; void* dev; size_t bytes; };
  std::vector<Out> outs;
  bool on_device;
  size_t used = 0;
  cudaError_t err = cudaSuccess;
  // upper_bound: bytes of every array the call can stage plus 256 per array of alignment slack
  Stager(lmpc_handle* h_, bool dev_, size_t upper_bound) : h(h_), on_device(dev_) {
    if (!on_device) err = h->d_aux.ensure(upper_bound);
  }
  void* carve(size_t bytes) {
    if (err != cudaSuccess) return nullptr;
    const size_t at = (used + 255) & ~(size_t)255;
    if (at + bytes > h->d_aux.cap) { err = cudaErrorMemoryAllocation; return nullptr; }
    used = at + bytes;
    return (char*)h->d_aux.p + at;
  }
  template <class T>
  T* in(const T* host, size_t n) {
    if (!host || on_device) return const_cast<T*>(host);
    void* d = carve(n * sizeof(T));
    if (d) err = cudaMemcpyAsync(d, host, n * sizeof(T), cudaMemcpyHostToDevice, h->stream);
    return (T*)d;
  }
  template <class T>
  T* out(T* host, size_t n, bool also_in = false) {
    if (!host || on_device) return host;
    void* d = carve(n * sizeof(T));
    if (d) {
      outs.push_back({host, d, n * sizeof(T)});
      if (also_in) err = cudaMemcpyAsync(d, host, n * sizeof(T), cudaMemcpyHostToDevice, h->stream);
    }
    return (T*)d;
  }
  cudaError_t finish() {
    for (const Out& o : outs)
      if (err == cudaSuccess) err = cudaMemcpyAsync(o.host, o.dev, o.bytes, cudaMemcpyDeviceToHost, h->stream);
    if (err == cudaSuccess) err = cudaStreamSynchronize(h->stream);
    return err;
  }
};

static GeomDev geom_dev(const lmpc_leg_geometry* g) {
  GeomDev G;
  for (int l = 0; l < 4; ++l)
    G.leg[l] = LegGeom{g->rho_fix[l][0], g->rho_fix[l][1], g->rho_fix[l][2], g->rho_fix[l][3], g->rho_fix[l][4],
                       g->rho_opt[l][0], g->rho_opt[l][1], g->rho_opt[l][2]};
  return G;
}

}  // namespace lmpc

using lmpc::set_err;

static lmpc_status fsm_call(lmpc_handle* h, int32_t batch, double dt, const lmpc_fsm_io* io, int do_reset) {
  if (!h || !io) { set_err("NULL handle/io"); return LMPC_ERR_INVALID_ARG; }
  if (batch < 1) { set_err("batch must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  if (!io->gait_phase || !io->fsm_state || !io->fsm_index || !io->fsm_start || !io->fsm_end || !io->gait || !io->gait_speed) {
    set_err("gait_phase, fsm_state, fsm_index, fsm_start, fsm_end, gait, gait_speed are required"); return LMPC_ERR_INVALID_ARG;
  }
  CK(cudaSetDevice(h->device));
  const size_t B = (size_t)batch, H = (size_t)h->kp.H;
  // 5 per-leg state arrays (<= 8 B), gait + speed, contact flag, plan, table; 16 arrays of slack
  lmpc::Stager s(h, io->on_device != 0, B * (4 * 8 * 4 + 4 * 4 * 2 + 4 + 8 + 4 * 8 + 4 + 4 * H) + 16 * 256);
  lmpc::FsmDev D;
  D.phase = s.out(io->gait_phase, 4 * B, true); D.state = s.out(io->fsm_state, 4 * B, true); D.index = s.out(io->fsm_index, 4 * B, true);
  D.start = s.out(io->fsm_start, 4 * B, true); D.end = s.out(io->fsm_end, 4 * B, true);
  D.gait = s.in(io->gait, B); D.speed = s.in(io->gait_speed, B); D.flag = s.in(io->foot_contact_flag, 4 * B);
  D.plan = s.out(io->plan_contacts, 4 * B); D.table = s.out(io->contact_table, 4 * H * B);
  D.B = batch; D.H = h->kp.H; D.mpc_dt = h->kp.dt;
  CK(s.err);
  lmpc::fsm_kernel<<<(4 * batch + 127) / 128, 128, 0, h->stream>>>(D, dt, do_reset);
  h->launches++;
  CK(cudaGetLastError());
  CK(s.finish());  // blocking for host and device pointers alike
  return LMPC_OK;
}
extern "C" lmpc_status lmpc_fsm_reset(lmpc_handle* h, int32_t batch, const lmpc_fsm_io* io) { return fsm_call(h, batch, 0.0, io, 1); }
extern "C" lmpc_status lmpc_fsm_update(lmpc_handle* h, int32_t batch, double dt, const lmpc_fsm_io* io) { return fsm_call(h, batch, dt, io, 0); }

extern "C" void lmpc_leg_geometry_a1(lmpc_leg_geometry* g) {  // BaseInterface.cpp:72-98
  if (!g) return;
  memset(g, 0, sizeof(*g));
  const double ox[4] = {0.1805, 0.1805, -0.1805, -0.1805}, oy[4] = {0.047, -0.047, 0.047, -0.047};
  const double mo[4] = {0.0838, -0.0838, 0.0838, -0.0838};
  for (int l = 0; l < 4; ++l) {
    g->rho_fix[l][0] = ox[l]; g->rho_fix[l][1] = oy[l]; g->rho_fix[l][2] = mo[l];
    g->rho_fix[l][3] = 0.21; g->rho_fix[l][4] = 0.21;  // upper_leg_length, LOWER_LEG_LENGTH (LeggedParams.h)
  }
}

static lmpc_status leg_call(lmpc_handle* h, int32_t batch, const lmpc_leg_geometry* g, const lmpc_leg_io* io, bool tau) {
  if (!h || !g || !io) { set_err("NULL handle/geometry/io"); return LMPC_ERR_INVALID_ARG; }
  if (batch < 1) { set_err("batch must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  if (!io->joint_pos || !io->joint_vel || !io->rot) { set_err("joint_pos, joint_vel, rot are required"); return LMPC_ERR_INVALID_ARG; }
  if (tau && (!io->grf || !io->foot_pos_target_world || !io->foot_vel_target_world || !io->movement_mode || !io->root_pos ||
              !io->root_lin_vel || !io->joint_tau_tgt || !io->joint_ang_tgt || !io->joint_vel_tgt)) {
    set_err("lmpc_tau_ctrl needs grf, foot_pos/vel_target_world, movement_mode, root_pos, root_lin_vel and the three joint targets");
    return LMPC_ERR_INVALID_ARG;
  }
  if (!tau && ((io->foot_pos_world && !io->root_pos) || (io->foot_vel_world && !io->root_lin_vel))) {
    set_err("foot_pos_world needs root_pos, foot_vel_world needs root_lin_vel"); return LMPC_ERR_INVALID_ARG;
  }
  CK(cudaSetDevice(h->device));
  const size_t B = (size_t)batch;
  // <= 16 arrays of 12 doubles, j_foot (36), rot (9), three 3-vectors, mode; 32 arrays of slack
  lmpc::Stager s(h, io->on_device != 0, B * 8 * (12 * 16 + 36 + 9 + 3 * 3 + 1) + 32 * 256);
  lmpc::LegDev D;
  memset(&D, 0, sizeof(D));
  D.joint_pos = s.in(io->joint_pos, 12 * B); D.joint_vel = s.in(io->joint_vel, 12 * B); D.rot = s.in(io->rot, 9 * B);
  D.root_pos = s.in(io->root_pos, 3 * B); D.root_lin_vel = s.in(io->root_lin_vel, 3 * B);
  if (tau) {
    D.grf = s.in(io->grf, 12 * B); D.ptgt = s.in(io->foot_pos_target_world, 12 * B); D.vtgt = s.in(io->foot_vel_target_world, 12 * B);
    D.mode = s.in(io->movement_mode, B);
    D.tau = s.out(io->joint_tau_tgt, 12 * B); D.ang_tgt = s.out(io->joint_ang_tgt, 12 * B, true);
    D.prev_ang_tgt = s.out(io->prev_joint_ang_tgt, 12 * B); D.vel_tgt = s.out(io->joint_vel_tgt, 12 * B);
  } else {
    D.imu_ang_vel = s.in(io->imu_ang_vel, 3 * B);
    D.foot_pos_rel = s.out(io->foot_pos_rel, 12 * B); D.j_foot = s.out(io->j_foot, 36 * B); D.foot_vel_rel = s.out(io->foot_vel_rel, 12 * B);
    D.foot_pos_abs = s.out(io->foot_pos_abs, 12 * B); D.foot_vel_abs = s.out(io->foot_vel_abs, 12 * B);
    D.foot_pos_world = s.out(io->foot_pos_world, 12 * B); D.foot_vel_world = s.out(io->foot_vel_world, 12 * B);
  }
  D.B = batch;
  CK(s.err);
  const lmpc::GeomDev G = lmpc::geom_dev(g);
  if (tau) lmpc::tau_ctrl_kernel<<<(4 * batch + 127) / 128, 128, 0, h->stream>>>(D, G);
  else lmpc::leg_kinematics_kernel<<<(4 * batch + 127) / 128, 128, 0, h->stream>>>(D, G);
  h->launches++;
  CK(cudaGetLastError());
  CK(s.finish());  // blocking for host and device pointers alike
  return LMPC_OK;
}
extern "C" lmpc_status lmpc_leg_kinematics(lmpc_handle* h, int32_t batch, const lmpc_leg_geometry* g, const lmpc_leg_io* io) {
  return leg_call(h, batch, g, io, false);
}
extern "C" lmpc_status lmpc_tau_ctrl(lmpc_handle* h, int32_t batch, const lmpc_leg_geometry* g, const lmpc_leg_io* io) {
  return leg_call(h, batch, g, io, true);
}
