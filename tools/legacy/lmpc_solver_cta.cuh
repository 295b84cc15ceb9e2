// lmpc_solver.cuh — device code of the fused convex-MPC solve (one CTA per QP).
//
// Stages per problem, all on chip (H never leaves shared memory unless the size class spills):
//   1. SRBD build:   I_w^-1 [r]x, yaw references          (ConvexQPSolver.cpp:198-228, 254-283)
//   2. discretise:   forward Euler (reference) or exact ZOH (Ac^2 = 0 => closed form)
//   3. condense:     swing-leg-eliminated Hessian H_r and gradient g_r in closed block form
//   4. QP solve:     feasible-start Mehrotra predictor-corrector interior point with an
//                    in-shared-memory Cholesky, then an exact active-set polish on the
//                    identified faces of the friction pyramids, KKT-verified
//   5. outputs:      u_0 (4x3 GRF), status, iteration counts, primal-active bitmask
//
// Matrix storage: packed row-major lower triangle, LIDX(i,j) = i(i+1)/2 + j (i >= j).  With one
// thread per row, lanes read element k of 16 consecutive rows at offsets T(i)+k, and triangular
// numbers are a permutation mod 16, so those accesses are bank-conflict free without padding.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "lmpc_types.cuh"

namespace lmpc {

#define LIDX(i, j) ((((i) * ((i) + 1)) >> 1) + (j))

__host__ __device__ inline int packed_size(int n) { return (n * (n + 1)) >> 1; }

// ---------------------------------------------------------------- block-wide reductions
template <int NT>
__device__ inline double block_reduce_sum(double v, double* red) {
  if (NT == 1) return v;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();  // protect `red` from the previous reduction's readers
  if (lane == 0) red[w] = v;
  __syncthreads();
  double r = 0;
#pragma unroll
  for (int i = 0; i < NT / 32; ++i) r += red[i];
  return r;
}
template <int NT>
__device__ inline double block_reduce_min(double v, double* red) {
  if (NT == 1) return v;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  double r = red[0];
#pragma unroll
  for (int i = 1; i < NT / 32; ++i) r = fmin(r, red[i]);
  return r;
}
template <int NT>
__device__ inline double block_reduce_max(double v, double* red) {
  return -block_reduce_min<NT>(-v, red);
}

// ---------------------------------------------------------------- dense kernels (packed lower)

// y = S x for symmetric S stored packed-lower.  One thread per row.
template <int NT>
__device__ inline void symv(const double* __restrict__ S, int n, const double* __restrict__ x,
                            double* __restrict__ y) {
  for (int i = threadIdx.x; i < n; i += NT) {
    const double* Si = S + LIDX(i, 0);
    double a0 = 0, a1 = 0;
    int j = 0;
    for (; j + 1 <= i; j += 2) { a0 += Si[j] * x[j]; a1 += Si[j + 1] * x[j + 1]; }
    if (j <= i) a0 += Si[j] * x[j];
    for (j = i + 1; j + 1 < n; j += 2) {
      a0 += S[LIDX(j, i)] * x[j];
      a1 += S[LIDX(j + 1, i)] * x[j + 1];
    }
    if (j < n) a0 += S[LIDX(j, i)] * x[j];
    y[i] = a0 + a1;
  }
}

// In-place Cholesky M = L L' (left-looking by columns, one thread per row).  The diagonal keeps
// the unnormalised pivot d_j; rinv[j] = 1/sqrt(d_j) = 1/L_jj.  Returns false on a non-positive
// pivot (uniform across the block).
template <int NT>
__device__ inline bool chol_factor(double* M, double* rinv, int n) {
  bool ok = true;
  for (int j = 0; j < n; ++j) {
    const double* Lj = M + LIDX(j, 0);
    for (int i = j + threadIdx.x; i < n; i += NT) {
      double* Li = M + LIDX(i, 0);
      double s0 = Li[j], s1 = 0.0;
      int k = 0;
      for (; k + 1 < j; k += 2) { s0 -= Li[k] * Lj[k]; s1 -= Li[k + 1] * Lj[k + 1]; }
      if (k < j) s0 -= Li[k] * Lj[k];
      Li[j] = s0 + s1;
    }
    __syncthreads();
    const double d = M[LIDX(j, j)];
    if (!(d > 0.0)) ok = false;
    const double ri = rsqrt(d > 0.0 ? d : 1.0);
    for (int i = j + 1 + threadIdx.x; i < n; i += NT) M[LIDX(i, j)] *= ri;
    if (threadIdx.x == 0) rinv[j] = ri;
    __syncthreads();
  }
  return ok;
}

// Solve L L' x = b in place (b in shared memory).
template <int NT>
__device__ inline void chol_solve(const double* __restrict__ M, const double* __restrict__ rinv,
                                  int n, double* b) {
  for (int j = 0; j < n; ++j) {  // forward, column oriented
    const double yj = b[j] * rinv[j];
    for (int i = j + 1 + threadIdx.x; i < n; i += NT) b[i] -= M[LIDX(i, j)] * yj;
    __syncthreads();
  }
  for (int i = threadIdx.x; i < n; i += NT) b[i] *= rinv[i];
  __syncthreads();
  for (int j = n - 1; j >= 0; --j) {  // backward
    const double xj = b[j] * rinv[j];
    const double* Lj = M + LIDX(j, 0);
    for (int i = threadIdx.x; i < j; i += NT) b[i] -= Lj[i] * xj;
    __syncthreads();
  }
  for (int i = threadIdx.x; i < n; i += NT) b[i] *= rinv[i];
  __syncthreads();
}

// ---------------------------------------------------------------- friction-pyramid blocks
// Rows of one stance block, written  C_b f <= d_b  with f = (fx, fy, fz):
//   c0: -fx - mu fz <= 0      (reference row 0:  0 <= fx + mu fz,   ConvexQPSolver.cpp:133,138)
//   c1:  fx - mu fz <= 0      (reference row 1:  fx - mu fz <= 0,   :134,139)
//   c2: -fy - mu fz <= 0      (reference row 2,                     :135,140)
//   c3:  fy - mu fz <= 0      (reference row 3,                     :136,141)
//   c4:  fz <= fz_max         (upper side of the bound row,         :163,171)
// (the lower side 0 <= fz is implied by c0+c1 and is reported through the apex face.)
__device__ inline void blk_slack(const double* f, double mu, double fmax, double* s) {
  s[0] = f[0] + mu * f[2];
  s[1] = mu * f[2] - f[0];
  s[2] = f[1] + mu * f[2];
  s[3] = mu * f[2] - f[1];
  s[4] = fmax - f[2];
}
// ds = -C_b df  (same linear form as the slack, without the constant)
__device__ inline void blk_dslack(const double* df, double mu, double* ds) {
  ds[0] = df[0] + mu * df[2];
  ds[1] = mu * df[2] - df[0];
  ds[2] = df[1] + mu * df[2];
  ds[3] = mu * df[2] - df[1];
  ds[4] = -df[2];
}
// r = C_b' w
__device__ inline void blk_ct(const double* w, double mu, double* r) {
  r[0] = w[1] - w[0];
  r[1] = w[3] - w[2];
  r[2] = w[4] - mu * (w[0] + w[1] + w[2] + w[3]);
}

// ---------------------------------------------------------------- shared-memory plan
struct SmemPlan {
  int ncap;      // matrix capacity (max reduced dimension of this launch), multiple of 3
  int H;
  bool mats_in_smem;
  // offsets in doubles
  int oH, oM, oU, oGS, oRD, oDUA, oDU, oTMP, oRINV, oPV, oZ, oBW, oCOS, oSIN, oCS, oSS, oE, oAJ,
      oBJ, oPRM, oRED, n_doubles;
  // offsets in bytes (after the doubles)
  int oBLK, oFACE, oBMAP, oINT, oACT, total_bytes;
};

__host__ __device__ inline SmemPlan make_plan(int H, int ncap, bool mats_in_smem) {
  SmemPlan p;
  p.ncap = ncap; p.H = H; p.mats_in_smem = mats_in_smem;
  int o = 0;
  const int ps = packed_size(ncap), nb = ncap / 3;
  p.oH = o; if (mats_in_smem) o += ps;
  p.oM = o; if (mats_in_smem) o += ps;
  p.oU = o; o += ncap;
  p.oGS = o; o += ncap;
  p.oRD = o; o += ncap;
  p.oDUA = o; o += ncap;
  p.oDU = o; o += ncap;
  p.oTMP = o; o += ncap;
  p.oRINV = o; o += ncap;
  p.oPV = o; o += ncap;
  p.oZ = o; o += 5 * nb;
  p.oBW = o; o += 36;
  p.oCOS = o; o += H;
  p.oSIN = o; o += H;
  p.oCS = o; o += H;
  p.oSS = o; o += H;
  p.oE = o; o += 12 * H;
  p.oAJ = o; o += 3 * H;
  p.oBJ = o; o += 3 * H;
  p.oPRM = o; o += 32;  // q[12], r[12], mu, fzmax, mass, sc, ...
  p.oRED = o; o += 32;
  p.n_doubles = o;
  int b = o * 8;
  p.oBLK = b; b += 2 * nb;          // uint8 k, uint8 leg per stance block
  p.oFACE = b; b += 4 * nb;         // int8 sx, sy, top, apex per block
  b = (b + 1) & ~1;
  p.oBMAP = b; b += 2 * 4 * H;      // int16 stance-block index per (step, leg), -1 for swing
  b = (b + 15) & ~15;
  p.oINT = b; b += 16 * 4;          // scratch ints
  p.oACT = b; b += 32 * 4;          // active-set words under construction
  p.total_bytes = b;
  return p;
}

enum { PRM_Q = 0, PRM_R = 12, PRM_MU = 24, PRM_FZ = 25, PRM_MASS = 26, PRM_SC = 27, PRM_VW = 28 };
enum { INT_NB = 0, INT_CHANGED = 1, INT_WORK = 2, INT_FAIL = 3 };

// ---------------------------------------------------------------- the fused solve
template <int NT>
__device__ void solve_one(const KParams& P, const KBatch& Bt, KStats* stats, const SmemPlan& pl,
                          unsigned char* smem, double* gH, double* gM, int pi) {
  const int tid = threadIdx.x;
  const int H = P.H;
  double* sd = reinterpret_cast<double*>(smem);
  double* Hs = pl.mats_in_smem ? sd + pl.oH : gH;
  double* M = pl.mats_in_smem ? sd + pl.oM : gM;
  double* u = sd + pl.oU;
  double* gs = sd + pl.oGS;
  double* rd = sd + pl.oRD;
  double* dua = sd + pl.oDUA;
  double* du = sd + pl.oDU;
  double* tmp = sd + pl.oTMP;
  double* rinv = sd + pl.oRINV;
  double* pv = sd + pl.oPV;
  double* z = sd + pl.oZ;
  double* Bw = sd + pl.oBW;   // 3 x 12 column-major: Bw[s + 3*c] = (I_w^-1 [r_leg]x)(s, c)
  double* cosv = sd + pl.oCOS;
  double* sinv = sd + pl.oSIN;
  double* Cs = sd + pl.oCS;   // prefix sums of cos / sin
  double* Ss = sd + pl.oSS;
  double* e = sd + pl.oE;     // free response minus reference, [H][12]
  double* aj = sd + pl.oAJ;
  double* bj = sd + pl.oBJ;
  double* prm = sd + pl.oPRM;
  double* red = sd + pl.oRED;
  unsigned char* blk = smem + pl.oBLK;
  signed char* face = reinterpret_cast<signed char*>(smem + pl.oFACE);
  int* si = reinterpret_cast<int*>(smem + pl.oINT);
  short* bmap = reinterpret_cast<short*>(smem + pl.oBMAP);
  unsigned int* aw = reinterpret_cast<unsigned int*>(smem + pl.oACT);

  const double dt = P.dt, dt2 = P.dt * P.dt, theta = P.zoh ? 0.5 : 0.0;
  const size_t p64 = (size_t)pi;
  const double* x0 = Bt.x0 + 12 * p64;
  const double* rot = Bt.rot + 9 * p64;
  const double* foot = Bt.foot + 12 * p64;
  const double* cmd = Bt.cmd + 9 * p64;
  const uint8_t* contact = Bt.contact + (size_t)(4 * H) * p64;

  // ---------------- stage 1: parameters, SRBD build, yaw references, stance-block list
  for (int t = tid; t < 12; t += NT) {
    prm[PRM_Q + t] = Bt.q ? Bt.q[12 * p64 + t] : P.q[t];
    prm[PRM_R + t] = Bt.r ? Bt.r[12 * p64 + t] : P.r[t];
  }
  if (tid == 0) {
    prm[PRM_MU] = Bt.mu ? Bt.mu[p64] : P.mu;
    prm[PRM_FZ] = Bt.fzmax ? Bt.fzmax[p64] : P.fz_max;
    prm[PRM_MASS] = Bt.mass ? Bt.mass[p64] : P.mass;
    // a1_trunk_inertia_world = R I R^T and its inverse  (ConvexQPSolver.cpp:201-205)
    double R[9], I[9], RI[9], Iw[9];
    for (int k = 0; k < 9; ++k) { R[k] = rot[k]; I[k] = Bt.inertia ? Bt.inertia[9 * p64 + k] : P.inertia[k]; }
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)
        RI[r + 3 * c] = R[r] * I[3 * c] + R[r + 3] * I[1 + 3 * c] + R[r + 6] * I[2 + 3 * c];
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)  // (RI R^T)(r,c) = sum_k RI(r,k) R(c,k)
        Iw[r + 3 * c] = RI[r] * R[c] + RI[r + 3] * R[c + 3] + RI[r + 6] * R[c + 6];
    const double a = Iw[0], b = Iw[3], c = Iw[6], d = Iw[1], ee = Iw[4], f = Iw[7], g = Iw[2],
                 h = Iw[5], i = Iw[8];
    const double c00 = ee * i - f * h, c01 = -(d * i - f * g), c02 = d * h - ee * g;
    const double c10 = -(b * i - c * h), c11 = a * i - c * g, c12 = -(a * h - b * g);
    const double c20 = b * f - c * ee, c21 = -(a * f - c * d), c22 = a * ee - b * d;
    const double id = 1.0 / (a * c00 + b * c01 + c * c02);
    double Ii[9];
    Ii[0] = c00 * id; Ii[3] = c10 * id; Ii[6] = c20 * id;
    Ii[1] = c01 * id; Ii[4] = c11 * id; Ii[7] = c21 * id;
    Ii[2] = c02 * id; Ii[5] = c12 * id; Ii[8] = c22 * id;
    for (int leg = 0; leg < 4; ++leg) {
      const double rx = foot[3 * leg], ry = foot[3 * leg + 1], rz = foot[3 * leg + 2];
      // skew(r) columns (Utils.cpp:89-95): (0, rz, -ry), (-rz, 0, rx), (ry, -rx, 0)
      for (int s = 0; s < 3; ++s) {
        Bw[s + 3 * (3 * leg + 0)] = Ii[s + 3] * rz - Ii[s + 6] * ry;
        Bw[s + 3 * (3 * leg + 1)] = -Ii[s] * rz + Ii[s + 6] * rx;
        Bw[s + 3 * (3 * leg + 2)] = Ii[s] * ry - Ii[s + 3] * rx;
      }
    }
    // root_lin_vel_d_world = R v_d_rel  (:260)
    for (int r = 0; r < 3; ++r) prm[PRM_VW + r] = R[r] * cmd[3] + R[r + 3] * cmd[4] + R[r + 6] * cmd[5];
    // stance-block list in (step, leg) order
    int nb = 0;
    for (int k = 0; k < H; ++k)
      for (int leg = 0; leg < 4; ++leg)
        if (contact[4 * k + leg]) { blk[2 * nb] = (unsigned char)k; blk[2 * nb + 1] = (unsigned char)leg; bmap[4 * k + leg] = (short)nb; ++nb; }
        else bmap[4 * k + leg] = -1;
    si[INT_NB] = nb;
    si[INT_FAIL] = 0;
  }
  for (int k = tid; k < H; k += NT) {
    // yaw of x_ref(k) = yaw + yaw_rate_d * dt * (k + offset)  (:267) ; Ad(k) linearised there (:279)
    const double yawk = x0[2] + cmd[8] * dt * (double)(k + P.xoff);
    double sv, cv;
    sincos(yawk, &sv, &cv);
    cosv[k] = cv; sinv[k] = sv;
  }
  __syncthreads();
  const int nb = si[INT_NB];
  const int n = 3 * nb;
  const double mu = prm[PRM_MU], fzm = prm[PRM_FZ], mass = prm[PRM_MASS];
  const double* qw = prm + PRM_Q;
  const double* rw = prm + PRM_R;

  // free response  x^f_{k+1} = Ad_k x^f_k + cd_k  and  e_k = x^f_{k+1} - x_ref(k)  (serial, tiny)
  if (tid == 0) {
    double th0 = x0[0], th1 = x0[1], th2 = x0[2], p0 = x0[3], p1 = x0[4], p2 = x0[5];
    const double w0 = x0[6], w1 = x0[7], w2 = x0[8], v0 = x0[9], v1 = x0[10];
    double v2 = x0[11];
    double ac = 0, as = 0;
    for (int k = 0; k < H; ++k) {
      const double c = cosv[k], s = sinv[k];
      ac += c; as += s; Cs[k] = ac; Ss[k] = as;
      th0 += dt * (c * w0 + s * w1);
      th1 += dt * (-s * w0 + c * w1);
      th2 += dt * w2;
      p0 += dt * v0; p1 += dt * v1;
      p2 += dt * v2 - theta * dt2 * P.gravity;
      v2 -= P.gravity * dt;
      const double t = dt * (double)(k + P.xoff);
      double* ek = e + 12 * k;  // x_ref(k) per ConvexQPSolver.cpp:264-276
      ek[0] = th0 - cmd[0]; ek[1] = th1 - cmd[1]; ek[2] = th2 - (x0[2] + cmd[8] * t);
      ek[3] = p0 - (x0[3] + prm[PRM_VW] * t); ek[4] = p1 - (x0[4] + prm[PRM_VW + 1] * t); ek[5] = p2 - cmd[2];
      ek[6] = w0 - cmd[6]; ek[7] = w1 - cmd[7]; ek[8] = w2 - cmd[8];
      ek[9] = v0 - prm[PRM_VW]; ek[10] = v1 - prm[PRM_VW + 1]; ek[11] = v2 - 0.0;
    }
  }
  __syncthreads();

  // ---------------- stage 3a: per-step gradient pieces  a_j (angular, 3) and b_j (linear, 3)
  // B_qp[k][j] = [E_kj Bw ; gam_kj Bv ; dt Bw ; dt Bv],  E_kj = dt^2 (sum_{i=j+1..k} T_i + theta T_j),
  // gam_kj = dt^2 (k - j + theta).  g_j = Bw' a_j + b_j / m.
  for (int j = tid; j < H; j += NT) {
    double a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0, b2 = 0, sw0 = 0, sw1 = 0, sw2 = 0, sv0 = 0, sv1 = 0, sv2 = 0;
    for (int k = j; k < H; ++k) {
      const double ea = dt2 * (Cs[k] - Cs[j] + theta * cosv[j]);
      const double eb = dt2 * (Ss[k] - Ss[j] + theta * sinv[j]);
      const double gam = dt2 * ((double)(k - j) + theta);
      const double* ek = e + 12 * k;
      const double w0 = qw[0] * ek[0], w1 = qw[1] * ek[1], w2 = qw[2] * ek[2];
      a0 += ea * w0 - eb * w1;   // E' w, E = [[ea, eb, 0], [-eb, ea, 0], [0, 0, gam]]
      a1 += eb * w0 + ea * w1;
      a2 += gam * w2;
      b0 += gam * qw[3] * ek[3]; b1 += gam * qw[4] * ek[4]; b2 += gam * qw[5] * ek[5];
      sw0 += ek[6]; sw1 += ek[7]; sw2 += ek[8];
      sv0 += ek[9]; sv1 += ek[10]; sv2 += ek[11];
    }
    aj[3 * j + 0] = a0 + dt * qw[6] * sw0; aj[3 * j + 1] = a1 + dt * qw[7] * sw1; aj[3 * j + 2] = a2 + dt * qw[8] * sw2;
    bj[3 * j + 0] = b0 + dt * qw[9] * sv0; bj[3 * j + 1] = b1 + dt * qw[10] * sv1; bj[3 * j + 2] = b2 + dt * qw[11] * sv2;
  }
  __syncthreads();

  if (n == 0) {  // every leg swings over the whole horizon: U = 0 (SURVEY KAT K3)
    if (tid == 0) {
      if (Bt.status) Bt.status[p64] = 0;
      if (Bt.iters) Bt.iters[p64] = 0;
      atomicAdd(&stats->problems, 1ull);
    }
    for (int i = tid; i < 12; i += NT) Bt.grf[12 * p64 + i] = 0.0;
    if (Bt.U) for (int i = tid; i < 12 * H; i += NT) Bt.U[(size_t)(12 * H) * p64 + i] = 0.0;
    if (Bt.active) {
      const int W = (24 * H + 31) / 32;
      for (int w = tid; w < W; w += NT) {
        const int lo = 32 * w, hi = min(32 * w + 32, 24 * H);
        Bt.active[(size_t)W * p64 + w] = (hi - lo == 32) ? 0xffffffffu : ((1u << (hi - lo)) - 1u);
      }
    }
    __syncthreads();
    return;
  }

  // ---------------- stage 3b: reduced Hessian, one 3x3 block pair per thread-iteration
  {
    const int npairs = (nb * (nb + 1)) >> 1;
    const double im2 = 1.0 / (mass * mass);
    for (int pr = tid; pr < npairs; pr += NT) {
      // pr -> (b1 >= b2)
      int b1 = (int)((sqrt(8.0 * (double)pr + 1.0) - 1.0) * 0.5);
      while (((b1 + 1) * (b1 + 2)) / 2 <= pr) ++b1;
      while ((b1 * (b1 + 1)) / 2 > pr) --b1;
      const int b2 = pr - (b1 * (b1 + 1)) / 2;
      const int j1 = blk[2 * b1], l1 = blk[2 * b1 + 1], j2 = blk[2 * b2], l2 = blk[2 * b2 + 1];
      // N = sum_{k>=j1} E_{k j1}' Q_theta E_{k j2}  (j1 >= j2), P = sum gam gam
      double n00 = 0, n01 = 0, n10 = 0, n11 = 0, n22 = 0;
      const double o1c = theta * cosv[j1] - Cs[j1], o1s = theta * sinv[j1] - Ss[j1];
      const double o2c = theta * cosv[j2] - Cs[j2], o2s = theta * sinv[j2] - Ss[j2];
      for (int k = j1; k < H; ++k) {
        const double a1 = dt2 * (Cs[k] + o1c), bb1 = dt2 * (Ss[k] + o1s);
        const double a2 = dt2 * (Cs[k] + o2c), bb2 = dt2 * (Ss[k] + o2s);
        const double g1 = dt2 * ((double)(k - j1) + theta), g2 = dt2 * ((double)(k - j2) + theta);
        n00 += a1 * qw[0] * a2 + bb1 * qw[1] * bb2;
        n01 += a1 * qw[0] * bb2 - bb1 * qw[1] * a2;
        n10 += bb1 * qw[0] * a2 - a1 * qw[1] * bb2;
        n11 += bb1 * qw[0] * bb2 + a1 * qw[1] * a2;
        n22 += g1 * g2;
      }
      const double cnt = (double)(H - j1) * dt2;
      const double Pg = n22;             // sum gam1 gam2 (shared by the yaw row and the positions)
      const double m00 = n00 + cnt * qw[6], m11 = n11 + cnt * qw[7], m22 = Pg * qw[2] + cnt * qw[8];
      // blk(a1,a2) = Bw[:,c1]' M3 Bw[:,c2] + delta (Pg Qp + cnt Qv)/m^2 (+ R on the diagonal block)
      double t[9];  // t = M3 * Bw[:, 3*l2 + a2]  -> t[s + 3*a2]
      for (int a2 = 0; a2 < 3; ++a2) {
        const double* bc = Bw + 3 * (3 * l2 + a2);
        t[0 + 3 * a2] = m00 * bc[0] + n01 * bc[1];
        t[1 + 3 * a2] = n10 * bc[0] + m11 * bc[1];
        t[2 + 3 * a2] = m22 * bc[2];
      }
      for (int a1 = 0; a1 < 3; ++a1) {
        const double* br = Bw + 3 * (3 * l1 + a1);
        const int row = 3 * b1 + a1;
        for (int a2 = 0; a2 < 3; ++a2) {
          const int col = 3 * b2 + a2;
          if (col > row) continue;
          double v = br[0] * t[0 + 3 * a2] + br[1] * t[1 + 3 * a2] + br[2] * t[2 + 3 * a2];
          if (a1 == a2) {
            v += (Pg * qw[3 + a1] + cnt * qw[9 + a1]) * im2;
            if (b1 == b2) v += rw[3 * l1 + a1];
          }
          Hs[LIDX(row, col)] = P.cs * v;
        }
      }
    }
    // gradient
    for (int b = tid; b < nb; b += NT) {
      const int j = blk[2 * b], l = blk[2 * b + 1];
      for (int a = 0; a < 3; ++a) {
        const double* bc = Bw + 3 * (3 * l + a);
        gs[3 * b + a] = P.cs * (bc[0] * aj[3 * j] + bc[1] * aj[3 * j + 1] + bc[2] * aj[3 * j + 2] + bj[3 * j + a] / mass);
      }
    }
  }
  __syncthreads();

  // ---------------- cost normalisation: Hs <- sc*H, gs <- sc*g with sc = 1/max diag(H)
  double sc;
  {
    double dmax = 0;
    for (int i = tid; i < n; i += NT) dmax = fmax(dmax, fabs(Hs[LIDX(i, i)]));
    dmax = block_reduce_max<NT>(dmax, red);
    sc = 1.0 / dmax;
    const int ps = packed_size(n);
    for (int i = tid; i < ps; i += NT) Hs[i] *= sc;
    for (int i = tid; i < n; i += NT) gs[i] *= sc;
  }
  __syncthreads();
  double gsinf = 0;
  for (int i = tid; i < n; i += NT) gsinf = fmax(gsinf, fabs(gs[i]));
  gsinf = block_reduce_max<NT>(gsinf, red);
  const double rd_tol = 1e-9 * (gsinf > 1.0 ? gsinf : 1.0);
  const double dual_tol = P.dual_rel * (gsinf > sc ? gsinf : sc);
  const double feas_tol = 1e-9;
  const int m = 5 * nb;

  // ---------------- stage 4a: interior point (feasible start)
  for (int b = tid; b < nb; b += NT) { u[3 * b] = 0.0; u[3 * b + 1] = 0.0; u[3 * b + 2] = 0.5 * fzm; }
  __syncthreads();
  symv<NT>(Hs, n, u, tmp);  // tmp = Hs u
  __syncthreads();
  {
    double ginf = 0, ssum = 0;
    for (int i = tid; i < n; i += NT) { const double v = fabs(tmp[i] + gs[i]); ginf = v > ginf ? v : ginf; }
    for (int b = tid; b < nb; b += NT) {
      double s[5];
      blk_slack(u + 3 * b, mu, fzm, s);
      ssum += s[0] + s[1] + s[2] + s[3] + s[4];
    }
    ginf = block_reduce_max<NT>(ginf, red);
    ssum = block_reduce_sum<NT>(ssum, red);
    const double zc = (ginf > 1e-6 ? ginf : 1e-6) * (ssum / (double)m);  // z_i s_i = const
    for (int b = tid; b < nb; b += NT) {
      double s[5];
      blk_slack(u + 3 * b, mu, fzm, s);
      for (int r = 0; r < 5; ++r) z[5 * b + r] = zc / s[r];
    }
  }
  __syncthreads();

  int n_ipm = 0, n_polish = 0, status = -1;
  const int max_iter = P.max_iter;
  bool solved = false;
  // three accuracy stages: polish after mu <= 1e-7; if the face guess does not verify, continue
  // to 1e-10 and retry; finally accept the interior-point answer at 1e-13.
  // warm start: see solve_one_warp
  uint8_t* wrec = Bt.warm ? Bt.warm + (size_t)warm_stride(H) * p64 : nullptr;
  const bool warm = wrec && (wrec[4 * H] != 0 || P.try_interior);
  for (int stage = warm ? -1 : 0; stage < 3 && !solved; ++stage) {
    const double mu_tol = stage <= 0 ? 1e-7 : (stage == 1 ? 1e-10 : 1e-13);
    bool ipm_converged = false, ipm_broken = false;
    while (stage >= 0 && n_ipm < max_iter) {
      // residuals at the current iterate
      symv<NT>(Hs, n, u, tmp);
      __syncthreads();
      double rdinf = 0, sz = 0;
      for (int b = tid; b < nb; b += NT) {
        double s[5], ct[3];
        blk_slack(u + 3 * b, mu, fzm, s);
        const double* zb = z + 5 * b;
        blk_ct(zb, mu, ct);
        // C' z with C rows (-1,0,-mu),(1,0,-mu),(0,-1,-mu),(0,1,-mu),(0,0,1):
        //   blk_ct gives [w1-w0, w3-w2, w4 - mu*sum] which is exactly C' w.
        for (int a = 0; a < 3; ++a) {
          const double v = tmp[3 * b + a] + gs[3 * b + a] + ct[a];
          rd[3 * b + a] = v;
          rdinf = fabs(v) > rdinf ? fabs(v) : rdinf;
        }
        for (int r = 0; r < 5; ++r) sz += s[r] * zb[r];
      }
      rdinf = block_reduce_max<NT>(rdinf, red);
      sz = block_reduce_sum<NT>(sz, red);
      const double mub = sz / (double)m;
      if (!(rdinf == rdinf) || !(mub == mub)) { ipm_broken = true; break; }
      if (rdinf <= rd_tol && mub <= mu_tol) { ipm_converged = true; break; }
      ++n_ipm;
      // M = Hs + C' W C
      {
        const int ps = packed_size(n);
        for (int i = tid; i < ps; i += NT) M[i] = Hs[i];
        __syncthreads();
        for (int b = tid; b < nb; b += NT) {
          double s[5];
          blk_slack(u + 3 * b, mu, fzm, s);
          const double* zb = z + 5 * b;
          const double w0 = zb[0] / s[0], w1 = zb[1] / s[1], w2 = zb[2] / s[2], w3 = zb[3] / s[3], w4 = zb[4] / s[4];
          const int r0 = 3 * b;
          M[LIDX(r0, r0)] += w0 + w1;
          M[LIDX(r0 + 1, r0 + 1)] += w2 + w3;
          M[LIDX(r0 + 2, r0 + 2)] += mu * mu * (w0 + w1 + w2 + w3) + w4;
          M[LIDX(r0 + 2, r0)] += mu * (w0 - w1);
          M[LIDX(r0 + 2, r0 + 1)] += mu * (w2 - w3);
        }
        __syncthreads();
      }
      if (!chol_factor<NT>(M, rinv, n)) { ipm_broken = true; break; }
      // predictor: du_aff = -M^-1 (Hs u + gs)
      for (int i = tid; i < n; i += NT) dua[i] = -(tmp[i] + gs[i]);
      __syncthreads();
      chol_solve<NT>(M, rinv, n, dua);
      double amax = 1.0;
      for (int b = tid; b < nb; b += NT) {
        double s[5], ds[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        blk_dslack(dua + 3 * b, mu, ds);
        const double* zb = z + 5 * b;
        for (int r = 0; r < 5; ++r) {
          const double dz = -zb[r] - (zb[r] / s[r]) * ds[r];
          if (ds[r] < 0) amax = fmin(amax, -s[r] / ds[r]);
          if (dz < 0) amax = fmin(amax, -zb[r] / dz);
        }
      }
      amax = block_reduce_min<NT>(amax, red);
      double szaff = 0;
      for (int b = tid; b < nb; b += NT) {
        double s[5], ds[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        blk_dslack(dua + 3 * b, mu, ds);
        const double* zb = z + 5 * b;
        for (int r = 0; r < 5; ++r) {
          const double dz = -zb[r] - (zb[r] / s[r]) * ds[r];
          szaff += (s[r] + amax * ds[r]) * (zb[r] + amax * dz);
        }
      }
      szaff = block_reduce_sum<NT>(szaff, red);
      const double rat = (szaff / (double)m) / mub;
      const double sigma = rat * rat * rat;
      // corrector: rhs = -rd - C' (rc / s),  rc = -s z + sigma mu - ds_aff dz_aff
      for (int b = tid; b < nb; b += NT) {
        double s[5], ds[5], w[5], ct[3];
        blk_slack(u + 3 * b, mu, fzm, s);
        blk_dslack(dua + 3 * b, mu, ds);
        const double* zb = z + 5 * b;
        for (int r = 0; r < 5; ++r) {
          const double dz = -zb[r] - (zb[r] / s[r]) * ds[r];
          w[r] = (-s[r] * zb[r] + sigma * mub - ds[r] * dz) / s[r];
        }
        blk_ct(w, mu, ct);
        for (int a = 0; a < 3; ++a) du[3 * b + a] = -rd[3 * b + a] - ct[a];
      }
      __syncthreads();
      chol_solve<NT>(M, rinv, n, du);
      double a2 = 1e300;
      for (int b = tid; b < nb; b += NT) {
        double s[5], dsa[5], ds[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        blk_dslack(dua + 3 * b, mu, dsa);
        blk_dslack(du + 3 * b, mu, ds);
        const double* zb = z + 5 * b;
        for (int r = 0; r < 5; ++r) {
          const double wr = zb[r] / s[r];
          const double dza = -zb[r] - wr * dsa[r];
          const double rc = -s[r] * zb[r] + sigma * mub - dsa[r] * dza;
          const double dz = rc / s[r] - wr * ds[r];
          if (ds[r] < 0) a2 = fmin(a2, -s[r] / ds[r]);
          if (dz < 0) a2 = fmin(a2, -zb[r] / dz);
        }
      }
      a2 = block_reduce_min<NT>(a2, red);
      const double alpha = fmin(1.0, 0.995 * a2);
      for (int b = tid; b < nb; b += NT) {
        double s[5], dsa[5], ds[5], znew[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        blk_dslack(dua + 3 * b, mu, dsa);
        blk_dslack(du + 3 * b, mu, ds);
        double* zb = z + 5 * b;
        for (int r = 0; r < 5; ++r) {
          const double wr = zb[r] / s[r];
          const double dza = -zb[r] - wr * dsa[r];
          const double rc = -s[r] * zb[r] + sigma * mub - dsa[r] * dza;
          const double dz = rc / s[r] - wr * ds[r];
          znew[r] = zb[r] + alpha * dz;
        }
        for (int r = 0; r < 5; ++r) zb[r] = znew[r];
        for (int a = 0; a < 3; ++a) u[3 * b + a] += alpha * du[3 * b + a];
      }
      __syncthreads();
    }
    if (stage == 2 && !ipm_broken) {  // (a broken interior point still gets one polish: see solve_one_warp)
      status = ipm_converged ? 1 : 2;
      solved = ipm_converged;
      break;
    }

    // ---------------- stage 4b: exact active-set polish on the identified pyramid faces
    // face per block: sx, sy in {-1, 0, +1} (fx = sx mu fz / fy = sy mu fz pinned), top (fz = fmax),
    // apex (f = 0).  u_b = T_b v_b + p_b with the shear T_b = I + mu (sx e_x + sy e_y) e_z'.
    if (stage < 0) {
      for (int b = tid; b < nb; b += NT) warm_face_load(wrec, H, Bt.warm_shift, blk[2 * b], blk[2 * b + 1], face + 4 * b);
      __syncthreads();
    } else {
      double smax = 0, zmax = 0;
      for (int b = tid; b < nb; b += NT) {
        double s[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        for (int r = 0; r < 5; ++r) { smax = fmax(smax, s[r]); zmax = fmax(zmax, z[5 * b + r]); }
      }
      smax = block_reduce_max<NT>(smax, red);
      zmax = block_reduce_max<NT>(zmax, red);
      const double bal = smax / (zmax > 1e-300 ? zmax : 1e-300);
      for (int b = tid; b < nb; b += NT) {
        double s[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        bool act[5];
        for (int r = 0; r < 5; ++r) act[r] = s[r] < z[5 * b + r] * bal;
        face[4 * b + 0] = act[0] ? -1 : (act[1] ? 1 : 0);
        face[4 * b + 1] = act[2] ? -1 : (act[3] ? 1 : 0);
        face[4 * b + 2] = act[4] ? 1 : 0;
        face[4 * b + 3] = ((act[0] && act[1]) || (act[2] && act[3])) ? 1 : 0;
      }
      __syncthreads();
    }
    bool polished = false;
    for (int round = 0; round < 6; ++round) {
      ++n_polish;
      // p (pinned particular solution) and tmp = Hs p
      for (int b = tid; b < nb; b += NT) {
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double pz = (top && !apex) ? fzm : 0.0;
        pv[3 * b] = sx * mu * pz; pv[3 * b + 1] = sy * mu * pz; pv[3 * b + 2] = pz;
      }
      __syncthreads();
      symv<NT>(Hs, n, pv, tmp);
      __syncthreads();
      // Mt = T' Hs T with pinned rows/cols replaced by identity, one 3x3 block pair at a time
      {
        const int npairs = (nb * (nb + 1)) >> 1;
        for (int pr = tid; pr < npairs; pr += NT) {
          int b1 = (int)((sqrt(8.0 * (double)pr + 1.0) - 1.0) * 0.5);
          while (((b1 + 1) * (b1 + 2)) / 2 <= pr) ++b1;
          while ((b1 * (b1 + 1)) / 2 > pr) --b1;
          const int b2 = pr - (b1 * (b1 + 1)) / 2;
          double X[9];  // X[a1 + 3*a2] = Hs(3b1+a1, 3b2+a2)
          for (int a1 = 0; a1 < 3; ++a1)
            for (int a2 = 0; a2 < 3; ++a2) {
              const int r = 3 * b1 + a1, c = 3 * b2 + a2;
              X[a1 + 3 * a2] = (c <= r) ? Hs[LIDX(r, c)] : Hs[LIDX(c, r)];
            }
          const double s1x = mu * face[4 * b1], s1y = mu * face[4 * b1 + 1];
          const double s2x = mu * face[4 * b2], s2y = mu * face[4 * b2 + 1];
          // rows: z-row += s1x * x-row + s1y * y-row   (T_b1')
          for (int a2 = 0; a2 < 3; ++a2) X[2 + 3 * a2] += s1x * X[0 + 3 * a2] + s1y * X[1 + 3 * a2];
          // cols: z-col += s2x * x-col + s2y * y-col   (T_b2)
          for (int a1 = 0; a1 < 3; ++a1) X[a1 + 6] += s2x * X[a1] + s2y * X[a1 + 3];
          bool pin1[3], pin2[3];
          {
            const int apex = face[4 * b1 + 3];
            pin1[0] = apex || face[4 * b1] != 0; pin1[1] = apex || face[4 * b1 + 1] != 0; pin1[2] = apex || face[4 * b1 + 2] != 0;
            const int apex2 = face[4 * b2 + 3];
            pin2[0] = apex2 || face[4 * b2] != 0; pin2[1] = apex2 || face[4 * b2 + 1] != 0; pin2[2] = apex2 || face[4 * b2 + 2] != 0;
          }
          for (int a1 = 0; a1 < 3; ++a1)
            for (int a2 = 0; a2 < 3; ++a2) {
              const int r = 3 * b1 + a1, c = 3 * b2 + a2;
              if (c > r) continue;
              double v = X[a1 + 3 * a2];
              if (pin1[a1] || pin2[a2]) v = (r == c) ? 1.0 : 0.0;
              M[LIDX(r, c)] = v;
            }
        }
        // rhs = -T'(gs + Hs p) on free components, 0 on pinned ones
        for (int b = tid; b < nb; b += NT) {
          const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
          const double gx = gs[3 * b] + tmp[3 * b], gy = gs[3 * b + 1] + tmp[3 * b + 1], gz = gs[3 * b + 2] + tmp[3 * b + 2];
          const double gzt = gz + mu * (sx * gx + sy * gy);
          du[3 * b] = (apex || sx != 0) ? 0.0 : -gx;
          du[3 * b + 1] = (apex || sy != 0) ? 0.0 : -gy;
          du[3 * b + 2] = (apex || top) ? 0.0 : -gzt;
        }
      }
      __syncthreads();
      if (!chol_factor<NT>(M, rinv, n)) { break; }
      chol_solve<NT>(M, rinv, n, du);
      // u = T v + p
      for (int b = tid; b < nb; b += NT) {
        const int sx = face[4 * b], sy = face[4 * b + 1];
        const double vx = du[3 * b], vy = du[3 * b + 1], vz = du[3 * b + 2];
        dua[3 * b] = vx + mu * sx * vz + pv[3 * b];
        dua[3 * b + 1] = vy + mu * sy * vz + pv[3 * b + 1];
        dua[3 * b + 2] = vz + pv[3 * b + 2];
      }
      __syncthreads();
      symv<NT>(Hs, n, dua, tmp);  // gradient = Hs u + gs
      __syncthreads();
      if (tid == 0) si[INT_CHANGED] = 0;
      __syncthreads();
      int changed = 0;
      for (int b = tid; b < nb; b += NT) {
        int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double gx = tmp[3 * b] + gs[3 * b], gy = tmp[3 * b + 1] + gs[3 * b + 1], gz = tmp[3 * b + 2] + gs[3 * b + 2];
        const double fx = dua[3 * b], fy = dua[3 * b + 1], fz = dua[3 * b + 2];
        if (apex) {
          // -grad must lie in the normal cone of the pyramid at its apex
          if (gz < mu * (fabs(gx) + fabs(gy)) - dual_tol) {
            apex = 0; top = 0;
            sx = gx > 0 ? -1 : (gx < 0 ? 1 : 0);
            sy = gy > 0 ? -1 : (gy < 0 ? 1 : 0);
            changed = 1;
          }
        } else {
          const double gvz = gz + mu * (sx * gx + sy * gy);
          // multipliers of pinned rows (must be >= 0), computed before any change
          const bool dropx = sx != 0 && ((sx == -1 ? gx : -gx) < -dual_tol);
          const bool dropy = sy != 0 && ((sy == -1 ? gy : -gy) < -dual_tol);
          const bool dropt = top && (-gvz < -dual_tol);
          int nsx = dropx ? 0 : sx, nsy = dropy ? 0 : sy, ntop = dropt ? 0 : top, napex = 0;
          if (fz > fzm + feas_tol) ntop = 1;
          if (fz < -feas_tol) napex = 1;
          else {
            if (sx == 0) { if (-fx - mu * fz > feas_tol) nsx = -1; else if (fx - mu * fz > feas_tol) nsx = 1; }
            if (sy == 0) { if (-fy - mu * fz > feas_tol) nsy = -1; else if (fy - mu * fz > feas_tol) nsy = 1; }
          }
          if (nsx != sx || nsy != sy || ntop != top || napex != apex) changed = 1;
          sx = nsx; sy = nsy; top = ntop; apex = napex;
        }
        face[4 * b] = (signed char)sx; face[4 * b + 1] = (signed char)sy; face[4 * b + 2] = (signed char)top; face[4 * b + 3] = (signed char)apex;
      }
      if (changed) atomicOr(&si[INT_CHANGED], 1);
      {  // NaN compares false everywhere above: a non-finite solve must never pass as verified
        double fsum = 0.0;
        for (int i = tid; i < n; i += NT) fsum += dua[i] + tmp[i] + gs[i];
        if (!isfinite(fsum)) atomicOr(&si[INT_CHANGED], 2);
      }
      __syncthreads();
      if (si[INT_CHANGED] & 2) break;
      if (si[INT_CHANGED] == 0) { polished = true; break; }
      __syncthreads();
    }
    if (polished) {
      for (int i = tid; i < n; i += NT) u[i] = dua[i];
      __syncthreads();
      status = 0;
      solved = true;
    } else if (ipm_broken) {
      status = 3;
      break;
    }
    // else: keep the interior iterate (u, z untouched by the polish) and tighten mu
  }
  if (status < 0) status = 2;

  // ---------------- stage 5: outputs
  const bool ok = status < 2;
  {
    double* Uout = Bt.U ? Bt.U + (size_t)(12 * H) * p64 : nullptr;
    if (Uout) {
      for (int i = tid; i < 12 * H; i += NT) Uout[i] = 0.0;
      __syncthreads();
    }
    for (int i = tid; i < 12; i += NT) Bt.grf[12 * p64 + i] = 0.0;
    __syncthreads();
    for (int b = tid; b < nb; b += NT) {
      const int k = blk[2 * b], l = blk[2 * b + 1];
      for (int a = 0; a < 3; ++a) {
        const double v = ok ? u[3 * b + a] : 0.0;  // failure -> zeros (ConvexQPSolver.cpp:321-326)
        if (Uout) Uout[12 * k + 3 * l + a] = v;
        if (k == 0) Bt.grf[12 * p64 + 3 * l + a] = v;
      }
    }
    if (Bt.active) {
      const int W = (24 * H + 31) / 32;
      for (int w = tid; w < W; w += NT) aw[w] = 0u;
      __syncthreads();
      for (int idx = tid; idx < 4 * H; idx += NT) {
        const int k = idx >> 2, l = idx & 3;
        double f0 = 0, f1 = 0, f2 = 0, ub = 0;
        const int b = bmap[idx];
        if (b >= 0) {
          ub = fzm;
          if (ok) { f0 = u[3 * b]; f1 = u[3 * b + 1]; f2 = u[3 * b + 2]; }
        }
        const double res[6] = {f0 + mu * f2, -(f0 - mu * f2), f1 + mu * f2, -(f1 - mu * f2), f2, ub - f2};
        const int bits[6] = {16 * k + 4 * l + 0, 16 * k + 4 * l + 1, 16 * k + 4 * l + 2, 16 * k + 4 * l + 3,
                             16 * H + 4 * k + l, 20 * H + 4 * k + l};
        for (int t = 0; t < 6; ++t)
          if (fabs(res[t]) <= P.act_tol) atomicOr(&aw[bits[t] >> 5], 1u << (bits[t] & 31));
      }
      __syncthreads();
      for (int w = tid; w < W; w += NT) Bt.active[(size_t)W * p64 + w] = aw[w];
    }
    if (wrec) {
      __syncthreads();  // every thread has read the record-valid byte
      for (int idx = tid; idx < 4 * H; idx += NT) {
        const int b = bmap[idx];
        wrec[idx] = (status == 0 && b >= 0) ? face_encode(face + 4 * b) : (unsigned char)0;
      }
      if (tid == 0) wrec[4 * H] = status == 0 ? 1 : 0;
    }
    if (tid == 0) {
      if (status == 0 && n_ipm == 0) atomicAdd(&stats->warm_hits, 1ull);
      const int nf = n_ipm + n_polish;
      if (Bt.status) Bt.status[p64] = status;
      if (Bt.iters) Bt.iters[p64] = (nf & 0xffff) | (n_polish << 16);
      atomicAdd(&stats->problems, 1ull);
      atomicAdd(&stats->sum_n, (unsigned long long)n);
      atomicAdd(&stats->sum_fact, (unsigned long long)nf);
      atomicAdd(&stats->sum_polish, (unsigned long long)n_polish);
      if (status >= 2) atomicAdd(&stats->not_ok, 1ull);
      if (status == 1) atomicAdd(&stats->ipm_only, 1ull);
      const double dn = (double)n;
      // executed flops: factorization n^3/3, two triangular solve pairs 2*2n^2 per IPM iteration
      // (one pair per polish round), symv 2n^2 each.
      atomicAdd(&stats->flops, (double)nf * (dn * dn * dn / 3.0) + (double)n_ipm * (4.0 * dn * dn + 2.0 * dn * dn) +
                                   (double)n_polish * (2.0 * dn * dn + 4.0 * dn * dn));
    }
  }
  __syncthreads();
}

}  // namespace lmpc
