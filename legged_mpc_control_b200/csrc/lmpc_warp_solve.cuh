// lmpc_warp_solve.cuh — one-warp-per-QP driver (n <= 64): SRBD build (ConvexQPSolver.cpp:198-228,
// 254-283) -> discretise (Euler as the reference, or exact ZOH) -> structured condense with swing
// legs eliminated -> feasible-start Mehrotra interior point -> exact active-set polish on the
// identified pyramid faces, KKT-verified -> outputs (u_0, status, iterations, active bitmask), all
// on a single warp with the FP64 tensor-core dense kernels of lmpc_wmma.cuh.
#pragma once
#include "lmpc_blocks.cuh"
#include "lmpc_wmma.cuh"

namespace lmpc {

// shared-memory plan of one warp (offsets in doubles, then bytes).  The setup scratch (SRBD
// build, free response, gradient pieces) is dead before the factor is first written and aliases it.
struct WarpPlan {
  int oLS, oU, oGS, oZ, oPRM, oHU, oRD, oDUA, oDU, oWROW, oXB;  // doubles
  int oBW, oCOS, oSIN, oCS, oSS, oE, oAJ, oBJ;             // setup, inside LS
  int n_doubles;
  int oBLK, oFACE, oBMAP, total_bytes;
};
__host__ __device__ inline WarpPlan make_warp_plan(int H) {
  WarpPlan p;
  int s = 0;
  p.oLS = 0;
  p.oBW = s; s += 36;
  p.oCOS = s; s += H; p.oSIN = s; s += H; p.oCS = s; s += H; p.oSS = s; s += H;
  p.oE = s; s += 12 * H; p.oAJ = s; s += 3 * H; p.oBJ = s; s += 3 * H;
  int o = s > WLS ? s : WLS;
  p.oU = o; o += WN;
  p.oGS = o; o += WN;
  p.oHU = o; o += WN;
  p.oRD = o; o += WN;
  p.oDUA = o; o += WN;
  p.oDU = o; o += WN;
  p.oWROW = o; o += 16;  // active-set words under construction (<= 24 x 32 bits)
  p.oZ = o;              // (multipliers live in registers)
  p.oPRM = o; o += 32;
#ifdef LMPC_DIAG8_SMEM
  p.oXB = o; o += 64;    // pivot-row exchange of the shared-memory diagonal-tile variant
#else
  p.oXB = o;             // wm_diag8_mma needs no exchange buffer
#endif
  p.n_doubles = o;
  int b = o * 8;
  p.oBLK = b; b += 2 * 24;
  p.oFACE = b; b += 4 * 24;
  p.oBMAP = b; b += 2 * 4 * H;
  p.total_bytes = (b + 15) & ~15;
  return p;
}

struct WarpStats {
  unsigned long long problems = 0, sum_n = 0, sum_fact = 0, sum_polish = 0, not_ok = 0, ipm_only = 0, warm_hits = 0, handed = 0;
  double flops = 0;
};

__device__ inline void flush_warp_stats(const WarpStats& ws, KStats* stats) {
  if ((threadIdx.x & 31) == 0 && (ws.problems | ws.handed | ws.sum_fact)) {  // (a warp may have handed over everything it touched)
    atomicAdd(&stats->problems, ws.problems);
    atomicAdd(&stats->sum_n, ws.sum_n);
    atomicAdd(&stats->sum_fact, ws.sum_fact);
    atomicAdd(&stats->sum_polish, ws.sum_polish);
    if (ws.not_ok) atomicAdd(&stats->not_ok, ws.not_ok);
    if (ws.ipm_only) atomicAdd(&stats->ipm_only, ws.ipm_only);
    if (ws.warm_hits) atomicAdd(&stats->warm_hits, ws.warm_hits);
    if (ws.handed) atomicAdd(&stats->handed, ws.handed);
    atomicAdd(&stats->flops, ws.flops);
  }
}

// 3x3 block (b1 >= b2 in block order, hence j1 >= j2) of cs*(B_qp' Qbar B_qp + Rbar), unscaled
// by the cost normalisation.  out[a1 + 3*a2].
static __device__ __noinline__ void wblock(int H, double dt2, double theta, double cs, double im2,
                                       const double* __restrict__ qw, const double* __restrict__ rw,
                                       const double* __restrict__ Bw, const double* __restrict__ cosv,
                                       const double* __restrict__ sinv, const double* __restrict__ Cs,
                                       const double* __restrict__ Ss, int j1, int l1, int j2, int l2,
                                       bool same, double* out) {
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
  const double m00 = n00 + cnt * qw[6], m11 = n11 + cnt * qw[7], m22 = n22 * qw[2] + cnt * qw[8];
  double t[9];
#pragma unroll
  for (int a2 = 0; a2 < 3; ++a2) {
    const double* bc = Bw + 3 * (3 * l2 + a2);
    t[0 + 3 * a2] = m00 * bc[0] + n01 * bc[1];
    t[1 + 3 * a2] = n10 * bc[0] + m11 * bc[1];
    t[2 + 3 * a2] = m22 * bc[2];
  }
#pragma unroll
  for (int a1 = 0; a1 < 3; ++a1) {
    const double* br = Bw + 3 * (3 * l1 + a1);
#pragma unroll
    for (int a2 = 0; a2 < 3; ++a2) {
      double v = br[0] * t[0 + 3 * a2] + br[1] * t[1 + 3 * a2] + br[2] * t[2 + 3 * a2];
      if (a1 == a2) {
        v += (n22 * qw[3 + a1] + cnt * qw[9 + a1]) * im2;
        if (same) v += rw[3 * l1 + a1];
      }
      out[a1 + 3 * a2] = cs * v;
    }
  }
}

__device__ __forceinline__ void pair_from_index(int pr, int& b1, int& b2) {
  b1 = (int)((sqrtf(8.0f * (float)pr + 1.0f) - 1.0f) * 0.5f);  // float guess, fixed up exactly below (no FP64 sqrt slow path)
  if (b1 < 0) b1 = 0;
  while (((b1 + 1) * (b1 + 2)) / 2 <= pr) ++b1;
  while ((b1 * (b1 + 1)) / 2 > pr) --b1;
  b2 = pr - (b1 * (b1 + 1)) / 2;
}

// x0_s / foot_s: optional copies of this problem's x0[12] / foot_pos_abs[12] records already staged in
// shared memory (TMA bulk prefetch, lmpc_k_warp.cu); nullptr -> read them from global memory.
static __device__ void solve_one_warp(const KParams& P, const KBatch& Bt, WarpStats& ws, const WarpPlan& pl,
                               unsigned char* smem, double* Hg, int pi, const double* x0_s = nullptr,
                               const double* foot_s = nullptr) {
  const int lane = threadIdx.x & 31;
  const int H = P.H;
  double* sd = reinterpret_cast<double*>(smem);
  double* Ls = sd + pl.oLS;
  double* u = sd + pl.oU;
  double* gs = sd + pl.oGS;
  double* prm = sd + pl.oPRM;
  double* hu = sd + pl.oHU;
  double* rd = sd + pl.oRD;
  double* dua = sd + pl.oDUA;
  double* du = sd + pl.oDU;
  double* wrow = sd + pl.oWROW;
  double* xb = sd + pl.oXB;
  double* Bw = sd + pl.oBW;
  double* cosv = sd + pl.oCOS;
  double* sinv = sd + pl.oSIN;
  double* Cs = sd + pl.oCS;
  double* Ss = sd + pl.oSS;
  double* e = sd + pl.oE;
  double* aj = sd + pl.oAJ;
  double* bj = sd + pl.oBJ;
  unsigned char* blk = smem + pl.oBLK;
  signed char* face = reinterpret_cast<signed char*>(smem + pl.oFACE);
  short* bmap = reinterpret_cast<short*>(smem + pl.oBMAP);

  const double dt = P.dt, dt2 = P.dt * P.dt, theta = P.zoh ? 0.5 : 0.0;
  const size_t p64 = (size_t)pi;
  const double* x0 = x0_s ? x0_s : Bt.x0 + 12 * p64;
  const double* rot = Bt.rot + 9 * p64;
  const double* foot = foot_s ? foot_s : Bt.foot + 12 * p64;
  const double* cmd = Bt.cmd + 9 * p64;
  const uint8_t* contact = Bt.contact + (size_t)(4 * H) * p64;

  __syncwarp();
  // ---------------- stage 1: parameters, SRBD build, yaw references, stance-block list
  if (lane < 12) {
    prm[PRM_Q + lane] = Bt.q ? Bt.q[12 * p64 + lane] : P.q[lane];
    prm[PRM_R + lane] = Bt.r ? Bt.r[12 * p64 + lane] : P.r[lane];
  }
  if (lane == 0) {
    prm[PRM_MU] = Bt.mu ? Bt.mu[p64] : P.mu;
    prm[PRM_FZ] = Bt.fzmax ? Bt.fzmax[p64] : P.fz_max;
    prm[PRM_MASS] = Bt.mass ? Bt.mass[p64] : P.mass;
    double R[9], I[9], RI[9], Iw[9];
    for (int k = 0; k < 9; ++k) { R[k] = rot[k]; I[k] = Bt.inertia ? Bt.inertia[9 * p64 + k] : P.inertia[k]; }
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)
        RI[r + 3 * c] = R[r] * I[3 * c] + R[r + 3] * I[1 + 3 * c] + R[r + 6] * I[2 + 3 * c];
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r)
        Iw[r + 3 * c] = RI[r] * R[c] + RI[r + 3] * R[c + 3] + RI[r + 6] * R[c + 6];
    const double a = Iw[0], b = Iw[3], c = Iw[6], d = Iw[1], ee = Iw[4], f = Iw[7], g = Iw[2], h = Iw[5], i = Iw[8];
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
      for (int s = 0; s < 3; ++s) {
        Bw[s + 3 * (3 * leg + 0)] = Ii[s + 3] * rz - Ii[s + 6] * ry;
        Bw[s + 3 * (3 * leg + 1)] = -Ii[s] * rz + Ii[s + 6] * rx;
        Bw[s + 3 * (3 * leg + 2)] = Ii[s] * ry - Ii[s + 3] * rx;
      }
    }
    for (int r = 0; r < 3; ++r) prm[PRM_VW + r] = R[r] * cmd[3] + R[r + 3] * cmd[4] + R[r + 6] * cmd[5];
  }
  int nb = 0;
  for (int base = 0; base < 4 * H; base += 32) {  // stance-block list by ballot + prefix popcount
    const int idx = base + lane;
    const bool fl = idx < 4 * H && contact[idx] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, fl);
    const int pos = nb + __popc(m & ((1u << lane) - 1u));
    if (idx < 4 * H) bmap[idx] = fl ? (short)pos : (short)-1;
    if (fl && pos < WNBMAX) { blk[2 * pos] = (unsigned char)(idx >> 2); blk[2 * pos + 1] = (unsigned char)(idx & 3); }
    nb += __popc(m);
  }
  for (int k = lane; k < H; k += 32) {
    const double yawk = x0[2] + cmd[8] * dt * (double)(k + P.xoff);
    double sv, cv;
    sincos(yawk, &sv, &cv);
    cosv[k] = cv; sinv[k] = sv;
  }
  __syncwarp();
  const int n = 3 * nb;
  const int np = WN;  // padded dimension (identity beyond n)
  const double mu = prm[PRM_MU], fzm = prm[PRM_FZ], mass = prm[PRM_MASS];
  const double* qw = prm + PRM_Q;
  const double* rw = prm + PRM_R;

  if (lane == 0) {  // free response and prefix sums (serial, H steps)
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
      double* ek = e + 12 * k;
      ek[0] = th0 - cmd[0]; ek[1] = th1 - cmd[1]; ek[2] = th2 - (x0[2] + cmd[8] * t);
      ek[3] = p0 - (x0[3] + prm[PRM_VW] * t); ek[4] = p1 - (x0[4] + prm[PRM_VW + 1] * t); ek[5] = p2 - cmd[2];
      ek[6] = w0 - cmd[6]; ek[7] = w1 - cmd[7]; ek[8] = w2 - cmd[8];
      ek[9] = v0 - prm[PRM_VW]; ek[10] = v1 - prm[PRM_VW + 1]; ek[11] = v2;
    }
  }
  __syncwarp();
  for (int j = lane; j < H; j += 32) {
    double a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0, b2 = 0, sw0 = 0, sw1 = 0, sw2 = 0, sv0 = 0, sv1 = 0, sv2 = 0;
    for (int k = j; k < H; ++k) {
      const double ea = dt2 * (Cs[k] - Cs[j] + theta * cosv[j]);
      const double eb = dt2 * (Ss[k] - Ss[j] + theta * sinv[j]);
      const double gam = dt2 * ((double)(k - j) + theta);
      const double* ek = e + 12 * k;
      const double w0 = qw[0] * ek[0], w1 = qw[1] * ek[1], w2 = qw[2] * ek[2];
      a0 += ea * w0 - eb * w1;
      a1 += eb * w0 + ea * w1;
      a2 += gam * w2;
      b0 += gam * qw[3] * ek[3]; b1 += gam * qw[4] * ek[4]; b2 += gam * qw[5] * ek[5];
      sw0 += ek[6]; sw1 += ek[7]; sw2 += ek[8];
      sv0 += ek[9]; sv1 += ek[10]; sv2 += ek[11];
    }
    aj[3 * j + 0] = a0 + dt * qw[6] * sw0; aj[3 * j + 1] = a1 + dt * qw[7] * sw1; aj[3 * j + 2] = a2 + dt * qw[8] * sw2;
    bj[3 * j + 0] = b0 + dt * qw[9] * sv0; bj[3 * j + 1] = b1 + dt * qw[10] * sv1; bj[3 * j + 2] = b2 + dt * qw[11] * sv2;
  }
  __syncwarp();

  const int W = (24 * H + 31) / 32;
  if (n == 0) {
    if (lane == 0) {
      if (Bt.status) Bt.status[p64] = 0;
      if (Bt.iters) Bt.iters[p64] = 0;
    }
    if (lane < 12) Bt.grf[12 * p64 + lane] = 0.0;
    if (Bt.U) for (int i = lane; i < 12 * H; i += 32) Bt.U[(size_t)(12 * H) * p64 + i] = 0.0;
    if (Bt.active)
      for (int w = lane; w < W; w += 32) {
        const int lo = 32 * w, hi = min(32 * w + 32, 24 * H);
        Bt.active[(size_t)W * p64 + w] = (hi - lo == 32) ? 0xffffffffu : ((1u << (hi - lo)) - 1u);
      }
    if (Bt.warm) {
      uint8_t* wrec = Bt.warm + (size_t)warm_stride(H) * p64;
      for (int i = lane; i < 4 * H; i += 32) wrec[i] = 0;
      if (lane == 0) wrec[4 * H] = 1;
    }
    ws.problems++;
    __syncwarp();
    return;
  }

  // ---------------- stage 3: reduced Hessian into the L2 workspace (full symmetric 64x64)
  const double im2 = 1.0 / (mass * mass);
  double sc;
  double hd[6] = {0, 0, 0, 0, 0, 0};  // own diagonal 3x3 block of the scaled Hessian: (0,0) (1,0) (2,0) (1,1) (2,1) (2,2)
  {
    double dmax = 0;
    if (lane < nb) {
      double o[9];
      const int j = blk[2 * lane], l = blk[2 * lane + 1];
      wblock(H, dt2, theta, P.cs, im2, qw, rw, Bw, cosv, sinv, Cs, Ss, j, l, j, l, true, o);
      dmax = fmax(fabs(o[0]), fmax(fabs(o[4]), fabs(o[8])));
      hd[0] = o[0]; hd[1] = o[1]; hd[2] = o[2]; hd[3] = o[4]; hd[4] = o[5]; hd[5] = o[8];
    }
    dmax = wmax(dmax);
    sc = 1.0 / dmax;
#pragma unroll
    for (int k = 0; k < 6; ++k) hd[k] *= sc;
    // identity padding for rows/cols >= n
    for (int idx = lane; idx < np * (np - n); idx += 32) {
      const int c = n + idx / np, r = idx % np;
      __stcg(Hg + hfull_off(r, c), (r == c) ? 1.0 : 0.0);
      __stcg(Hg + hfull_off(c, r), (r == c) ? 1.0 : 0.0);
    }
    const int npairs = (nb * (nb + 1)) >> 1;
    for (int pr = lane; pr < npairs; pr += 32) {
      int b1, b2;
      pair_from_index(pr, b1, b2);
      double o[9];
      wblock(H, dt2, theta, P.cs, im2, qw, rw, Bw, cosv, sinv, Cs, Ss, blk[2 * b1], blk[2 * b1 + 1], blk[2 * b2],
             blk[2 * b2 + 1], b1 == b2, o);
#pragma unroll
      for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
        for (int a2 = 0; a2 < 3; ++a2) {
          const int r = 3 * b1 + a1, c = 3 * b2 + a2;
          const double v = sc * o[a1 + 3 * a2];
          __stcg(Hg + hfull_off(r, c), v);
          __stcg(Hg + hfull_off(c, r), v);  // (diagonal blocks are symmetric, so the double write agrees)
        }
    }
    if (lane < nb) {
      const int j = blk[2 * lane], l = blk[2 * lane + 1];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double* bc = Bw + 3 * (3 * l + a);
        gs[3 * lane + a] = sc * P.cs * (bc[0] * aj[3 * j] + bc[1] * aj[3 * j + 1] + bc[2] * aj[3 * j + 2] + bj[3 * j + a] / mass);
      }
    }
    for (int i = n + lane; i < WN; i += 32) gs[i] = 0.0;
  }
  __threadfence_block();
  __syncwarp();  // Hg visible to every lane; setup scratch (aliased with hu/rd/dua/du/wrow) is dead from here on
  double gsinf = wmax(fmax(fabs(gs[lane]), fabs(gs[lane + 32])));
  const double rd_tol = 1e-9 * (gsinf > 1.0 ? gsinf : 1.0);
  const double rd_tol0 = P.rdtol0 * (gsinf > 1.0 ? gsinf : 1.0);
  const double dual_tol = P.dual_rel * (gsinf > sc ? gsinf : sc);
  const double feas_tol = 1e-9;
  const int m = 5 * nb;

  // ---------------- stage 4a: interior point (feasible start)
  // Lane b < nb owns stance block b: its force u_b, multipliers z_b, slacks, gradient g_b and
  // (H u)_b stay in registers.  The factorization reads M = H + C'WC from the L2 workspace: every
  // iteration the owners overwrite their 3x3 diagonal block with H_bb + W_b (C'WC is block
  // diagonal) and restore H_bb before the polish.  (H u)_b is updated from the Newton system,
  // H du = rhs - C'WC du, and re-computed exactly (one H*x) only when convergence is claimed.
  // Ratio form of the step rules (q = ds/s):  -ds/s = -q,  -dz_aff/z = 1 + q,  so the predictor
  // needs no division; reciprocals come from frcp (rcp.approx + 2 Newton steps).
  const bool act = lane < nb;
  const double inv_m = 1.0 / (double)m;
  double fz_init = P.fz0 * fzm;
  if (P.fz0 <= 0.0 && act) {  // weight shared by the stance legs of this block's step
    const uint8_t* ck = contact + 4 * blk[2 * lane];
    const int ns = (ck[0] != 0) + (ck[1] != 0) + (ck[2] != 0) + (ck[3] != 0);
    fz_init = fmin(0.9 * fzm, fmax(0.05 * fzm, mass * P.gravity / (double)ns));
  }
  double uf[3] = {0.0, 0.0, act ? fz_init : 0.0};
  double zr[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  double huf[3] = {0.0, 0.0, 0.0}, gsf[3] = {0.0, 0.0, 0.0};
  int moff[9];  // workspace offsets of the own diagonal block, [a1 + 3 a2]
#pragma unroll
  for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
    for (int a2 = 0; a2 < 3; ++a2) moff[a1 + 3 * a2] = act ? hfull_off(3 * lane + a1, 3 * lane + a2) : 0;
  u[lane] = 0.0; u[lane + 32] = 0.0;
  hu[lane] = 0.0; hu[lane + 32] = 0.0;
  dua[lane] = 0.0; dua[lane + 32] = 0.0;
  du[lane] = 0.0; du[lane + 32] = 0.0;
  __syncwarp();
  if (act) u[3 * lane + 2] = uf[2];
  __syncwarp();
  int n_ipm = 0, n_polish = 0, status = -1;
  const int max_iter = P.max_iter;
  bool solved = false;
  // Warm start: stage -1 skips the interior point and hands the faces stored by the previous solve
  // of this problem (shifted along the horizon) straight to the exact polish; if they do not verify
  // within its repair rounds the cold path runs as usual.
  uint8_t* wrec = Bt.warm ? Bt.warm + (size_t)warm_stride(H) * p64 : nullptr;
  const bool warm = wrec && (wrec[4 * H] != 0 || P.try_interior);
  {  // interior-point start
    if (!warm) wm_symv(Hg, u, hu);  // a warm start defers this product until its faces fail to verify (hu = 0 here)
    __syncwarp();
    double ssum = 0, gi = 0;
    double s[5];
    blk_slack(uf, mu, fzm, s);
    if (act) {
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        huf[a] = hu[3 * lane + a];
        gsf[a] = gs[3 * lane + a];
        gi = fmax(gi, fabs(huf[a] + gsf[a]));
      }
      ssum = s[0] + s[1] + s[2] + s[3] + s[4];
    }
    const double ginf = wmax(gi);
    ssum = wsum(ssum);
    const double zc = P.z0_scale * (ginf > 1e-6 ? ginf : 1e-6) * (ssum * inv_m);
    if (act)
#pragma unroll
      for (int r = 0; r < 5; ++r) zr[r] = zc * frcp(s[r]);
  }
#pragma unroll 1
  for (int stage = warm ? -1 : 0; stage < 3 && !solved; ++stage) {
    const double mu_tol = stage <= 0 ? P.tol0 : (stage == 1 ? 1e-10 : 1e-13);
    const double rdt = stage <= 0 ? rd_tol0 : rd_tol;
    bool ipm_converged = false, ipm_broken = false;
    if (stage == 0 && warm) {  // the stored faces did not verify: cold path after all, H u0 is needed now
      wm_symv(Hg, u, hu);
      __syncwarp();
      if (act) {
#pragma unroll
        for (int a = 0; a < 3; ++a) huf[a] = hu[3 * lane + a];
      }
    }
#pragma unroll 1
    while (stage >= 0 && n_ipm < max_iter) {
      double s[5], is[5], w[5], rdv[3] = {0.0, 0.0, 0.0}, ctz[3] = {0.0, 0.0, 0.0};
      double wb[6] = {0, 0, 0, 0, 0, 0};  // C_b' W C_b: (0,0) (2,0) (1,1) (2,1) (2,2); (1,0) = 0
      double rdinf = 0, sz = 0;
      blk_slack(uf, mu, fzm, s);
      if (act) {
        blk_ct(zr, mu, ctz);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          rdv[a] = huf[a] + gsf[a] + ctz[a];
          rdinf = fmax(rdinf, fabs(rdv[a]));
        }
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          is[r] = frcp(s[r]);
          w[r] = zr[r] * is[r];
          sz = fma(s[r], zr[r], sz);
        }
        wb[0] = w[0] + w[1]; wb[1] = mu * (w[0] - w[1]);
        wb[2] = w[2] + w[3]; wb[3] = mu * (w[2] - w[3]);
        wb[4] = mu * mu * (w[0] + w[1] + w[2] + w[3]) + w[4];
      }
      rdinf = wmax(rdinf);
      sz = wsum(sz);
      const double mub = sz * inv_m;
      if (!(rdinf == rdinf) || !(mub == mub)) { ipm_broken = true; break; }
      if (rdinf <= rdt && mub <= mu_tol) {
        // Stages 0 and 1 hand over to the exact polish, which verifies KKT on H itself: the claim needs no
        // confirmation there (one H*x saved per solve).  The last stage returns the interior-point iterate
        // as the answer, so there convergence claimed on the recursively updated H u is confirmed with an
        // exact product (subtracting W u from M u would cancel catastrophically: W ~ z/s is huge here).
        if (stage < 2) { ipm_converged = true; break; }
        if (act) {
          __stcg(Hg + moff[0], hd[0]); __stcg(Hg + moff[4], hd[3]); __stcg(Hg + moff[8], hd[5]);
          __stcg(Hg + moff[1], hd[1]); __stcg(Hg + moff[3], hd[1]);
          __stcg(Hg + moff[2], hd[2]); __stcg(Hg + moff[6], hd[2]);
          __stcg(Hg + moff[5], hd[4]); __stcg(Hg + moff[7], hd[4]);
        }
        __syncwarp();
        wm_symv(Hg, u, hu);
        __syncwarp();
        double ri = 0;
        if (act) {
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            huf[a] = hu[3 * lane + a];
            rdv[a] = huf[a] + gsf[a] + ctz[a];
            ri = fmax(ri, fabs(rdv[a]));
          }
        }
        rdinf = wmax(ri);
        if (rdinf <= rdt) { ipm_converged = true; break; }
      }
      ++n_ipm;
      // predictor: (H + C'WC) du_aff = -(H u + g)
      if (act) {
#pragma unroll
        for (int a = 0; a < 3; ++a) dua[3 * lane + a] = -(huf[a] + gsf[a]);
        // M_bb = H_bb + W_b into the workspace (both triangles; the block may straddle tiles)
        __stcg(Hg + moff[0], hd[0] + wb[0]);
        __stcg(Hg + moff[4], hd[3] + wb[2]);
        __stcg(Hg + moff[8], hd[5] + wb[4]);
        __stcg(Hg + moff[1], hd[1]); __stcg(Hg + moff[3], hd[1]);
        __stcg(Hg + moff[2], hd[2] + wb[1]); __stcg(Hg + moff[6], hd[2] + wb[1]);
        __stcg(Hg + moff[5], hd[4] + wb[3]); __stcg(Hg + moff[7], hd[4] + wb[3]);
      }
      __syncwarp();
      if (!wm_factor<true>(Hg, Ls, xb)) { ipm_broken = true; break; }
      wm_solve(Ls, dua);
      __syncwarp();
      double q[5], tau = 0.0;
      if (act) {
        double dsa[5];
        blk_dslack(dua + 3 * lane, mu, dsa);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          q[r] = dsa[r] * is[r];
          tau = fmax(tau, fmax(-q[r], 1.0 + q[r]));
        }
      }
      tau = wmax(tau);
      const double amax = tau > 1.0 ? frcp(tau) : 1.0;
      double szaff = 0;
      if (act)
#pragma unroll
        for (int r = 0; r < 5; ++r)
          szaff = fma(s[r] * zr[r], fma(amax, q[r], 1.0) * (1.0 - amax * (1.0 + q[r])), szaff);
      szaff = wsum(szaff);
      const double rat = szaff * frcp(sz);
      const double smu = (P.sigma_pow == 2 ? rat * rat : (P.sigma_pow == 4 ? rat * rat * rat * rat : rat * rat * rat)) * mub;  // sigma * mu
      // corrector: rhs = -rd - C'((sigma mu - ds_aff dz_aff - s z)/s)
      double w2[5], rhs[3] = {0.0, 0.0, 0.0};
      if (act) {
        double ct[3];
#pragma unroll
        for (int r = 0; r < 5; ++r) w2[r] = fma(zr[r], fma(q[r], 1.0 + q[r], -1.0), smu * is[r]);
        blk_ct(w2, mu, ct);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          rhs[a] = -rdv[a] - ct[a];
          du[3 * lane + a] = rhs[a];
        }
      }
      __syncwarp();
      wm_solve(Ls, du);
      __syncwarp();
      double dz[5], duf[3] = {0.0, 0.0, 0.0}, tau2 = 0.0;
      if (act) {
        double ds[5];
#pragma unroll
        for (int a = 0; a < 3; ++a) duf[a] = du[3 * lane + a];
        blk_dslack(duf, mu, ds);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          dz[r] = fma(-w[r], ds[r], w2[r]);
          tau2 = fmax(tau2, fmax(-ds[r] * is[r], -dz[r] * frcp(zr[r])));
        }
      }
      tau2 = wmax(tau2);
      const double alpha = tau2 > P.step_frac ? P.step_frac * frcp(tau2) : 1.0;
      if (act) {
#pragma unroll
        for (int r = 0; r < 5; ++r) zr[r] = fma(alpha, dz[r], zr[r]);
        // H du = rhs - W_b du  (Newton system), so H u follows without a matrix product
        const double hd0 = rhs[0] - (wb[0] * duf[0] + wb[1] * duf[2]);
        const double hd1 = rhs[1] - (wb[2] * duf[1] + wb[3] * duf[2]);
        const double hd2 = rhs[2] - (wb[1] * duf[0] + wb[3] * duf[1] + wb[4] * duf[2]);
        huf[0] = fma(alpha, hd0, huf[0]); huf[1] = fma(alpha, hd1, huf[1]); huf[2] = fma(alpha, hd2, huf[2]);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          uf[a] = fma(alpha, duf[a], uf[a]);
          u[3 * lane + a] = uf[a];
        }
      }
      __syncwarp();
    }
    // the polish works on H itself: put the own diagonal block back
    if (act) {
      __stcg(Hg + moff[0], hd[0]); __stcg(Hg + moff[4], hd[3]); __stcg(Hg + moff[8], hd[5]);
      __stcg(Hg + moff[1], hd[1]); __stcg(Hg + moff[3], hd[1]);
      __stcg(Hg + moff[2], hd[2]); __stcg(Hg + moff[6], hd[2]);
      __stcg(Hg + moff[5], hd[4]); __stcg(Hg + moff[7], hd[4]);
    }
    __syncwarp();
    // a broken interior point (non-finite iterate, non-positive pivot) still gets one exact polish
    // from the last good iterate's faces: the polish verifies KKT itself, so this can only help
    if (stage == 2 && !ipm_broken) {
      status = ipm_converged ? 1 : 2;
      solved = ipm_converged;
      break;
    }

    // ---------------- stage 4b: exact active-set polish on the identified pyramid faces
    if (stage < 0) {
      if (act) warm_face_load(wrec, H, Bt.warm_shift, blk[2 * lane], blk[2 * lane + 1], face + 4 * lane);
      __syncwarp();
    } else {
      double smax = 0, zmax = 0;
      double s[5];
      blk_slack(uf, mu, fzm, s);
      if (act)
        for (int r = 0; r < 5; ++r) { smax = fmax(smax, s[r]); zmax = fmax(zmax, zr[r]); }
      smax = wmax(smax);
      zmax = wmax(zmax);
      const double bal = smax / (zmax > 1e-300 ? zmax : 1e-300);
      if (lane < nb) {
        bool actr[5];
        for (int r = 0; r < 5; ++r) actr[r] = s[r] < zr[r] * bal;
        face[4 * lane + 0] = actr[0] ? -1 : (actr[1] ? 1 : 0);
        face[4 * lane + 1] = actr[2] ? -1 : (actr[3] ? 1 : 0);
        face[4 * lane + 2] = actr[4] ? 1 : 0;
        face[4 * lane + 3] = ((actr[0] && actr[1]) || (actr[2] && actr[3])) ? 1 : 0;
      }
      __syncwarp();
    }
    bool polished = false;
    for (int round = 0; round < 10; ++round) {
      ++n_polish;
      // p (pinned particular solution) in rd ; hu = Hs p
      rd[lane] = 0.0; rd[lane + 32] = 0.0;
      __syncwarp();
      bool pinned_top = false;
      if (lane < nb) {
        const int b = lane;
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        pinned_top = top && !apex;
        const double pz = pinned_top ? fzm : 0.0;
        rd[3 * b] = sx * mu * pz; rd[3 * b + 1] = sy * mu * pz; rd[3 * b + 2] = pz;
      }
      __syncwarp();
      if (__any_sync(0xffffffffu, pinned_top)) {
        wm_symv(Hg, rd, hu);
      } else {  // p = 0 (no force pinned at fz_max, the usual case): H p = 0 without reading H
        hu[lane] = 0.0; hu[lane + 32] = 0.0;
      }
      __syncwarp();
      // Mt = T' Hs T with pinned rows/cols replaced by identity, written into the panel storage
      {
        const int npairs = (nb * (nb + 1)) >> 1;
        for (int pr = lane; pr < npairs; pr += 32) {
          int b1, b2;
          pair_from_index(pr, b1, b2);
          double X[9];
#pragma unroll
          for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
            for (int a2 = 0; a2 < 3; ++a2) X[a1 + 3 * a2] = __ldcg(Hg + hfull_off(3 * b1 + a1, 3 * b2 + a2));
          const double s1x = mu * face[4 * b1], s1y = mu * face[4 * b1 + 1];
          const double s2x = mu * face[4 * b2], s2y = mu * face[4 * b2 + 1];
#pragma unroll
          for (int a2 = 0; a2 < 3; ++a2) X[2 + 3 * a2] += s1x * X[0 + 3 * a2] + s1y * X[1 + 3 * a2];
#pragma unroll
          for (int a1 = 0; a1 < 3; ++a1) X[a1 + 6] += s2x * X[a1] + s2y * X[a1 + 3];
          const int ap1 = face[4 * b1 + 3], ap2 = face[4 * b2 + 3];
          const bool pin1[3] = {ap1 || face[4 * b1] != 0, ap1 || face[4 * b1 + 1] != 0, ap1 || face[4 * b1 + 2] != 0};
          const bool pin2[3] = {ap2 || face[4 * b2] != 0, ap2 || face[4 * b2 + 1] != 0, ap2 || face[4 * b2 + 2] != 0};
#pragma unroll
          for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
            for (int a2 = 0; a2 < 3; ++a2) {
              const int r = 3 * b1 + a1, c = 3 * b2 + a2;
              if (c <= r) {
                double v = X[a1 + 3 * a2];
                if (pin1[a1] || pin2[a2]) v = (r == c) ? 1.0 : 0.0;
                Ls[lfac_off(r, c)] = v;
                if ((r >> 3) == (c >> 3)) Ls[lfac_off(c, r)] = v;  // diagonal tiles are kept full
              }
            }
        }
        // identity padding rows n..63
        for (int r = n; r < np; ++r)
          for (int c = lane; c <= r; c += 32) {
            Ls[lfac_off(r, c)] = (r == c) ? 1.0 : 0.0;
            if ((r >> 3) == (c >> 3)) Ls[lfac_off(c, r)] = (r == c) ? 1.0 : 0.0;
          }
      }
      __syncwarp();
      du[lane] = 0.0; du[lane + 32] = 0.0;
      __syncwarp();
      if (lane < nb) {
        const int b = lane;
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double gx = gs[3 * b] + hu[3 * b], gy = gs[3 * b + 1] + hu[3 * b + 1], gz = gs[3 * b + 2] + hu[3 * b + 2];
        const double gzt = gz + mu * (sx * gx + sy * gy);
        du[3 * b] = (apex || sx != 0) ? 0.0 : -gx;
        du[3 * b + 1] = (apex || sy != 0) ? 0.0 : -gy;
        du[3 * b + 2] = (apex || top) ? 0.0 : -gzt;
      }
      __syncwarp();
      if (!wm_factor<false>(Hg, Ls, xb)) break;
      wm_solve(Ls, du);
      dua[lane] = 0.0; dua[lane + 32] = 0.0;
      __syncwarp();
      if (lane < nb) {
        const int b = lane;
        const int sx = face[4 * b], sy = face[4 * b + 1];
        const double vx = du[3 * b], vy = du[3 * b + 1], vz = du[3 * b + 2];
        dua[3 * b] = vx + mu * sx * vz + rd[3 * b];
        dua[3 * b + 1] = vy + mu * sy * vz + rd[3 * b + 1];
        dua[3 * b + 2] = vz + rd[3 * b + 2];
      }
      __syncwarp();
      wm_symv(Hg, dua, hu);
      __syncwarp();
      // Face repair (face_step, lmpc_blocks.cuh): rounds 0-2 apply every indicated change at once; from round 3
      // on, adds while any block is primal infeasible, else the single most violated multiplier is dropped.
      int changed = 0;
      FaceStep fs = {0, 0, 0, 0, false, 0.0, -1};
      double gxb = 0.0, gyb = 0.0;
      if (lane < nb) {
        const int b = lane;
        gxb = hu[3 * b] + gs[3 * b]; gyb = hu[3 * b + 1] + gs[3 * b + 1];
        fs = face_step(face[4 * b], face[4 * b + 1], face[4 * b + 2], face[4 * b + 3], gxb, gyb, hu[3 * b + 2] + gs[3 * b + 2],
                       dua[3 * b], dua[3 * b + 1], dua[3 * b + 2], mu, fzm, dual_tol, feas_tol, round < 3);
      }
      if (round >= 3) {
        const bool any_add = __any_sync(0xffffffffu, fs.add);
        const double dmax = wmax(fs.dscore);
        if (!any_add && lane < nb && fs.dwhich >= 0 && fs.dscore == dmax) face_single_drop(fs, gxb, gyb);
      }
      if (lane < nb) {
        const int b = lane;
        if (fs.sx != face[4 * b] || fs.sy != face[4 * b + 1] || fs.top != face[4 * b + 2] || fs.apex != face[4 * b + 3]) changed = 1;
        face[4 * b] = (signed char)fs.sx; face[4 * b + 1] = (signed char)fs.sy; face[4 * b + 2] = (signed char)fs.top; face[4 * b + 3] = (signed char)fs.apex;
      }
      const bool any_changed = __any_sync(0xffffffffu, changed);
      // NaN compares false everywhere above: a non-finite solve must never pass as verified
      const bool finite = !__any_sync(0xffffffffu, !isfinite(dua[lane] + dua[lane + 32] + hu[lane] + hu[lane + 32] + gs[lane] + gs[lane + 32]));
      __syncwarp();
      if (!finite) break;
      if (!any_changed) { polished = true; break; }
    }
    if (polished) {
      // One step of iterative refinement on the verified face.  An unrefined Cholesky solve carries
      // a forward error of ~ n eps cond(H) (6e-8 of the force scale at cond 4e6 on the parameter
      // sweep); hu = H dua is already here from the verification, so the stationarity residual of the
      // free coordinates, r = -T'(H u + g), costs nothing and one more triangular solve with the factor
      // still in Ls brings the answer to ~ eps cond(H).
      du[lane] = 0.0; du[lane + 32] = 0.0;
      __syncwarp();
      if (lane < nb) {
        const int b = lane;
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double gx = gs[3 * b] + hu[3 * b], gy = gs[3 * b + 1] + hu[3 * b + 1], gz = gs[3 * b + 2] + hu[3 * b + 2];
        const double gzt = gz + mu * (sx * gx + sy * gy);
        du[3 * b] = (apex || sx != 0) ? 0.0 : -gx;
        du[3 * b + 1] = (apex || sy != 0) ? 0.0 : -gy;
        du[3 * b + 2] = (apex || top) ? 0.0 : -gzt;
      }
      __syncwarp();
      wm_solve(Ls, du);
      __syncwarp();
      if (lane < nb) {
        const int b = lane;
        const int sx = face[4 * b], sy = face[4 * b + 1];
        const double vz = du[3 * b + 2];
        dua[3 * b] += du[3 * b] + mu * sx * vz;
        dua[3 * b + 1] += du[3 * b + 1] + mu * sy * vz;
        dua[3 * b + 2] += vz;
      }
      __syncwarp();
      u[lane] = dua[lane]; u[lane + 32] = dua[lane + 32];
      __syncwarp();
      status = 0;
      solved = true;
    } else if (ipm_broken) {
      status = 3;
      break;
    }
  }
  if (status < 0) status = 2;

  // ---------------- stage 5: outputs
  const bool ok = status < 2;
  {
    double* Uout = Bt.U ? Bt.U + (size_t)(12 * H) * p64 : nullptr;
    if (Uout) {
      for (int i = lane; i < 12 * H; i += 32) {
        const int b = bmap[i / 3];
        Uout[i] = (b >= 0 && ok) ? u[3 * b + i % 3] : 0.0;
      }
    }
    if (lane < 12) {
      const int b = bmap[lane / 3];  // step 0, leg lane/3
      Bt.grf[12 * p64 + lane] = (b >= 0 && ok) ? u[3 * b + lane % 3] : 0.0;
    }
    if (Bt.active) {
      unsigned int* aw = reinterpret_cast<unsigned int*>(wrow);  // scratch (W <= 24 words)
      __syncwarp();
      for (int w = lane; w < W; w += 32) aw[w] = 0u;
      __syncwarp();
      for (int idx = lane; idx < 4 * H; idx += 32) {
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
      __syncwarp();
      for (int w = lane; w < W; w += 32) Bt.active[(size_t)W * p64 + w] = aw[w];
    }
    if (wrec) {  // faces of the verified solution for the next solve of this problem
      for (int idx = lane; idx < 4 * H; idx += 32) {
        const int b = bmap[idx];
        wrec[idx] = (status == 0 && b >= 0) ? face_encode(face + 4 * b) : (unsigned char)0;
      }
      if (lane == 0) wrec[4 * H] = status == 0 ? 1 : 0;
    }
    if (lane == 0) {
      const int nf = n_ipm + n_polish;
      if (Bt.status) Bt.status[p64] = status;
      if (Bt.iters) Bt.iters[p64] = (nf & 0xffff) | (n_polish << 16);
    }
    const int nf = n_ipm + n_polish;
    if (status == 0 && n_ipm == 0) ws.warm_hits++;
    // no answer from the small-complementarity start: hand the problem to the retry launch (conservative start),
    // whose outputs replace the zeros written above; it is counted there
    const bool handed = status >= 2 && Bt.retry_list != nullptr;
    if (handed && lane == 0) Bt.retry_list[atomicAdd(Bt.retry_count, 1)] = pi;
    if (handed) ws.handed++;
    if (!handed) { ws.problems++; ws.sum_n += n; }
    ws.sum_fact += nf; ws.sum_polish += n_polish;
    if (status >= 2 && !handed) ws.not_ok++;
    if (status == 1) ws.ipm_only++;
    const double dn = (double)n;
    ws.flops += (double)nf * (dn * dn * dn / 3.0) + (double)n_ipm * (4.0 * dn * dn + 2.0 * dn * dn) +
                (double)n_polish * (2.0 * dn * dn + 4.0 * dn * dn);
  }
  __syncwarp();
}

}  // namespace lmpc
