// lmpc_tile_solve.cuh — one-warp-per-QP driver for reduced dimensions 64 <= n <= 8 NT (NT = 15:
// H = 20 schedules, NT = 24: the reference's H = 30 trot).  Same algorithm and formulas as
// solve_one_warp() (lmpc_warp_solve.cuh): SRBD build -> discretise -> structured condense ->
// Mehrotra interior point (M = H + C'WC kept in the L2 workspace, H u updated from the Newton
// system) -> exact active-set polish; a lane owns the stance blocks b = lane, lane + 32, ... and
// the per-block state lives in shared memory.  Reference citations as in lmpc_solver.cuh.
#pragma once
#include "lmpc_tile.cuh"
#include "lmpc_warp_solve.cuh"

namespace lmpc {

struct TilePlan {
  int oLS, oU, oGS, oHU, oRD, oDUA, oDU, oZ, oQ, oIS, oPRM, oXB, oTT;  // doubles
  int oBW, oCOS, oSIN, oCS, oSS, oE, oAJ, oBJ;           // setup scratch, inside LS
  int n_doubles;
  int oBLK, oFACE, oBMAP, total_bytes;
};
// NT above this keeps the factor in the CTA's global (L2 / HBM) workspace slot instead of shared memory.
constexpr int kTileSmemMaxNT = 24;
template <int NT>
__host__ __device__ inline size_t tile_slot_doubles() {
  using G = TileGeom<NT>;
  return (size_t)G::HS + 6 * G::NBMAX + 2 + (NT > kTileSmemMaxNT ? (size_t)G::LS : 0);
}
// factor storage of this CTA
template <int NT>
__device__ __forceinline__ double* tile_ls(const TilePlan& pl, unsigned char* smem, double* Hg) {
  using G = TileGeom<NT>;
  if (NT > kTileSmemMaxNT) return Hg + G::HS + 6 * G::NBMAX + 2;
  return reinterpret_cast<double*>(smem) + pl.oLS;
}
template <int NT>
__host__ __device__ inline TilePlan make_tile_plan(int H) {
  using G = TileGeom<NT>;
  TilePlan p;
  int s = 0;
  p.oLS = 0;
  p.oBW = s; s += 36;
  p.oCOS = s; s += H; p.oSIN = s; s += H; p.oCS = s; s += H; p.oSS = s; s += H;
  p.oE = s; s += 12 * H; p.oAJ = s; s += 3 * H; p.oBJ = s; s += 3 * H;
  int o = (NT > kTileSmemMaxNT || s > G::LS) ? s : G::LS;  // the setup scratch aliases the factor when that lives here
  p.oU = o; o += G::N;
  p.oGS = o; o += G::N;
  p.oHU = o; o += G::N;
  p.oRD = o; o += G::N;
  p.oDUA = o; o += G::N;
  p.oDU = o; o += G::N;
  p.oZ = o; o += 5 * G::NBMAX;
  p.oQ = o; o += 5 * G::NBMAX;
  p.oIS = o; o += 5 * G::NBMAX + (G::NBMAX & 1);
  p.oPRM = o; o += 32;
  p.oXB = o; o += 64;
  p.oTT = o; if (NT >= kTileSmemMaxNT) o += G::N;  // scratch vector of the multi-warp solve (classes that use it)
  p.n_doubles = o;
  int b = o * 8;
  p.oBLK = b; b += 2 * G::NBMAX;
  p.oFACE = b; b += 4 * G::NBMAX;
  b = (b + 1) & ~1;
  p.oBMAP = b; b += 2 * 4 * H;
  p.total_bytes = (b + 15) & ~15;
  return p;
}

// W_b = C_b' diag(w) C_b, entries (0,0) (2,0) (1,1) (2,1) (2,2); (1,0) = 0
__device__ __forceinline__ void blk_w(const double* w, double mu, double* wb) {
  wb[0] = w[0] + w[1]; wb[1] = mu * (w[0] - w[1]);
  wb[2] = w[2] + w[3]; wb[3] = mu * (w[2] - w[3]);
  wb[4] = mu * mu * (w[0] + w[1] + w[2] + w[3]) + w[4];
}
// store a symmetric 3x3 block v = (0,0) (1,0) (2,0) (1,1) (2,1) (2,2) at rows/cols 3b.. of the workspace
template <int NT>
__device__ __forceinline__ void blk_store(double* Hg, int b, const double* v) {
  using G = TileGeom<NT>;
  const int r = 3 * b;
  __stcg(Hg + G::hoff(r, r), v[0]);
  __stcg(Hg + G::hoff(r + 1, r + 1), v[3]);
  __stcg(Hg + G::hoff(r + 2, r + 2), v[5]);
  __stcg(Hg + G::hoff(r + 1, r), v[1]); __stcg(Hg + G::hoff(r, r + 1), v[1]);
  __stcg(Hg + G::hoff(r + 2, r), v[2]); __stcg(Hg + G::hoff(r, r + 2), v[2]);
  __stcg(Hg + G::hoff(r + 2, r + 1), v[4]); __stcg(Hg + G::hoff(r + 1, r + 2), v[4]);
}

// Reduced Hessian blocks (+ identity padding) into the L2 workspace, pair index strided over `nthr`
// threads: one warp, or every warp of the CTA when helper warps exist (job 3).
template <int NT>
__device__ void tile_hessian_blocks(const KParams& P, const TilePlan& pl, unsigned char* smem, double* Hg, int nb, int n,
                                    double sc, double im2, int tid, int nthr) {
  using G = TileGeom<NT>;
  constexpr int N = G::N;
  const int H = P.H;
  double* sd = reinterpret_cast<double*>(smem);
  const double* prm = sd + pl.oPRM;
  const double* qw = prm + PRM_Q;
  const double* rw = prm + PRM_R;
  const double* Bw = sd + pl.oBW;
  const double* cosv = sd + pl.oCOS;
  const double* sinv = sd + pl.oSIN;
  const double* Cs = sd + pl.oCS;
  const double* Ss = sd + pl.oSS;
  const unsigned char* blk = smem + pl.oBLK;
  double* Hd = Hg + G::HS;
  const double dt2 = P.dt * P.dt, theta = P.zoh ? 0.5 : 0.0;
  for (int idx = tid; idx < N * (N - n); idx += nthr) {  // identity padding for rows/cols >= n
    const int c = n + idx / N, r = idx % N;
    __stcg(Hg + G::hoff(r, c), (r == c) ? 1.0 : 0.0);
    __stcg(Hg + G::hoff(c, r), (r == c) ? 1.0 : 0.0);
  }
  const int npairs = (nb * (nb + 1)) >> 1;
  for (int pr = tid; pr < npairs; pr += nthr) {
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
        __stcg(Hg + G::hoff(r, c), v);
        __stcg(Hg + G::hoff(c, r), v);
      }
    if (b1 == b2) {  // keep the diagonal block: the interior point overwrites it with H_bb + W_b
      double* hd = Hd + 6 * b1;
      __stcg(hd + 0, sc * o[0]); __stcg(hd + 1, sc * o[1]); __stcg(hd + 2, sc * o[2]);
      __stcg(hd + 3, sc * o[4]); __stcg(hd + 4, sc * o[5]); __stcg(hd + 5, sc * o[8]);
    }
  }
}

// Reduced matrix of the polish, Mt = T' H T with pinned rows/cols replaced by identity, from the L2
// workspace into the factor storage; same striding (job 4).
template <int NT>
__device__ void tile_reduced_matrix(const TilePlan& pl, unsigned char* smem, const double* Hg, int nb, int n, double mu,
                                    int tid, int nthr) {
  using G = TileGeom<NT>;
  constexpr int N = G::N;
  double* Ls = tile_ls<NT>(pl, smem, const_cast<double*>(Hg));
  const signed char* face = reinterpret_cast<const signed char*>(smem + pl.oFACE);
  const int npairs = (nb * (nb + 1)) >> 1;
  for (int pr = tid; pr < npairs; pr += nthr) {
    int b1, b2;
    pair_from_index(pr, b1, b2);
    double X[9];
#pragma unroll
    for (int a1 = 0; a1 < 3; ++a1)
#pragma unroll
      for (int a2 = 0; a2 < 3; ++a2) X[a1 + 3 * a2] = __ldcg(Hg + G::hoff(3 * b1 + a1, 3 * b2 + a2));
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
          Ls[G::loff(r, c)] = v;
          if ((r >> 3) == (c >> 3)) Ls[G::loff(c, r)] = v;  // diagonal tiles are kept full
        }
      }
  }
  for (int r = n; r < N; ++r)  // identity padding rows n..N-1
    for (int c = tid; c <= r; c += nthr) {
      Ls[G::loff(r, c)] = (r == c) ? 1.0 : 0.0;
      if ((r >> 3) == (c >> 3)) Ls[G::loff(c, r)] = (r == c) ? 1.0 : 0.0;
    }
}

// Mailbox arguments of the shared jobs.
struct TileJobArgs { int nb, n, nte; double sc, im2, mu; const double* x; double* y; };
__device__ __forceinline__ TileJobArgs* tile_job_args() {
  __shared__ TileJobArgs a;
  return &a;
}
// Warp 0's call: runs job 3 / 4 alone (NW = 1) or together with the helper warps.
template <int NT, int NW>
__device__ __forceinline__ void tile_shared(int job, const KParams& P, const TilePlan& pl, unsigned char* smem, double* Hg,
                                            int nb, int n, double sc, double im2, double mu) {
  if (NW == 1) {
    if (job == 3) tile_hessian_blocks<NT>(P, pl, smem, Hg, nb, n, sc, im2, threadIdx.x & 31, 32);
    else tile_reduced_matrix<NT>(pl, smem, Hg, nb, n, mu, threadIdx.x & 31, 32);
    return;
  }
  if ((threadIdx.x & 31) == 0) {
    TileJobArgs* a = tile_job_args();
    a->nb = nb; a->n = n; a->sc = sc; a->im2 = im2; a->mu = mu;
    *tile_job() = job;
  }
  __syncthreads();  // release the helpers (also publishes warp 0's shared-memory state)
  if (job == 3) tile_hessian_blocks<NT>(P, pl, smem, Hg, nb, n, sc, im2, threadIdx.x, 32 * NW);
  else tile_reduced_matrix<NT>(pl, smem, Hg, nb, n, mu, threadIdx.x, 32 * NW);
  __syncthreads();  // every warp's share is written
}
// y = H x with the tile rows dealt over the warps of the CTA (job 5); x, y in shared memory.
template <int NT, int NW>
__device__ __forceinline__ void tile_symv(const double* Hg, const double* x, double* y, int nte) {
  if (NW == 1) { gm_symv<NT>(Hg, x, y, nte); return; }
  if ((threadIdx.x & 31) == 0) {
    TileJobArgs* a = tile_job_args();
    a->x = x; a->y = y; a->nte = nte;
    *tile_job() = 5;
  }
  __syncthreads();
  gm_symv<NT>(Hg, x, y, nte, 0, NW);
  __syncthreads();
}
// Does this class run the triangular solves on all warps of the CTA (diagonal blocks of the factor inverted)?
// Measured (solves/s, one-warp solve vs plain blocked substitution over the CTA): NT=15 C3 618k vs 535k, NT=24
// with four warps trot H=30 139k vs 137k (shared-memory factor: the barriers cost what the parallel rows save);
// NT=36/48 all-gait H=30 43.6k vs 50.6k (factor in L2).  With inverted diagonal blocks: see DESIGN.md §6.
template <int NT, int NW>
__host__ __device__ constexpr bool tile_mw_solve() { return NW > 1 && !(NT <= 24 && NW < 8); }

// What warp 0 calls in place of gm_factor: the factorization on all warps (job 1 / 2), followed by the in-place
// inversion of the diagonal blocks when the class solves on all warps.
template <int NT, int NW, bool FROM_G>
__device__ __forceinline__ bool tile_factor(const double* Mg, double* Ls, double* xb, int nte) {
  if (NW == 1) return gm_factor<NT, FROM_G>(Mg, Ls, xb, nte);
  if ((threadIdx.x & 31) == 0) { *tile_job_nte() = nte; *tile_job() = FROM_G ? 1 : 2; }
  __syncthreads();  // releases the helper warps; also publishes warp 0's writes of M
  const bool ok = gm_factor_mw<NT, FROM_G, NW>(Mg, Ls, xb, 0, nte);
  if (tile_mw_solve<NT, NW>()) gm_invert_blocks_mw<NT, NW>(Ls, nte, 0);
  __syncwarp();
  return ok;
}
// L L' x = b (job 6 on all warps, or one warp); x and the scratch vector tt in shared memory.
template <int NT, int NW>
__device__ __forceinline__ void tile_solve(const double* Ls, double* x, double* tt, int nte) {
  if (!tile_mw_solve<NT, NW>()) { gm_solve<NT>(Ls, x, nte); return; }
  if ((threadIdx.x & 31) == 0) {
    TileJobArgs* a = tile_job_args();
    a->y = x; a->nte = nte;
    *tile_job() = 6;
  }
  __syncthreads();
  gm_solve_mw<NT, NW>(Ls, x, tt, nte, 0);
}
// Helper warps (1 .. NW-1) live here until warp 0 posts job 0.
template <int NT, int NW>
__device__ __forceinline__ void tile_helper_loop(const KParams& P, const TilePlan& pl, unsigned char* smem, double* Hg) {
  double* sd = reinterpret_cast<double*>(smem);
  const int w = threadIdx.x >> 5;
  for (;;) {
    __syncthreads();
    const int job = *tile_job();
    if (job == 0) return;
    if (job == 1 || job == 2) {
      if (job == 1) gm_factor_mw<NT, true, NW>(Hg, tile_ls<NT>(pl, smem, Hg), sd + pl.oXB, w, *tile_job_nte());
      else gm_factor_mw<NT, false, NW>(Hg, tile_ls<NT>(pl, smem, Hg), sd + pl.oXB, w, *tile_job_nte());
      if (tile_mw_solve<NT, NW>()) gm_invert_blocks_mw<NT, NW>(tile_ls<NT>(pl, smem, Hg), *tile_job_nte(), w);
    } else {
      const TileJobArgs a = *tile_job_args();
      if (job == 3) tile_hessian_blocks<NT>(P, pl, smem, Hg, a.nb, a.n, a.sc, a.im2, threadIdx.x, 32 * NW);
      else if (job == 4) tile_reduced_matrix<NT>(pl, smem, Hg, a.nb, a.n, a.mu, threadIdx.x, 32 * NW);
      else if (job == 5) gm_symv<NT>(Hg, a.x, a.y, a.nte, w, NW);
      else { gm_solve_mw<NT, NW>(tile_ls<NT>(pl, smem, Hg), a.y, sd + pl.oTT, a.nte, w); continue; }  // job 6 ends on its own barrier
      __syncthreads();
    }
  }
}

template <int NT, int NW = 1>
__device__ void solve_one_tile(const KParams& P, const KBatch& Bt, WarpStats& ws, const TilePlan& pl,
                               unsigned char* smem, double* Hg, int pi) {
  using G = TileGeom<NT>;
  constexpr int N = G::N;
  const int lane = threadIdx.x & 31;
  const int H = P.H;
  double* sd = reinterpret_cast<double*>(smem);
  double* Ls = tile_ls<NT>(pl, smem, Hg);
  double* u = sd + pl.oU;
  double* gs = sd + pl.oGS;
  double* hu = sd + pl.oHU;
  double* rd = sd + pl.oRD;
  double* dua = sd + pl.oDUA;
  double* du = sd + pl.oDU;
  double* z = sd + pl.oZ;
  double* qv = sd + pl.oQ;    // per-row scratch of the interior point: q = ds_aff / s, then dz
  double* isv = sd + pl.oIS;  // 1 / s
  double* prm = sd + pl.oPRM;
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
  double* Hd = Hg + G::HS;  // 6 doubles per block: the diagonal 3x3 blocks of the scaled Hessian

  const double dt = P.dt, dt2 = P.dt * P.dt, theta = P.zoh ? 0.5 : 0.0;
  const size_t p64 = (size_t)pi;
  const double* x0 = Bt.x0 + 12 * p64;
  const double* rot = Bt.rot + 9 * p64;
  const double* foot = Bt.foot + 12 * p64;
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
    // a1_trunk_inertia_world = R I R' and its inverse (ConvexQPSolver.cpp:201-205)
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
    for (int leg = 0; leg < 4; ++leg) {  // I_w^-1 skew(r_leg)  (Utils.cpp:89-95)
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
    if (fl && pos < G::NBMAX) { blk[2 * pos] = (unsigned char)(idx >> 2); blk[2 * pos + 1] = (unsigned char)(idx & 3); }
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
  const int nte = max(1, (n + 7) >> 3);  // 8x8 tiles this problem fills; the rest of NT is identity padding
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
      double* ek = e + 12 * k;  // x_ref(k) per ConvexQPSolver.cpp:264-276
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
  if (n == 0 || n > N) {  // nothing to solve / does not fit this size class (dispatch error): zeros
    const bool misfit = n > N;
    if (lane == 0) {
      if (Bt.status) Bt.status[p64] = misfit ? 2 : 0;
      if (Bt.iters) Bt.iters[p64] = 0;
    }
    if (lane < 12) Bt.grf[12 * p64 + lane] = 0.0;
    if (Bt.U) for (int i = lane; i < 12 * H; i += 32) Bt.U[(size_t)(12 * H) * p64 + i] = 0.0;
    if (Bt.active)
      for (int w = lane; w < W; w += 32) {
        const int lo = 32 * w, hi = min(32 * w + 32, 24 * H);
        Bt.active[(size_t)W * p64 + w] = (hi - lo == 32) ? 0xffffffffu : ((1u << (hi - lo)) - 1u);
      }
    ws.problems++;
    if (misfit) ws.not_ok++;
    __syncwarp();
    return;
  }

  // ---------------- stage 3: reduced Hessian into the L2 workspace (full symmetric, N x N)
  const double im2 = 1.0 / (mass * mass);
  double sc;
  {
    double dmax = 0;
    for (int b = lane; b < nb; b += 32) {
      double o[9];
      const int j = blk[2 * b], l = blk[2 * b + 1];
      wblock(H, dt2, theta, P.cs, im2, qw, rw, Bw, cosv, sinv, Cs, Ss, j, l, j, l, true, o);
      dmax = fmax(dmax, fmax(fabs(o[0]), fmax(fabs(o[4]), fabs(o[8]))));
    }
    dmax = wmax(dmax);
    sc = 1.0 / dmax;
    tile_shared<NT, NW>(3, P, pl, smem, Hg, nb, n, sc, im2, mu);
    for (int b = lane; b < nb; b += 32) {
      const int j = blk[2 * b], l = blk[2 * b + 1];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double* bc = Bw + 3 * (3 * l + a);
        gs[3 * b + a] = sc * P.cs * (bc[0] * aj[3 * j] + bc[1] * aj[3 * j + 1] + bc[2] * aj[3 * j + 2] + bj[3 * j + a] / mass);
      }
    }
    for (int i = n + lane; i < N; i += 32) gs[i] = 0.0;
  }
  __threadfence_block();
  __syncwarp();  // workspace visible to every lane; the setup scratch inside Ls is dead from here on
  double gsinf = 0;
  for (int i = lane; i < N; i += 32) gsinf = fmax(gsinf, fabs(gs[i]));
  gsinf = wmax(gsinf);
  const double rd_tol = 1e-9 * (gsinf > 1.0 ? gsinf : 1.0);
  const double rd_tol0 = P.rdtol0 * (gsinf > 1.0 ? gsinf : 1.0);
  const double dual_tol = P.dual_rel * (gsinf > sc ? gsinf : sc);
  const double feas_tol = 1e-9;
  const int m = 5 * nb;
  const double inv_m = 1.0 / (double)m;

  // ---------------- stage 4a: interior point (feasible start); see solve_one_warp for the rules
  for (int i = lane; i < N; i += 32) { u[i] = 0.0; hu[i] = 0.0; dua[i] = 0.0; du[i] = 0.0; rd[i] = 0.0; }
  __syncwarp();
  for (int b = lane; b < nb; b += 32) {
    double fz_init = P.fz0 * fzm;
    if (P.fz0 <= 0.0) {  // weight shared by the stance legs of this block's step
      const uint8_t* ck = contact + 4 * blk[2 * b];
      const int ns = (ck[0] != 0) + (ck[1] != 0) + (ck[2] != 0) + (ck[3] != 0);
      fz_init = fmin(0.9 * fzm, fmax(0.05 * fzm, mass * P.gravity / (double)ns));
    }
    u[3 * b + 2] = fz_init;
  }
  __syncwarp();
  int n_ipm = 0, n_polish = 0, status = -1;
  const int max_iter = P.max_iter;
  bool solved = false, ipm_inited = false;
  // warm start: see solve_one_warp
  uint8_t* wrec = Bt.warm ? Bt.warm + (size_t)warm_stride(H) * p64 : nullptr;
  const bool warm = wrec && (wrec[4 * H] != 0 || P.try_interior);
#pragma unroll 1
  for (int stage = warm ? -1 : 0; stage < 3 && !solved; ++stage) {
    const double mu_tol = stage <= 0 ? P.tol0 : (stage == 1 ? 1e-10 : 1e-13);
    const double rdt = stage <= 0 ? rd_tol0 : rd_tol;
    bool ipm_converged = false, ipm_broken = false;
    if (stage >= 0 && !ipm_inited) {  // interior-point start (skipped entirely by a verified warm start)
      ipm_inited = true;
      {
        tile_symv<NT, NW>(Hg, u, hu, nte);
        __syncwarp();
        double gi = 0, ssum = 0;
        for (int i = lane; i < N; i += 32) gi = fmax(gi, fabs(hu[i] + gs[i]));
        for (int b = lane; b < nb; b += 32) {
          double s[5];
          blk_slack(u + 3 * b, mu, fzm, s);
          ssum += s[0] + s[1] + s[2] + s[3] + s[4];
        }
        const double ginf = wmax(gi);
        ssum = wsum(ssum);
        const double zc = P.z0_scale * (ginf > 1e-6 ? ginf : 1e-6) * (ssum * inv_m);
        for (int b = lane; b < nb; b += 32) {
          double s[5];
          blk_slack(u + 3 * b, mu, fzm, s);
#pragma unroll
          for (int r = 0; r < 5; ++r) z[5 * b + r] = zc * frcp(s[r]);
        }
      }
      __syncwarp();
    }
#pragma unroll 1
    while (stage >= 0 && n_ipm < max_iter) {
      double rdinf = 0, sz = 0;
      for (int b = lane; b < nb; b += 32) {
        double s[5], ct[3];
        blk_slack(u + 3 * b, mu, fzm, s);
        const double* zb = z + 5 * b;
        blk_ct(zb, mu, ct);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          const double v = hu[3 * b + a] + gs[3 * b + a] + ct[a];
          rd[3 * b + a] = v;
          rdinf = fmax(rdinf, fabs(v));
        }
#pragma unroll
        for (int r = 0; r < 5; ++r) sz = fma(s[r], zb[r], sz);
      }
      rdinf = wmax(rdinf);
      sz = wsum(sz);
      const double mub = sz * inv_m;
      if (!(rdinf == rdinf) || !(mub == mub)) { ipm_broken = true; break; }
      if (rdinf <= rdt && mub <= mu_tol) {
        // stages 0 and 1 hand over to the exact polish, which verifies KKT on H itself (see solve_one_warp);
        // the last stage confirms the claim made on the recursively updated H u with an exact product on H
        if (stage < 2) { ipm_converged = true; break; }
        for (int b = lane; b < nb; b += 32) {
          double hd[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) hd[k] = __ldcg(Hd + 6 * b + k);
          blk_store<NT>(Hg, b, hd);
        }
        __syncwarp();
        tile_symv<NT, NW>(Hg, u, hu, nte);
        __syncwarp();
        double ri = 0;
        for (int b = lane; b < nb; b += 32) {
          double ct[3];
          blk_ct(z + 5 * b, mu, ct);
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            const double v = hu[3 * b + a] + gs[3 * b + a] + ct[a];
            rd[3 * b + a] = v;
            ri = fmax(ri, fabs(v));
          }
        }
        rdinf = wmax(ri);
        if (rdinf <= rdt) { ipm_converged = true; break; }
      }
      ++n_ipm;
      // M_bb = H_bb + W_b into the workspace; predictor rhs -(H u + g)
      for (int b = lane; b < nb; b += 32) {
        double s[5], w[5], wb[5], hd[6];
        blk_slack(u + 3 * b, mu, fzm, s);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const double is = frcp(s[r]);
          isv[5 * b + r] = is;
          w[r] = z[5 * b + r] * is;
        }
        blk_w(w, mu, wb);
#pragma unroll
        for (int k = 0; k < 6; ++k) hd[k] = __ldcg(Hd + 6 * b + k);
        hd[0] += wb[0]; hd[2] += wb[1]; hd[3] += wb[2]; hd[4] += wb[3]; hd[5] += wb[4];
        blk_store<NT>(Hg, b, hd);
#pragma unroll
        for (int a = 0; a < 3; ++a) dua[3 * b + a] = -(hu[3 * b + a] + gs[3 * b + a]);
      }
      __syncwarp();
      if (!tile_factor<NT, NW, true>(Hg, Ls, xb, nte)) { ipm_broken = true; break; }
      tile_solve<NT, NW>(Ls, dua, sd + pl.oTT, nte);
      __syncwarp();
      double tau = 0.0;
      for (int b = lane; b < nb; b += 32) {
        double dsa[5];
        blk_dslack(dua + 3 * b, mu, dsa);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const double q = dsa[r] * isv[5 * b + r];
          qv[5 * b + r] = q;
          tau = fmax(tau, fmax(-q, 1.0 + q));
        }
      }
      tau = wmax(tau);
      const double amax = tau > 1.0 ? frcp(tau) : 1.0;
      double szaff = 0;
      for (int b = lane; b < nb; b += 32) {
        double s[5];
        blk_slack(u + 3 * b, mu, fzm, s);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const double q = qv[5 * b + r];
          szaff = fma(s[r] * z[5 * b + r], fma(amax, q, 1.0) * (1.0 - amax * (1.0 + q)), szaff);
        }
      }
      szaff = wsum(szaff);
      const double rat = szaff * frcp(sz);
      const double smu = (P.sigma_pow == 2 ? rat * rat : (P.sigma_pow == 4 ? rat * rat * rat * rat : rat * rat * rat)) * mub;  // sigma * mu
      // corrector: rhs = -rd - C'((sigma mu - ds_aff dz_aff - s z)/s); rd keeps the rhs
      for (int b = lane; b < nb; b += 32) {
        double w2[5], ct[3];
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const double q = qv[5 * b + r];
          w2[r] = fma(z[5 * b + r], fma(q, 1.0 + q, -1.0), smu * isv[5 * b + r]);
        }
        blk_ct(w2, mu, ct);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          const double rhs = -rd[3 * b + a] - ct[a];
          rd[3 * b + a] = rhs;
          du[3 * b + a] = rhs;
        }
      }
      __syncwarp();
      tile_solve<NT, NW>(Ls, du, sd + pl.oTT, nte);
      __syncwarp();
      double tau2 = 0.0;
      for (int b = lane; b < nb; b += 32) {
        double ds[5];
        blk_dslack(du + 3 * b, mu, ds);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const double q = qv[5 * b + r], is = isv[5 * b + r], zz = z[5 * b + r];
          const double w2 = fma(zz, fma(q, 1.0 + q, -1.0), smu * is);
          const double dz = fma(-zz * is, ds[r], w2);
          qv[5 * b + r] = dz;
          tau2 = fmax(tau2, fmax(-ds[r] * is, -dz * frcp(zz)));
        }
      }
      tau2 = wmax(tau2);
      const double alpha = tau2 > P.step_frac ? P.step_frac * frcp(tau2) : 1.0;
      for (int b = lane; b < nb; b += 32) {
        double w[5], wb[5], d3[3];
#pragma unroll
        for (int r = 0; r < 5; ++r) w[r] = z[5 * b + r] * isv[5 * b + r];
        blk_w(w, mu, wb);
#pragma unroll
        for (int a = 0; a < 3; ++a) d3[a] = du[3 * b + a];
        // H du = rhs - W_b du  (Newton system), so H u follows without a matrix product
        const double h0 = rd[3 * b] - (wb[0] * d3[0] + wb[1] * d3[2]);
        const double h1 = rd[3 * b + 1] - (wb[2] * d3[1] + wb[3] * d3[2]);
        const double h2 = rd[3 * b + 2] - (wb[1] * d3[0] + wb[3] * d3[1] + wb[4] * d3[2]);
        hu[3 * b] = fma(alpha, h0, hu[3 * b]);
        hu[3 * b + 1] = fma(alpha, h1, hu[3 * b + 1]);
        hu[3 * b + 2] = fma(alpha, h2, hu[3 * b + 2]);
#pragma unroll
        for (int r = 0; r < 5; ++r) z[5 * b + r] = fma(alpha, qv[5 * b + r], z[5 * b + r]);
#pragma unroll
        for (int a = 0; a < 3; ++a) u[3 * b + a] = fma(alpha, d3[a], u[3 * b + a]);
      }
      __syncwarp();
    }
    // the polish works on H itself: put the diagonal blocks back
    for (int b = lane; b < nb; b += 32) {
      double hd[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) hd[k] = __ldcg(Hd + 6 * b + k);
      blk_store<NT>(Hg, b, hd);
    }
    __syncwarp();
    if (stage == 2 && !ipm_broken) {  // (a broken interior point still gets one polish: see solve_one_warp)
      status = ipm_converged ? 1 : 2;
      solved = ipm_converged;
      break;
    }

    // ---------------- stage 4b: exact active-set polish on the identified pyramid faces
    if (stage < 0) {
      for (int b = lane; b < nb; b += 32) warm_face_load(wrec, H, Bt.warm_shift, blk[2 * b], blk[2 * b + 1], face + 4 * b);
      __syncwarp();
    } else {
      double smax = 0, zmax = 0;
      for (int b = lane; b < nb; b += 32) {
        double s[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        for (int r = 0; r < 5; ++r) { smax = fmax(smax, s[r]); zmax = fmax(zmax, z[5 * b + r]); }
      }
      smax = wmax(smax);
      zmax = wmax(zmax);
      const double bal = smax / (zmax > 1e-300 ? zmax : 1e-300);
      for (int b = lane; b < nb; b += 32) {
        double s[5];
        blk_slack(u + 3 * b, mu, fzm, s);
        bool actr[5];
        for (int r = 0; r < 5; ++r) actr[r] = s[r] < z[5 * b + r] * bal;
        face[4 * b + 0] = actr[0] ? -1 : (actr[1] ? 1 : 0);
        face[4 * b + 1] = actr[2] ? -1 : (actr[3] ? 1 : 0);
        face[4 * b + 2] = actr[4] ? 1 : 0;
        face[4 * b + 3] = ((actr[0] && actr[1]) || (actr[2] && actr[3])) ? 1 : 0;
      }
      __syncwarp();
    }
    bool polished = false;
    for (int round = 0; round < 10; ++round) {
      ++n_polish;
      // p (pinned particular solution) in rd ; hu = H p
      for (int i = lane; i < N; i += 32) rd[i] = 0.0;
      __syncwarp();
      bool pinned_top = false;
      for (int b = lane; b < nb; b += 32) {
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        pinned_top = pinned_top || (top && !apex);
        const double pz = (top && !apex) ? fzm : 0.0;
        rd[3 * b] = sx * mu * pz; rd[3 * b + 1] = sy * mu * pz; rd[3 * b + 2] = pz;
      }
      __syncwarp();
      if (__any_sync(0xffffffffu, pinned_top)) {
        tile_symv<NT, NW>(Hg, rd, hu, nte);
      } else {  // p = 0: H p = 0 without reading H
        for (int i = lane; i < N; i += 32) hu[i] = 0.0;
      }
      __syncwarp();
      // Mt = T' H T with pinned rows/cols replaced by identity, written into the factor storage
      tile_shared<NT, NW>(4, P, pl, smem, Hg, nb, n, sc, im2, mu);
      __syncwarp();
      for (int i = lane; i < N; i += 32) du[i] = 0.0;
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double gx = gs[3 * b] + hu[3 * b], gy = gs[3 * b + 1] + hu[3 * b + 1], gz = gs[3 * b + 2] + hu[3 * b + 2];
        const double gzt = gz + mu * (sx * gx + sy * gy);
        du[3 * b] = (apex || sx != 0) ? 0.0 : -gx;
        du[3 * b + 1] = (apex || sy != 0) ? 0.0 : -gy;
        du[3 * b + 2] = (apex || top) ? 0.0 : -gzt;
      }
      __syncwarp();
      if (!tile_factor<NT, NW, false>(Hg, Ls, xb, nte)) break;
      tile_solve<NT, NW>(Ls, du, sd + pl.oTT, nte);
      for (int i = lane; i < N; i += 32) dua[i] = 0.0;
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {
        const int sx = face[4 * b], sy = face[4 * b + 1];
        const double vx = du[3 * b], vy = du[3 * b + 1], vz = du[3 * b + 2];
        dua[3 * b] = vx + mu * sx * vz + rd[3 * b];
        dua[3 * b + 1] = vy + mu * sy * vz + rd[3 * b + 1];
        dua[3 * b + 2] = vz + rd[3 * b + 2];
      }
      __syncwarp();
      tile_symv<NT, NW>(Hg, dua, hu, nte);
      __syncwarp();
      // Face repair (face_step, lmpc_blocks.cuh; same rule as solve_one_warp): rounds 0-2 apply every indicated
      // change at once; from round 3 on, adds while any block is primal infeasible, else the single most
      // violated multiplier of the whole problem is dropped.  A lane owns the blocks b = lane, lane + 32, ...
      int changed = 0;
      bool any_add = false;
      double dmax = 0.0;
      if (round >= 3) {  // first pass: is any block infeasible, and what is the largest multiplier violation
        bool add = false;
        double dloc = 0.0;
        for (int b = lane; b < nb; b += 32) {
          const FaceStep fs = face_step(face[4 * b], face[4 * b + 1], face[4 * b + 2], face[4 * b + 3], hu[3 * b] + gs[3 * b],
                                        hu[3 * b + 1] + gs[3 * b + 1], hu[3 * b + 2] + gs[3 * b + 2], dua[3 * b], dua[3 * b + 1],
                                        dua[3 * b + 2], mu, fzm, dual_tol, feas_tol, false);
          add = add || fs.add;
          dloc = fmax(dloc, fs.dscore);
        }
        any_add = __any_sync(0xffffffffu, add);
        dmax = wmax(dloc);
      }
      for (int b = lane; b < nb; b += 32) {
        const double gx = hu[3 * b] + gs[3 * b], gy = hu[3 * b + 1] + gs[3 * b + 1];
        FaceStep fs = face_step(face[4 * b], face[4 * b + 1], face[4 * b + 2], face[4 * b + 3], gx, gy, hu[3 * b + 2] + gs[3 * b + 2],
                                dua[3 * b], dua[3 * b + 1], dua[3 * b + 2], mu, fzm, dual_tol, feas_tol, round < 3);
        if (round >= 3 && !any_add && fs.dwhich >= 0 && fs.dscore == dmax) face_single_drop(fs, gx, gy);
        if (fs.sx != face[4 * b] || fs.sy != face[4 * b + 1] || fs.top != face[4 * b + 2] || fs.apex != face[4 * b + 3]) changed = 1;
        face[4 * b] = (signed char)fs.sx; face[4 * b + 1] = (signed char)fs.sy; face[4 * b + 2] = (signed char)fs.top; face[4 * b + 3] = (signed char)fs.apex;
      }
      const bool any_changed = __any_sync(0xffffffffu, changed);
      double fsum = 0.0;  // NaN compares false everywhere above: a non-finite solve must never pass as verified
      for (int i = lane; i < N; i += 32) fsum += dua[i] + hu[i] + gs[i];
      const bool finite = !__any_sync(0xffffffffu, !isfinite(fsum));
      __syncwarp();
      if (!finite) break;
      if (!any_changed) { polished = true; break; }
    }
    if (polished) {
      // one step of iterative refinement on the verified face (see solve_one_warp): r = -T'(H u + g)
      // on the free coordinates from the hu the verification just computed, factor still in Ls
      for (int i = lane; i < N; i += 32) du[i] = 0.0;
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {
        const int sx = face[4 * b], sy = face[4 * b + 1], top = face[4 * b + 2], apex = face[4 * b + 3];
        const double gx = gs[3 * b] + hu[3 * b], gy = gs[3 * b + 1] + hu[3 * b + 1], gz = gs[3 * b + 2] + hu[3 * b + 2];
        const double gzt = gz + mu * (sx * gx + sy * gy);
        du[3 * b] = (apex || sx != 0) ? 0.0 : -gx;
        du[3 * b + 1] = (apex || sy != 0) ? 0.0 : -gy;
        du[3 * b + 2] = (apex || top) ? 0.0 : -gzt;
      }
      __syncwarp();
      tile_solve<NT, NW>(Ls, du, sd + pl.oTT, nte);
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {
        const int sx = face[4 * b], sy = face[4 * b + 1];
        const double vz = du[3 * b + 2];
        dua[3 * b] += du[3 * b] + mu * sx * vz;
        dua[3 * b + 1] += du[3 * b + 1] + mu * sy * vz;
        dua[3 * b + 2] += vz;
      }
      __syncwarp();
      for (int i = lane; i < N; i += 32) u[i] = dua[i];
      __syncwarp();
      status = 0;
      solved = true;
    } else if (ipm_broken) {
      status = 3;
      break;
    } else {
      // back to the interior point at a tighter mu: H u of the interior iterate is needed again
      tile_symv<NT, NW>(Hg, u, hu, nte);
      __syncwarp();
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
      unsigned int* aw = reinterpret_cast<unsigned int*>(qv);  // scratch (W <= 24 words)
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
    if (wrec) {
      for (int idx = lane; idx < 4 * H; idx += 32) {
        const int b = bmap[idx];
        wrec[idx] = (status == 0 && b >= 0) ? face_encode(face + 4 * b) : (unsigned char)0;
      }
      if (lane == 0) wrec[4 * H] = status == 0 ? 1 : 0;
    }
    const int nf = n_ipm + n_polish;
    if (lane == 0) {
      if (Bt.status) Bt.status[p64] = status;
      if (Bt.iters) Bt.iters[p64] = (nf & 0xffff) | (n_polish << 16);
    }
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
