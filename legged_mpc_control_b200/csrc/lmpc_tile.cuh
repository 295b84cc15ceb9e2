// lmpc_tile.cuh — dense FP64 kernels of the one-warp-per-QP path for ANY tile count NT (padded
// dimension 8 NT; used for 64 <= n <= 192, i.e. H = 20 / H = 30 schedules).  Same fragment-ordered
// 8x8 tiles, DMMA trailing updates and t-layout vectors as lmpc_wmma.cuh (read its header first),
// with the triangular solves and H*x rolled over the tile index (vectors stay in shared memory)
// and the factorization unrolled over the tile column only (static accumulator indices).
//
// Factor storage: lower tiles, COLUMN-major packed, tile (I, K) (I >= K) at index
//   lcol(K) + (I - K),   lcol(K) = K NT - K (K - 1) / 2,
// so the tiles of one column are contiguous (immediate address offsets in the unrolled i loops).
// The Hessian workspace in global memory (L2) is full symmetric, tile (I, J) at index I NT + J.
#pragma once
#include "lmpc_wmma.cuh"

namespace lmpc {

template <int NT>
struct TileGeom {
  static constexpr int N = 8 * NT;                   // padded dimension
  static constexpr int LTILES = NT * (NT + 1) / 2;   // tiles of the factor
  static constexpr int LS = LTILES * 64;             // doubles of factor storage
  static constexpr int HS = NT * NT * 64;            // doubles of the Hessian workspace
  static constexpr int NBMAX = (8 * NT) / 3;         // stance blocks that fit
  __host__ __device__ static constexpr int lcol(int K) { return K * NT - (K * (K - 1)) / 2; }
  // element offsets (in doubles)
  __device__ static __forceinline__ int hoff(int r, int c) { return frag_off((r >> 3) * NT + (c >> 3), r, c); }
  __device__ static __forceinline__ int loff(int r, int c) { return frag_off(lcol(c >> 3) + (r >> 3) - (c >> 3), r, c); }
};

#ifndef LMPC_LATE_M
#define LMPC_LATE_M 1
#endif
constexpr bool kLateM = LMPC_LATE_M != 0;

// One tile column of the factorization with a compile-time tile count CNT = NT - J (static
// accumulator indices, no predicated tensor instructions); the K loop stays rolled.
template <int NT, int CNT, bool FROM_G>
__device__ __forceinline__ bool gm_column(const double2* Mf, double2* Lf, double* xb, int nte) {
  using G = TileGeom<NT>;
  constexpr int J = NT - CNT;
  if (J >= nte) return false;  // identity padding beyond the problem's own tile count: never read back
  double2* colL = Lf + 32 * G::lcol(J);              // tile (J + i, J) at colL[32 i]
  const double2* colM = Mf + 32 * (J * NT + J);      // tile (J + i, J) at colM[32 NT i]
  // FROM_G: the tiles of M come from L2 (~400 cycles).  They are requested here and SUBTRACTED after the K loop,
  // so the latency runs under the trailing updates instead of in front of them.
  double2 acc[CNT], mv[CNT];
#pragma unroll
  for (int i = 0; i < CNT; ++i) {
    mv[i] = FROM_G ? __ldcg(colM + 32 * NT * i) : colL[32 * i];
    acc[i] = kLateM && FROM_G && NT <= 24 ? make_double2(0.0, 0.0) : make_double2(-mv[i].x, -mv[i].y);
  }
  if (J > 0) {
    const double2* p = Lf + 32 * J;  // tile (J, 0); column K + 1 starts NT - K - 1 tiles further
#pragma unroll 1
    for (int K = 0; K < J; ++K) {  // acc[i] += L_{J+i,K} L_{J,K}'
      const double2 bj = p[0];
      dmma_xyt(acc[0], bj, bj);
#pragma unroll
      for (int i = 1; i < CNT; ++i) dmma_xyt(acc[i], p[32 * i], bj);
      p += 32 * (NT - K - 1);
    }
  }
  if (kLateM && FROM_G && NT <= 24) {  // (no gain where the factor itself streams from L2)
#pragma unroll
    for (int i = 0; i < CNT; ++i) { acc[i].x -= mv[i].x; acc[i].y -= mv[i].y; }
  }
  const double2 li = wm_diag8u(-acc[0].x, -acc[0].y, xb);
  colL[0] = li;
  const double2 nli = make_double2(-li.x, -li.y);
#pragma unroll
  for (int i = 1; i < CNT; ++i) {
    double2 x = make_double2(0.0, 0.0);
    dmma_xyt(x, acc[i], nli);
    colL[32 * i] = x;
  }
  return !(li.x == li.x) || !(li.y == li.y);
}
template <int NT, int CNT, bool FROM_G>
struct GmColumns {
  __device__ static __forceinline__ bool run(const double2* Mf, double2* Lf, double* xb, int nte) {
    const bool b = gm_column<NT, CNT, FROM_G>(Mf, Lf, xb, nte);
    return GmColumns<NT, CNT - 1, FROM_G>::run(Mf, Lf, xb, nte) || b;
  }
};
template <int NT, bool FROM_G>
struct GmColumns<NT, 0, FROM_G> {
  __device__ static __forceinline__ bool run(const double2*, double2*, double*, int) { return false; }
};

// Factor M = L L' on NT x NT tiles.  FROM_G: lower tiles of M read from the workspace Mg (L2);
// !FROM_G: M's lower tiles are already in Ls (diagonal tiles full symmetric).  On exit Ls holds
// L_IJ (I > J) and inv(L_JJ) in the diagonal slots.  Returns false on a non-positive pivot.
template <int NT, bool FROM_G>
static __device__ __noinline__ bool gm_factor(const double* Mg, double* Ls, double* xb, int nte) {
  const int lane = threadIdx.x & 31;
  double2* Lf = reinterpret_cast<double2*>(Ls) + lane;
  const double2* Mf = reinterpret_cast<const double2*>(Mg) + lane;
  const bool bad = GmColumns<NT, NT, FROM_G>::run(Mf, Lf, xb, nte);
  const bool ok = !__any_sync(0xffffffffu, bad);
  __syncwarp();
  return ok;
}

// ---- the same factorization spread over the NW warps of a CTA (one QP per CTA) ----
// Warp W owns the tiles i = W, W + NW, ... of every tile column; all warps meet at two CTA barriers
// per column (inverse of the diagonal tile published; column final).  W is a template parameter so
// that accumulator indices stay static; a helper warp whose share of a late column is empty only
// passes the barriers.  Everything else of the solve stays on warp 0.
// Tile ownership inside a column.  Round-robin from the diagonal down (tile i -> warp i % NW) leaves warp 0 with
// its share of the trailing update IN FRONT of the serial diagonal tile; with kDiagWarp (classes whose factor is
// in shared memory) warp 0 owns ONLY the diagonal tile — its chain is one split-accumulator product row plus
// wm_diag8 — and the sub-diagonal tiles go round-robin over warps 1 .. NW-1, whose updates run beside it.
#ifndef LMPC_DIAG_WARP
#define LMPC_DIAG_WARP 1
#endif
template <int NT, int NW>
__host__ __device__ constexpr bool diag_warp() { return LMPC_DIAG_WARP != 0 && NW >= 2 && NT <= 24; }
// With eight warps two warps share a scheduler (warp w runs on scheduler w % 4) and the FP64 tensor pipe issues one
// DMMA per 16 cycles per scheduler, in order: the dependent DMMAs of wm_diag8 on warp 0 queue behind the update
// stream of warp 4.  LMPC_DIAG_ALONE leaves warp 4 out of the factorization, so the diagonal warp has its pipe alone.
#ifndef LMPC_DIAG_ALONE
#define LMPC_DIAG_ALONE 1
#endif
template <int NT, int NW>
__host__ __device__ constexpr bool diag_alone() { return LMPC_DIAG_ALONE != 0 && diag_warp<NT, NW>() && NW == 8; }
template <int NT, int NW>
__host__ __device__ constexpr int mw_workers() { return diag_alone<NT, NW>() ? 6 : NW - 1; }  // warps that share the sub-diagonal tiles
template <int NT, int NW, int W>
__host__ __device__ constexpr int mw_worker() {  // index of warp W among those workers, -1: none
  return diag_alone<NT, NW>() ? (W == 0 || W == 4 ? -1 : (W < 4 ? W - 1 : W - 2)) : W - 1;
}
template <int NT, int NW, int W>
__host__ __device__ constexpr int mw_tile(int m) {  // m-th tile (offset from the diagonal) of warp W
  return diag_warp<NT, NW>() ? (W == 0 ? 0 : 1 + mw_worker<NT, NW, W>() + mw_workers<NT, NW>() * m) : W + NW * m;
}
template <int NT, int NW, int W>
__host__ __device__ constexpr int mw_count(int CNT) {  // tiles of a CNT-tile column owned by warp W
  return diag_warp<NT, NW>()
             ? (W == 0 ? 1
                       : (mw_worker<NT, NW, W>() >= 0 && CNT - 1 > mw_worker<NT, NW, W>()
                              ? (CNT - 1 - mw_worker<NT, NW, W>() + mw_workers<NT, NW>() - 1) / mw_workers<NT, NW>()
                              : 0))
             : (CNT > W ? (CNT - W + NW - 1) / NW : 0);
}
template <int NT, int CNT, bool FROM_G, int NW, int W>
__device__ __forceinline__ bool gm_column_mw(const double2* Mf, double2* Lf, double* xb, int nte) {
  using G = TileGeom<NT>;
  constexpr int J = NT - CNT;
  if (J >= nte) return false;  // CTA-uniform: every warp skips the padded columns and their barriers
  constexpr int CW = mw_count<NT, NW, W>(CNT);  // tiles of this column owned by warp W
  double2* colL = Lf + 32 * G::lcol(J);
  const double2* colM = Mf + 32 * (J * NT + J);
  double2 acc[CW > 0 ? CW : 1], mv[CW > 0 ? CW : 1];  // (late subtraction of M: see gm_column)
#pragma unroll
  for (int m = 0; m < CW; ++m) {
    const int i = mw_tile<NT, NW, W>(m);
    mv[m] = FROM_G ? __ldcg(colM + 32 * NT * i) : colL[32 * i];
    acc[m] = kLateM && FROM_G && NT <= 24 ? make_double2(0.0, 0.0) : make_double2(-mv[m].x, -mv[m].y);
  }
  if (J > 0 && CW > 0) {
    const double2* p = Lf + 32 * J;
    if (kLateM && CW == 1) {
      // a single tile per warp (the late, narrow columns): the two k-chunks of every product run as separate
      // accumulation chains, which halves the dependent DMMA chain (26 instead of 52 cycles per K)
      double2 acb = make_double2(0.0, 0.0);
#pragma unroll 1
      for (int K = 0; K < J; ++K) {
        const double2 bj = p[0];
        const double2 a = p[32 * mw_tile<NT, NW, W>(0)];
        dmma(acc[0], a.x, bj.x);
        dmma(acb, a.y, bj.y);
        p += 32 * (NT - K - 1);
      }
      acc[0].x += acb.x; acc[0].y += acb.y;
    } else {
#pragma unroll 1
      for (int K = 0; K < J; ++K) {  // acc[m] += L_{J+i,K} L_{J,K}'
        const double2 bj = p[0];
#pragma unroll
        for (int m = 0; m < CW; ++m) dmma_xyt(acc[m], p[32 * mw_tile<NT, NW, W>(m)], bj);
        p += 32 * (NT - K - 1);
      }
    }
  }
  if (kLateM && FROM_G && NT <= 24) {  // (no gain where the factor itself streams from L2)
#pragma unroll
    for (int m = 0; m < CW; ++m) { acc[m].x -= mv[m].x; acc[m].y -= mv[m].y; }
  }
  bool bad = false;
  if (W == 0) {
    const double2 li0 = wm_diag8u(-acc[0].x, -acc[0].y, xb);
    colL[0] = li0;
    bad = !(li0.x == li0.x) || !(li0.y == li0.y);
  }
  __syncthreads();  // inv(L_JJ) visible to every warp
  if (CW > (W == 0 ? 1 : 0)) {
    const double2 li = colL[0];
    const double2 nli = make_double2(-li.x, -li.y);
#pragma unroll
    for (int m = (W == 0 ? 1 : 0); m < CW; ++m) {
      double2 x = make_double2(0.0, 0.0);
      dmma_xyt(x, acc[m], nli);
      colL[32 * mw_tile<NT, NW, W>(m)] = x;
    }
  }
  __syncthreads();  // column J final
  return bad;
}
template <int NT, int CNT, bool FROM_G, int NW, int W>
struct GmColumnsMW {
  __device__ static __forceinline__ bool run(const double2* Mf, double2* Lf, double* xb, int nte) {
    const bool b = gm_column_mw<NT, CNT, FROM_G, NW, W>(Mf, Lf, xb, nte);
    return GmColumnsMW<NT, CNT - 1, FROM_G, NW, W>::run(Mf, Lf, xb, nte) || b;
  }
};
template <int NT, bool FROM_G, int NW, int W>
struct GmColumnsMW<NT, 0, FROM_G, NW, W> {
  __device__ static __forceinline__ bool run(const double2*, double2*, double*, int) { return false; }
};
template <int NT, bool FROM_G, int NW, int W>
static __device__ __noinline__ bool gm_factor_mw_w(const double* Mg, double* Ls, double* xb, int nte) {
  const int lane = threadIdx.x & 31;
  double2* Lf = reinterpret_cast<double2*>(Ls) + lane;
  const double2* Mf = reinterpret_cast<const double2*>(Mg) + lane;
  const bool bad = GmColumnsMW<NT, NT, FROM_G, NW, W>::run(Mf, Lf, xb, nte);
  return !__any_sync(0xffffffffu, bad);
}
// Entry for warp `w` of the CTA (every warp of the CTA must call it; only warp 0's result counts).
template <int NT, bool FROM_G, int NW>
__device__ __forceinline__ bool gm_factor_mw(const double* Mg, double* Ls, double* xb, int w, int nte) {
  if (NW >= 2 && w == 1) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 2 ? 1 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 4 && w == 2) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 4 ? 2 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 4 && w == 3) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 4 ? 3 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 8 && w == 4) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 8 ? 4 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 8 && w == 5) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 8 ? 5 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 8 && w == 6) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 8 ? 6 : 0)>(Mg, Ls, xb, nte);
  if (NW >= 8 && w == 7) return gm_factor_mw_w<NT, FROM_G, NW, (NW >= 8 ? 7 : 0)>(Mg, Ls, xb, nte);
  return gm_factor_mw_w<NT, FROM_G, NW, 0>(Mg, Ls, xb, nte);
}
// Job mailbox of the CTA: 0 exit, 1 factor from the L2 workspace, 2 factor in place, 3 Hessian
// blocks, 4 reduced matrix of the polish (lmpc_tile_solve.cuh).
__device__ __forceinline__ int* tile_job() {
  __shared__ int job;
  return &job;
}
__device__ __forceinline__ int* tile_job_nte() {
  __shared__ int nte;
  return &nte;
}
// Solve L L' x = b in place on the (8 NT)-vector x in shared memory.  Row-oriented block
// substitution in t-layout; the solved blocks are read back from x itself (every lane re-reads the
// pair it wrote, so no barrier is needed inside).  Caller syncs before and after.
template <int NT>
__device__ __forceinline__ void gm_solve_impl(const double* Ls, double* x, const int nte) {
  using G = TileGeom<NT>;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const double2* Lf = reinterpret_cast<const double2*>(Ls) + lane;
  double2* xt = reinterpret_cast<double2*>(x) + t;  // block K pair at xt[4 K]
  // forward: y_J = inv(L_JJ) (b_J - sum_{K<J} L_JK y_K)
#pragma unroll 1
  for (int J = 0; J < nte; ++J) {  // blocks beyond the problem's own tile count are identity with zero right-hand sides
    // NT > 24 (factor in the CTA's L2 slot): eight independent accumulators, i.e. four 400-cycle loads in
    // flight per dependent chain (all-gait H=30 batch +17 %); the shared-memory classes keep the two-chain
    // loop, which measured faster there (C3 624k vs 554k, trot H=30 139k vs 127k)
    double pa[4] = {0.0, 0.0, 0.0, 0.0}, pb[4] = {0.0, 0.0, 0.0, 0.0};
    const double2* p = Lf + 32 * J;  // tile (J, 0); tile (J, K+1) is NT - K - 1 tiles further
    int K = 0;
    for (; NT > 24 && K + 4 <= J; K += 4) {  // (the factor of these classes lives in L2: four loads in flight per chain)
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const double2 l = *p;
        const double2 y = xt[4 * (K + u)];
        pa[u] = fma(l.x, y.x, pa[u]);
        pb[u] = fma(l.y, y.y, pb[u]);
        p += 32 * (NT - (K + u) - 1);
      }
    }
#pragma unroll 4
    for (; K < J; ++K) {
      const double2 l = *p;
      const double2 y = xt[4 * K];
      pa[0] = fma(l.x, y.x, pa[0]);
      pb[0] = fma(l.y, y.y, pb[0]);
      p += 32 * (NT - K - 1);
    }
    const double p0 = (pa[0] + pa[1]) + (pa[2] + pa[3]), p1 = (pb[0] + pb[1]) + (pb[2] + pb[3]);
    double2 r = xt[4 * J];
    if (J > 0) {
      const double2 s = quad_reduce_t(p0 + p1);
      r.x -= s.x; r.y -= s.y;
    }
    const double2 li = Lf[32 * G::lcol(J)];
    xt[4 * J] = quad_reduce_t(fma(li.x, r.x, li.y * r.y));
  }
  // backward: x_J = inv(L_JJ)' (y_J - sum_{I>J} L_IJ' x_I)
#pragma unroll 1
  for (int J = nte - 1; J >= 0; --J) {
    const double* ct = Ls + 64 * G::lcol(J) + 16 * t + g;  // transposed reads of column J: tile i at ct[64 i], ct[64 i + 8]
    const double* cs = Ls + 64 * G::lcol(J);               // ... and the two-way-conflict form for the off-diagonal tiles
    double pa[4] = {0.0, 0.0, 0.0, 0.0}, pb[4] = {0.0, 0.0, 0.0, 0.0};
    int i = 1;
    for (; NT > 24 && i + 4 <= nte - J; i += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const double2 xi = vec_ld_tr<kTransS && NT <= 24>(x, J + i + u, t);
        const double2 c = tile_ld_tr<kTransS && NT <= 24>(cs + 64 * (i + u), g, t);
        pa[u] = fma(c.x, xi.x, pa[u]);
        pb[u] = fma(c.y, xi.y, pb[u]);
      }
    }
#pragma unroll 4
    for (; i < nte - J; ++i) {
      const double2 xi = vec_ld_tr<kTransS && NT <= 24>(x, J + i, t);
      const double2 c = tile_ld_tr<kTransS && NT <= 24>(cs + 64 * i, g, t);
      pa[0] = fma(c.x, xi.x, pa[0]);
      pb[0] = fma(c.y, xi.y, pb[0]);
    }
    const double p0 = (pa[0] + pa[1]) + (pa[2] + pa[3]), p1 = (pb[0] + pb[1]) + (pb[2] + pb[3]);
    double2 r = xt[4 * J];
    if (J < nte - 1) {
      const double2 s = quad_reduce_t(p0 + p1);
      r.x -= s.x; r.y -= s.y;
    }
    xt[4 * J] = quad_reduce_t(fma(ct[0], r.x, ct[8] * r.y));
  }
}
// nte = tiles this problem fills (<= NT).  The full-size case keeps its compile-time trip counts.
template <int NT>
static __device__ __noinline__ void gm_solve(const double* Ls, double* x, int nte) {
  if (nte == NT) gm_solve_impl<NT>(Ls, x, NT);
  else gm_solve_impl<NT>(Ls, x, nte);
}

// ---- the same solve spread over the NW warps of a CTA: blocked substitution with BS tile rows per
// block and EXPLICITLY INVERTED diagonal blocks.  A plain blocked substitution leaves the BS x BS
// triangle of every block to one warp (BS dependent rows, each two DMMA round trips) while the other
// warps wait at the barrier: at NT = 24 with eight warps that was 42 % of all stall samples
// (profiles/r02_v8_ncu_stalls_solve_tile24_kernel.txt).  So right after the factorization every
// diagonal block L_bb (BS x BS tiles, lower triangular) is replaced IN PLACE by its inverse — one warp
// per block, ~BS^3/6 tile products, no barrier inside — and the triangle becomes BS independent rows of
// a small matrix-vector product, one per warp.
//
// Storage after gm_invert_blocks_mw: tiles (I, K) outside the diagonal blocks still hold L_IK; the
// diagonal slots still hold inv(L_II); a slot (I, K), I > K, INSIDE a diagonal block holds the
// TRANSPOSE of tile (I, K) of inv(L_bb) (that is what the DMMA product X * Y' delivers without a
// layout change: V_KI = -(sum_{k=K}^{I-1} V_Kk L_Ik') inv(L_II)',  V_KK = inv(L_KK)').
#ifndef LMPC_TILE_BS
#define LMPC_TILE_BS 4
#endif
constexpr int kTileSolveBS = LMPC_TILE_BS;

template <int NT, int BS>
__device__ __forceinline__ void gm_block_inverse(double* Ls, int d0, int d1) {
  using G = TileGeom<NT>;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  double2* Lf = reinterpret_cast<double2*>(Ls) + lane;
#pragma unroll
  for (int a = 0; a < BS - 1; ++a) {  // tile column K = d0 + a of the block, left to right (columns > K still hold L)
    const int K = d0 + a;
    if (K < d1 - 1) {
      double2 v[BS];  // v[c] = V_{K, d0 + c}
      v[a] = tile_ld_t(Ls + 64 * G::lcol(K), g, t);
#pragma unroll
      for (int b = a + 1; b < BS; ++b) {
        const int I = d0 + b;
        if (I < d1) {
          double2 acc = make_double2(0.0, 0.0);
#pragma unroll
          for (int c = a; c < b; ++c) dmma_xyt(acc, v[c], Lf[32 * (G::lcol(d0 + c) + I - (d0 + c))]);
          double2 x = make_double2(0.0, 0.0);
          dmma_xyt(x, make_double2(-acc.x, -acc.y), Lf[32 * G::lcol(I)]);
          v[b] = x;
          Lf[32 * (G::lcol(K) + I - K)] = x;
        }
      }
    }
  }
}
// Every warp of the CTA calls it right after gm_factor_mw (whose last barrier made the factor visible).
template <int NT, int NW>
__device__ __forceinline__ void gm_invert_blocks_mw(double* Ls, int nte, int w) {
  constexpr int BS = kTileSolveBS;
  for (int d0 = BS * w; d0 < nte; d0 += BS * NW) gm_block_inverse<NT, BS>(Ls, d0, d0 + BS < nte ? d0 + BS : nte);
  __syncthreads();
}

// L L' x = b on a factor prepared by gm_invert_blocks_mw; x and the scratch vector tt (8 NT doubles
// each) in shared memory, t-layout.  Forward: y goes to tt, the running right-hand side stays in x;
// backward: the running right-hand side is tt, the solution lands in x.  Two CTA barriers per block
// and direction; every warp of the CTA must call it.
template <int NT, int NW>
__device__ __forceinline__ void gm_solve_mw_impl(const double* Ls, double* x, double* tt, const int nte, const int w) {
  using G = TileGeom<NT>;
  constexpr int BS = kTileSolveBS;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const double2* Lf = reinterpret_cast<const double2*>(Ls) + lane;
  double2* xt = reinterpret_cast<double2*>(x) + t;
  double2* yt = reinterpret_cast<double2*>(tt) + t;
  // forward
  for (int d0 = 0; d0 < nte; d0 += BS) {
    const int d1 = d0 + BS < nte ? d0 + BS : nte;
    for (int J = d0 + w; J < d1; J += NW) {  // y_J = inv(L_bb)[J, :] r   (row J of the block inverse: transposed slots)
      const double2 li = Lf[32 * G::lcol(J)];
      const double2 rj = xt[4 * J];
      double p0 = li.x * rj.x, p1 = li.y * rj.y;
#pragma unroll
      for (int u = 0; u < BS - 1; ++u) {
        const int K = d0 + u;
        if (K < J) {
          const double2 c = tile_ld_tr<kTransS && NT <= 24>(Ls + 64 * (G::lcol(K) + J - K), g, t);
          const double2 r = vec_ld_tr<kTransS && NT <= 24>(x, K, t);
          p0 = fma(c.x, r.x, p0);
          p1 = fma(c.y, r.y, p1);
        }
      }
      yt[4 * J] = quad_reduce_t(p0 + p1);
    }
    __syncthreads();
    for (int I = d1 + w; I < nte; I += NW) {  // r_I -= sum_{K in block} L_IK y_K
      double p0 = 0.0, p1 = 0.0;
#pragma unroll
      for (int u = 0; u < BS; ++u) {
        const int K = d0 + u;
        if (K < d1) {
          const double2 l = Lf[32 * (G::lcol(K) + I - K)];
          const double2 y = yt[4 * K];
          p0 = fma(l.x, y.x, p0);
          p1 = fma(l.y, y.y, p1);
        }
      }
      const double2 s = quad_reduce_t(p0 + p1);
      double2 r = xt[4 * I];
      r.x -= s.x; r.y -= s.y;
      xt[4 * I] = r;
    }
    __syncthreads();
  }
  // backward: blocks from the bottom; right-hand side in tt
  const int dlast = ((nte - 1) / BS) * BS;
  for (int d0 = dlast; d0 >= 0; d0 -= BS) {
    const int d1 = d0 + BS < nte ? d0 + BS : nte;
    for (int J = d0 + w; J < d1; J += NW) {  // x_J = inv(L_bb)'[J, :] y   (column J of the block inverse: direct slots)
      const double2 c = tile_ld_tr<kTransS && NT <= 24>(Ls + 64 * G::lcol(J), g, t);
      const double2 yj = vec_ld_tr<kTransS && NT <= 24>(tt, J, t);
      double p0 = c.x * yj.x, p1 = c.y * yj.y;
#pragma unroll
      for (int u = 1; u < BS; ++u) {
        const int I = d0 + u;
        if (I > J && I < d1) {
          const double2 l = Lf[32 * (G::lcol(J) + I - J)];
          const double2 y = yt[4 * I];
          p0 = fma(l.x, y.x, p0);
          p1 = fma(l.y, y.y, p1);
        }
      }
      xt[4 * J] = quad_reduce_t(p0 + p1);
    }
    __syncthreads();
    for (int J = w; J < d0; J += NW) {  // y_J -= sum_{I in block} L_IJ' x_I
      const double* cs = Ls + 64 * G::lcol(J);
      double p0 = 0.0, p1 = 0.0;
#pragma unroll
      for (int u = 0; u < BS; ++u) {
        const int I = d0 + u;
        if (I < d1) {
          const double2 xi = vec_ld_tr<kTransS && NT <= 24>(x, I, t);
          const double2 c = tile_ld_tr<kTransS && NT <= 24>(cs + 64 * (I - J), g, t);
          p0 = fma(c.x, xi.x, p0);
          p1 = fma(c.y, xi.y, p1);
        }
      }
      const double2 s = quad_reduce_t(p0 + p1);
      double2 r = yt[4 * J];
      r.x -= s.x; r.y -= s.y;
      yt[4 * J] = r;
    }
    __syncthreads();
  }
}
template <int NT, int NW>
static __device__ __noinline__ void gm_solve_mw(const double* Ls, double* x, double* tt, int nte, int w) {
  gm_solve_mw_impl<NT, NW>(Ls, x, tt, nte, w);
}

// y = H x for the full-symmetric workspace (global, L2); x, y (8 NT)-vectors in shared memory.
// Tile rows I = i0, i0 + di, ...: one warp does them all (i0 = 0, di = 1) or the warps of the CTA
// take every NW-th row (job 5).
template <int NT>
__device__ __forceinline__ void gm_symv_impl(const double* Hg, const double* x, double* y, const int nte, int i0, int di) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const double2* Hf = reinterpret_cast<const double2*>(Hg) + lane;
  const double2* xt = reinterpret_cast<const double2*>(x) + t;
#pragma unroll 1
  for (int I = i0; I < nte; I += di) {  // rows / columns beyond nte are identity padding times zeros: y stays 0 there
    double p0 = 0.0, p1 = 0.0;
    const double2* row = Hf + 32 * (I * NT);
#pragma unroll 8
    for (int J = 0; J < nte; ++J) {
      const double2 h = __ldcg(row + 32 * J);
      const double2 v = xt[4 * J];
      p0 = fma(h.x, v.x, p0);
      p1 = fma(h.y, v.y, p1);
    }
    const double2 s = quad_reduce_t(p0 + p1);
    if (g == 0) *reinterpret_cast<double2*>(y + 8 * I + 2 * t) = s;
  }
}
template <int NT>
static __device__ __noinline__ void gm_symv(const double* Hg, const double* x, double* y, int nte, int i0 = 0, int di = 1) {
  if (nte == NT) gm_symv_impl<NT>(Hg, x, y, NT, i0, di);
  else gm_symv_impl<NT>(Hg, x, y, nte, i0, di);
}

}  // namespace lmpc
