// lmpc_wmma.cuh — dense FP64 kernels of the one-warp-per-QP path (reduced dimension n <= 64) on
// the FP64 tensor path (DMMA, mma.sync.aligned.m8n8k4.f64).
//
// FRAGMENT LAYOUT.  The padded 64x64 matrices are cut into 8x8 tiles and every tile is kept in the
// accumulator ("C") layout of the m8n8k4 DMMA:  lane = 4*g + t  (g = lane>>2, t = lane&3) owns
// the two adjacent elements  tile[g][2t], tile[g][2t+1]  as one double2.  A tile is 32 x 16 B =
// 512 contiguous bytes ("fragment order"), so a warp moves a tile with ONE conflict-free 128-bit
// load or store per lane, and a lane only ever re-reads the 16 bytes it wrote itself: the shared
// memory that holds the Cholesky factor is a private extension of the register file and the
// factorization needs no warp barrier at all.
//
// K-PERMUTATION.  A DMMA sums over 4 k-slots and slot t of operand A meets slot t of operand B.
// With the slot->k map  k = 2t (first DMMA) and k = 2t+1 (second DMMA), a C-layout register pair
// of tile X is DIRECTLY the A operand X[g][k] (.x then .y) and DIRECTLY the B operand
// B[k][n=g] = Y[g][k] of the product X * Y^T.  Every product the Cholesky needs has that shape
// (L_IK * L_JK^T, A_IJ * inv(L_JJ)^T), so tiles flow accumulator -> operand with no shuffle.
//
// Factorization (left-looking by tile columns): per tile column J the 8-J accumulator tiles are
// initialised with -(H + W), updated with 2 DMMAs per earlier tile column, the diagonal tile is
// factored and inverted by row elimination with quad/oct shuffles (wm_diag8), and the panel is
// L_IJ = A_IJ * inv(L_JJ)^T, two more DMMAs per tile.  Diagonal slots store inv(L_JJ)^T, which
// makes both triangular solves 8x8 block mat-vecs with no layout conversion:
//   "t-layout": lane holds v[8K+2t], v[8K+2t+1];  "g-layout": lane holds v[8K+g].
//   tile * t-layout vector -> (sum over t: 2 shuffle stages) -> g-layout
//   tile^T * g-layout vector -> (sum over g: 3 shuffle stages) -> t-layout
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "lmpc_types.cuh"

namespace lmpc {

constexpr int WN = 64;            // padded dimension of the warp path
constexpr int WNT = 8;            // tiles per dimension
constexpr int WNBMAX = 21;        // max stance blocks (3*21 = 63 <= 64)
constexpr int WLTILES = 36;       // lower-triangular tiles of the factor
constexpr int WLS = WLTILES * 64; // doubles of factor storage (18 KB)
constexpr int WHS = 64 * 64;      // doubles of the full-symmetric Hessian workspace (all 64 tiles)

__host__ __device__ constexpr int tri(int r, int c) { return (r * (r + 1)) / 2 + c; }

// offset (in doubles) of element (r, c) inside a fragment-ordered tile array with `tile` index
__device__ __forceinline__ int frag_off(int tile, int r, int c) {
  return tile * 64 + (((r & 7) << 2) + ((c & 7) >> 1)) * 2 + (c & 1);
}
// full-symmetric Hessian workspace: tile (I, J) at index 8I + J
__device__ __forceinline__ int hfull_off(int r, int c) { return frag_off(((r >> 3) << 3) + (c >> 3), r, c); }
// factor storage: lower tiles only, tile (I, J), I >= J at index tri(I, J)
__device__ __forceinline__ int lfac_off(int r, int c) { return frag_off(tri(r >> 3, c >> 3), r, c); }

// c += A-slot a x B-slot b  (one k-chunk of an 8x8x4 FP64 tensor-core product)
__device__ __forceinline__ void dmma(double2& c, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c.x), "+d"(c.y)
      : "d"(a), "d"(b));
}
// Sum over the warp on the tensor pipe: ones x v gives the quad sums (lane (g,t) holds those of
// quads 2t, 2t+1), a second ones-product adds the eight quads.  Every lane gets the total.
__device__ __forceinline__ double wsum(double v) {
  double2 q = make_double2(0.0, 0.0);
  dmma(q, 1.0, v);
  double2 s = make_double2(0.0, 0.0);
  dmma(s, 1.0, q.x + q.y);
  return s.x;
}
// Max over the warp of NON-NEGATIVE doubles (IEEE order == unsigned order of the bit pattern; a
// NaN compares above everything and propagates): two 32-bit redux.sync steps.
__device__ __forceinline__ double wmax(double v) {
  v = fabs(v);  // callers pass non-negative values; this also folds -0.0
  const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
  const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
  const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
  return __hiloint2double((int)mh, (int)ml);
}
// sum over the 4 lanes of a quad (over t)
__device__ __forceinline__ double qsum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}
// sum over the 8 quads (over g)
__device__ __forceinline__ double gsum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  return v;
}

// 1/a and 1/sqrt(a) from the 20-bit hardware seed + two Newton steps (full double accuracy up to
// the last ulp; no slow-path branch, a must be a positive normal number).
__device__ __forceinline__ double frcp(double a) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
  double e = fma(-a, x, 1.0);
  x = fma(x, e, x);
  e = fma(-a, x, 1.0);
  return fma(x, e, x);
}
__device__ __forceinline__ double frsqrt(double a) {
  double x;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
  double e = fma(-a * x, x, 1.0);
  x = fma(0.5 * x, e, x);
  e = fma(-a * x, x, 1.0);
  return fma(0.5 * x, e, x);
}

// c += X * Y^T for two C-layout tiles (both k-chunks)
__device__ __forceinline__ void dmma_xyt(double2& c, const double2& x, const double2& y) {
  dmma(c, x.x, y.x);
  dmma(c, x.y, y.y);
}
// Wider PTX shapes, same convention (fragment layouts verified on a B200 by tools/probe/mma_probe.cu).
// NOT used on the hot path: ptxas lowers m16n8k8 / m16n8k16 to 4 / 8 DMMA.8x8x4 on sm_100a (the only
// FP64 tensor SASS shape), so they save no issue slots; kept for reference.
// [c1; c2] += [X1; X2] * Y^T (m16n8k8) ...
__device__ __forceinline__ void dmma2_xyt(double2& c1, double2& c2, const double2& x1, const double2& x2,
                                          const double2& y) {
  asm("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+d"(c1.x), "+d"(c1.y), "+d"(c2.x), "+d"(c2.y)
      : "d"(x1.x), "d"(x2.x), "d"(x1.y), "d"(x2.y), "d"(y.x), "d"(y.y));
}
// ... and [c1; c2] += [X1a X1b; X2a X2b] * [Ya Yb]^T (two k-tiles) in ONE m16n8k16 instruction.
__device__ __forceinline__ void dmma4_xyt(double2& c1, double2& c2, const double2& x1a, const double2& x2a,
                                          const double2& x1b, const double2& x2b, const double2& ya,
                                          const double2& yb) {
  asm("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, "
      "{%12,%13,%14,%15}, {%0,%1,%2,%3};"
      : "+d"(c1.x), "+d"(c1.y), "+d"(c2.x), "+d"(c2.y)
      : "d"(x1a.x), "d"(x2a.x), "d"(x1a.y), "d"(x2a.y), "d"(x1b.x), "d"(x2b.x), "d"(x1b.y), "d"(x2b.y),
        "d"(ya.x), "d"(ya.y), "d"(yb.x), "d"(yb.y));
}
// t-layout result of  sum over the quad of the lane-local partial dot products p  (see "vectors")
__device__ __forceinline__ double2 quad_reduce_t(double p) {
  double2 c = make_double2(0.0, 0.0);
  dmma(c, 1.0, p);
  return c;
}


// ---- vectors.  A 64-vector block v[8K..8K+7] lives in registers in "t-layout": lane (g, t) holds
// the pair v[8K+2t], v[8K+2t+1], replicated over g.  tile * v:  every lane multiplies its two tile
// elements (row g, columns 2t, 2t+1) with its pair (two DFMAs, partials of several tiles
// accumulate first), then ONE DMMA  ones x p  adds the four partials of each row AND returns the
// result in t-layout (C[.][n] = sum_k p(lane (n,k))).  tile' * v works the same on a transposed
// read (two LDS.64).  Triangular solves and H*x therefore need no shuffle and no layout change.

// transposed read of a fragment-ordered tile: returns {T[2t][g], T[2t+1][g]}
__device__ __forceinline__ double2 tile_ld_t(const double* tile, int g, int t) {
  const double* p = tile + 16 * t + g;
  return make_double2(p[0], p[8]);
}

// The same transposed product with HALF the bank conflicts.  A tile row is 64 bytes, so rows of equal parity sit in
// the same 16 banks: tile_ld_t (rows 2t and 2t+1 of column g, one LDS.64 each) makes the four lanes of a quad hit the
// same banks — a four-way conflict, 8 wavefronts per load.  Reading rows t and t+4 instead puts two parities into
// every load (two-way, 4 wavefronts); the lane then multiplies with the vector elements 8K+t and 8K+t+4 and the quad
// sum over t is the same column-g result.  For vectors that live in shared memory.
#ifndef LMPC_TRANS_S
#define LMPC_TRANS_S 1
#endif
constexpr bool kTransS = LMPC_TRANS_S != 0;
// S (the factor is in shared memory): returns {T[t][g], T[t+4][g]}; else {T[2t][g], T[2t+1][g]} (a factor in the
// global slot has no banks, and the rows-t/t+4 pattern measured 5 % slower on class 36)
template <bool S>
__device__ __forceinline__ double2 tile_ld_tr(const double* tile, int g, int t) {
  const double* p = tile + (S ? 8 * t : 16 * t) + g;
  return make_double2(p[0], p[S ? 32 : 8]);
}
// the matching pair of block K of the vector x: {x[8K+t], x[8K+t+4]} or {x[8K+2t], x[8K+2t+1]}
template <bool S>
__device__ __forceinline__ double2 vec_ld_tr(const double* x, int K, int t) {
  if (S) return make_double2(x[8 * K + t], x[8 * K + t + 4]);
  return *reinterpret_cast<const double2*>(x + 8 * K + 2 * t);
}

// Factor the SPD 8x8 tile A (C-layout, full symmetric) and return G = inv(L) (lower triangular,
// C-layout; G A G' = I).  Elimination by 2x2 BLOCK pivots (rows 2s, 2s+1 at step s = 0, 1, 2):
//   rows g > 2s+1:  row_g -= [A[g][2s] A[g][2s+1]] inv(P_s) [row_2s; row_2s+1],   P_s = A[2s..2s+1][2s..2s+1],
// applied to A and to E = I, gives E = inv(L_b) of the block factorization A = L_b D_b L_b'; then
// G = blockdiag(inv chol(P_s)) E.  Two pivots cost ONE reciprocal (of det P_s) on the sequential
// chain instead of two, and the four 2x2 Cholesky inverses (two rsqrt each) run side by side at
// the end.  The pivot rows travel through the shared-memory buffer `xb` (64 doubles: per row the A
// part then the E part, two rows per step, double-buffered by step parity); the trailing matrix
// stays symmetric, so A[g][c] is read as A[c][g] from the broadcast row.  A non-positive pivot
// block poisons its rows with NaN.
template <bool UNROLL>
__device__ __forceinline__ double2 wm_diag8_impl(double ax, double ay, double* xb) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  double ex = (g == 2 * t) ? 1.0 : 0.0, ey = (g == 2 * t + 1) ? 1.0 : 0.0;
  double2* xs = reinterpret_cast<double2*>(xb) + t;   // this lane's column pair inside a row
  const double* xg = xb + g;
#pragma unroll(UNROLL ? 3 : 1)
  for (int s = 0; s < 3; ++s) {
    const int c = 2 * s;
    const int o = 32 * (s & 1);        // doubles: row c at o (A) / o+8 (E), row c+1 at o+16 / o+24
    if ((g >> 1) == s) {
      const int q = (o >> 1) + 8 * (g & 1);
      xs[q] = make_double2(ax, ay);
      xs[q + 4] = make_double2(ex, ey);
    }
    __syncwarp();
    const double2 ab = *reinterpret_cast<const double2*>(xb + o + c);            // A[c][c], A[c][c+1]
    const double cc = xb[o + 16 + c + 1];                                        // A[c+1][c+1]
    const double rdet = frcp(fma(ab.x, cc, -ab.y * ab.y));
    const double p0 = xg[o], p1 = xg[o + 16];                                    // A[g][c], A[g][c+1]
    double m0 = fma(p0, cc, -p1 * ab.y) * rdet;
    double m1 = fma(p1, ab.x, -p0 * ab.y) * rdet;
    const bool below = g > c + 1;
    m0 = below ? m0 : 0.0;
    m1 = below ? m1 : 0.0;
    const double2 r0 = xs[o >> 1], f0 = xs[(o >> 1) + 4], r1 = xs[(o >> 1) + 8], f1 = xs[(o >> 1) + 12];
    ax = fma(-m1, r1.x, fma(-m0, r0.x, ax)); ay = fma(-m1, r1.y, fma(-m0, r0.y, ay));
    ex = fma(-m1, f1.x, fma(-m0, f0.x, ex)); ey = fma(-m1, f1.y, fma(-m0, f0.y, ey));
  }
  // 2x2 Cholesky inverse of every pivot block: rows 2s, 2s+1 are final after step s-1.
  //   a = A[2s][2s], b = A[2s][2s+1] sit in lane (g = 2s, t = s); d = A[2s+1][2s+1] in lane (2s+1, s).y
  const int sb = g >> 1, top = 8 * sb + sb;          // lane (2 sb, sb)
  const double a = __shfl_sync(0xffffffffu, ax, top);
  const double bq = __shfl_sync(0xffffffffu, ay, top);
  const double d = __shfl_sync(0xffffffffu, ay, top + 4);
  const double ux = __shfl_up_sync(0xffffffffu, ex, 4), uy = __shfl_up_sync(0xffffffffu, ey, 4);  // E row above
  const double nan_ = __longlong_as_double(0x7ff8000000000000ll);
  const double ra = (a > 0.0) ? frsqrt(a) : nan_;    // 1 / l11
  const double ia = ra * ra;                         // 1 / a
  const double sc = fma(-bq * ia, bq, d);            // l22^2 = d - b^2 / a
  const double rs = (sc > 0.0) ? frsqrt(sc) : nan_;  // 1 / l22
  if (g & 1) {
    const double k = bq * ia;                        // l21 / l11
    return make_double2(rs * fma(-k, ux, ex), rs * fma(-k, uy, ey));
  }
  return make_double2(ra * ex, ra * ey);
}
// Tensor-pipe variant of the same elimination (the default; -DLMPC_DIAG8_SMEM selects the
// shared-memory one above).  Nothing travels through shared memory and there is no warp barrier:
//   * a step only broadcasts the three numbers of its 2x2 pivot block P_s (three SHFLs); the
//     multipliers X[g][0..1] = [A[g][2s] A[g][2s+1]] inv(P_s) are computed by the lanes that OWN
//     those two elements (t == s) from their own registers;
//   * the rank-2 updates are DMMAs whose k-slots are masked to slot s:  A -= X C'  uses the old
//     A registers as the B operand (C = A[:, 2s..2s+1], by symmetry of the trailing matrix), and
//     the inverse is accumulated TRANSPOSED, F = E':  F -= F[:, 2s..2s+1] X', so that both
//     operands are again registers the lane already holds;
//   * G' = F D' (2x2 Cholesky inverse of every pivot block, per-lane column pair) and one
//     identity-DMMA pair transposes G' into the C-layout of G.
// 16 DMMAs + 12 SHFLs per tile instead of ~400 instructions, and the F chain runs beside the A chain.
#ifndef LMPC_DIAG8_UNROLL
#define LMPC_DIAG8_UNROLL 3
#endif
constexpr int kDiag8Unroll = LMPC_DIAG8_UNROLL;
__device__ __forceinline__ double2 wm_diag8_mma(double ax, double ay) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  double fx = (g == 2 * t) ? 1.0 : 0.0, fy = (g == 2 * t + 1) ? 1.0 : 0.0;
#pragma unroll kDiag8Unroll
  for (int s = 0; s < 3; ++s) {
    const int src = 9 * s;  // lane (g = 2s, t = s)
    const double a = __shfl_sync(0xffffffffu, ax, src);
    const double b = __shfl_sync(0xffffffffu, ay, src);
    const double d = __shfl_sync(0xffffffffu, ay, src + 4);
    const double rdet = frcp(fma(a, d, -b * b));
    const double c0 = ax, c1 = ay;
    const bool act = (t == s) && (g > 2 * s + 1);
    const double x0 = act ? -(fma(c0, d, -c1 * b) * rdet) : 0.0;   // -X[g][0]
    const double x1 = act ? -(fma(c1, a, -c0 * b) * rdet) : 0.0;   // -X[g][1]
    double2 A0 = make_double2(ax, ay), A1 = make_double2(0.0, 0.0);
    dmma(A0, x0, c0);
    dmma(A1, x1, c1);
    double2 F0 = make_double2(fx, fy), F1 = make_double2(0.0, 0.0);
    dmma(F0, fx, x0);
    dmma(F1, fy, x1);
    ax = A0.x + A1.x; ay = A0.y + A1.y;
    fx = F0.x + F1.x; fy = F0.y + F1.y;
  }
  // final pivot block of this lane's column pair: rows 2t, 2t+1
  const double a = __shfl_sync(0xffffffffu, ax, 9 * t);
  const double b = __shfl_sync(0xffffffffu, ay, 9 * t);
  const double d = __shfl_sync(0xffffffffu, ay, 9 * t + 4);
  const double nan_ = __longlong_as_double(0x7ff8000000000000ll);
  // 1/l11 = rsqrt(a);  l22^2 = det / a with det = a d - b^2, so 1/l22 = rsqrt(det) sqrt(a) = rsqrt(det) a rsqrt(a):
  // the two reciprocal square roots do not depend on each other and run side by side
  const double det = fma(a, d, -b * b);
  const double ra = (a > 0.0) ? frsqrt(a) : nan_;      // 1 / l11
  const double rd = (det > 0.0) ? frsqrt(det) : nan_;
  const double ia = ra * ra;                           // 1 / a
  const double rs = rd * (a * ra);                     // 1 / l22
  const double gx = ra * fx;                           // G'[n][2t]   = G[2t][n]
  const double gy = rs * fma(-(b * ia), fx, fy);       // G'[n][2t+1] = G[2t+1][n]
  double2 G = make_double2(0.0, 0.0);
  dmma(G, (g == 2 * t) ? 1.0 : 0.0, gx);
  dmma(G, (g == 2 * t + 1) ? 1.0 : 0.0, gy);
  return G;
}
#ifdef LMPC_DIAG8_SMEM
static __device__ __noinline__ double2 wm_diag8(double ax, double ay, double* xb) { return wm_diag8_impl<true>(ax, ay, xb); }
static __device__ __noinline__ double2 wm_diag8u(double ax, double ay, double* xb) { return wm_diag8_impl<true>(ax, ay, xb); }
#else
static __device__ __noinline__ double2 wm_diag8(double ax, double ay, double*) { return wm_diag8_mma(ax, ay); }
static __device__ __noinline__ double2 wm_diag8u(double ax, double ay, double*) { return wm_diag8_mma(ax, ay); }
#endif

// Factor M = L L' (64 x 64, 8 x 8 tiles).  FROM_G: the lower tiles of M are read from the
// full-symmetric fragment-ordered workspace Mg in global memory (L2; the interior-point loop keeps
// H + C'WC there), each tile column fetched one column ahead of its use.  !FROM_G: the lower tiles
// of M are already in Ls (diagonal tiles full symmetric).  On exit Ls holds L_IJ (I > J) and
// inv(L_JJ) in the diagonal slots.  Left-looking by tile columns: J unrolled (static accumulator
// indices), the K loop rolled to keep the hot loop inside the instruction cache; for the narrow
// late columns the two k-chunks of a tile product accumulate in separate chains.
// Returns false on a non-positive pivot.
template <bool FROM_G>
static __device__ __noinline__ bool wm_factor(const double* Mg, double* Ls, double* xb) {
  const int lane = threadIdx.x & 31;
  double2* Lf = reinterpret_cast<double2*>(Ls) + lane;
  const double2* Mf = reinterpret_cast<const double2*>(Mg) + lane;
  bool bad = false;
  double2 nxt[WNT];
#pragma unroll
  for (int I = 0; I < WNT; ++I) nxt[I] = FROM_G ? __ldcg(Mf + 32 * (8 * I)) : Lf[32 * tri(I, 0)];
#pragma unroll
  for (int J = 0; J < WNT; ++J) {
    double2 acc[WNT], acb[WNT];
#pragma unroll
    for (int I = J; I < WNT; ++I) {
      acc[I] = make_double2(-nxt[I].x, -nxt[I].y);
      acb[I] = make_double2(0.0, 0.0);
    }
    if (J + 1 < WNT) {
#pragma unroll
      for (int I = J + 1; I < WNT; ++I)
        nxt[I] = FROM_G ? __ldcg(Mf + 32 * (8 * I + J + 1)) : Lf[32 * tri(I, J + 1)];
    }
    if (J > 0) {  // acc[I] += L_IK L_JK' over K < J
      const double2* rowJ = Lf + 32 * tri(J, 0);
#pragma unroll 1
      for (int K = 0; K < J; ++K) {
        const double2 bj = rowJ[32 * K];
        if (J >= 4) {
          dmma(acc[J], bj.x, bj.x);
          dmma(acb[J], bj.y, bj.y);
#pragma unroll
          for (int I = J + 1; I < WNT; ++I) {
            const double2 a = Lf[32 * (tri(I, 0) + K)];
            dmma(acc[I], a.x, bj.x);
            dmma(acb[I], a.y, bj.y);
          }
        } else {
          dmma_xyt(acc[J], bj, bj);
#pragma unroll
          for (int I = J + 1; I < WNT; ++I) dmma_xyt(acc[I], Lf[32 * (tri(I, 0) + K)], bj);
        }
      }
      if (J >= 4) {
#pragma unroll
        for (int I = J; I < WNT; ++I) { acc[I].x += acb[I].x; acc[I].y += acb[I].y; }
      }
    }
    const double2 li = wm_diag8(-acc[J].x, -acc[J].y, xb);
    bad = bad || !(li.x == li.x) || !(li.y == li.y);
    Lf[32 * tri(J, J)] = li;
    const double2 nli = make_double2(-li.x, -li.y);
#pragma unroll
    for (int I = J + 1; I < WNT; ++I) {
      double2 x = make_double2(0.0, 0.0);
      dmma_xyt(x, acc[I], nli);
      Lf[32 * tri(I, J)] = x;
    }
  }
  const bool ok = !__any_sync(0xffffffffu, bad);
  __syncwarp();  // the solves read tiles transposed, i.e. words written by other lanes
  return ok;
}

// Solve L L' x = b in place on the 64-vector `x` in shared memory.  Row-oriented block
// substitution in t-layout: lane-local DFMA partials over the tiles of a block row, one ones-DMMA
// per reduction (16 per triangular solve), straight-line code.
// The caller must __syncwarp() between writing b and this call, and again before reading x.
static __device__ __noinline__ void wm_solve(const double* Ls, double* x) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const double2* Lf = reinterpret_cast<const double2*>(Ls) + lane;
  const double* Lt = Ls + 16 * t + g;  // transposed reads: {Lt[64 T], Lt[64 T + 8]} = {T[2t][g], T[2t+1][g]}
  double2 r[WNT];
#pragma unroll
  for (int I = 0; I < WNT; ++I) r[I] = *reinterpret_cast<const double2*>(x + 8 * I + 2 * t);
  // forward: y_J = inv(L_JJ) (b_J - sum_{K<J} L_JK y_K)
#pragma unroll
  for (int J = 0; J < WNT; ++J) {
    if (J > 0) {
      double p0 = 0.0, p1 = 0.0;
#pragma unroll
      for (int K = 0; K < J; ++K) {
        const double2 l = Lf[32 * tri(J, K)];
        p0 = fma(l.x, r[K].x, p0);
        p1 = fma(l.y, r[K].y, p1);
      }
      const double2 s = quad_reduce_t(p0 + p1);
      r[J].x -= s.x; r[J].y -= s.y;
    }
    const double2 li = Lf[32 * tri(J, J)];
    r[J] = quad_reduce_t(fma(li.x, r[J].x, li.y * r[J].y));
  }
  // backward: x_J = inv(L_JJ)' (y_J - sum_{I>J} L_IJ' x_I)
#pragma unroll
  for (int J = WNT - 1; J >= 0; --J) {
    if (J < WNT - 1) {
      double p0 = 0.0, p1 = 0.0;
#pragma unroll
      for (int I = J + 1; I < WNT; ++I) {
        p0 = fma(Lt[64 * tri(I, J)], r[I].x, p0);
        p1 = fma(Lt[64 * tri(I, J) + 8], r[I].y, p1);
      }
      const double2 s = quad_reduce_t(p0 + p1);
      r[J].x -= s.x; r[J].y -= s.y;
    }
    r[J] = quad_reduce_t(fma(Lt[64 * tri(J, J)], r[J].x, Lt[64 * tri(J, J) + 8] * r[J].y));
  }
  if (g == 0) {
#pragma unroll
    for (int I = 0; I < WNT; ++I) *reinterpret_cast<double2*>(x + 8 * I + 2 * t) = r[I];
  }
}

// y = H x for the full-symmetric fragment-ordered Hessian workspace (global, L2 resident);
// x and y are 64-vectors in shared memory (x != y).  Caller syncs as for wm_solve.  Hg is
// written earlier in the same kernel: loads go through L2 (__ldcg).
static __device__ __noinline__ void wm_symv(const double* Hg, const double* x, double* y) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const double2* Hf = reinterpret_cast<const double2*>(Hg) + lane;
  double2 xt[WNT];
#pragma unroll
  for (int J = 0; J < WNT; ++J) xt[J] = *reinterpret_cast<const double2*>(x + 8 * J + 2 * t);
#pragma unroll 2
  for (int I = 0; I < WNT; ++I) {
    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
    for (int J = 0; J < WNT; J += 2) {
      const double2 h0 = __ldcg(Hf + 32 * (8 * I + J));
      const double2 h1 = __ldcg(Hf + 32 * (8 * I + J + 1));
      p0 = fma(h0.x, xt[J].x, p0);
      p1 = fma(h0.y, xt[J].y, p1);
      p2 = fma(h1.x, xt[J + 1].x, p2);
      p3 = fma(h1.y, xt[J + 1].y, p3);
    }
    const double2 s = quad_reduce_t((p0 + p1) + (p2 + p3));
    if (g == 0) *reinterpret_cast<double2*>(y + 8 * I + 2 * t) = s;
  }
}

}  // namespace lmpc
