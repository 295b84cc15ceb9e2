// lmpc_warp.cuh — fast path for reduced dimension n <= 64 (e.g. trot at H=10: n = 60):
// ONE WARP PER QP, no block-wide barrier anywhere.
//
// Layout and mapping
//   * padded dimension 64; lane l owns rows l and l+32 ("lane-rows"): vectors live in registers
//     or are read conflict-free from shared memory with consecutive lanes -> consecutive rows.
//   * the Cholesky factor lives in shared memory in PANEL storage: 8 panels of 8 columns; panel J
//     holds rows 8J..63 column-major with column stride (64 - 8J) + 2.  The "+2" makes three
//     access patterns bank-conflict free at once: (a) consecutive rows of one column (64-bit,
//     consecutive lanes), (b) 8 consecutive rows of one column as broadcast 128-bit loads,
//     (c) per-lane 128-bit loads of 8 consecutive rows in per-lane columns (stride = 2 or 10 mod 16
//     doubles, so the 8 lanes of a quarter-warp hit 8 distinct 16-byte bank groups).
//   * the scaled Hessian H stays in an L2-resident global workspace (full symmetric 64x64,
//     column-major): it is read once per factorization (accumulator init) and once per H*u.
//   * factorization: left-looking by panels with a 2x8 register accumulator tile per lane; the
//     8x8 diagonal block is gathered to every lane and factored AND inverted redundantly in
//     registers (right-looking, dependency depth = one rsqrt + 2 FMAs per column), so the panel
//     solve and both triangular solves are 8x8 mat-vecs with the stored inverse blocks instead of
//     64-step substitution chains.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "lmpc_types.cuh"

namespace lmpc {

constexpr int WN = 64;        // padded dimension of the warp path
constexpr int WLP = 2432;     // doubles of panel storage
constexpr int WNBMAX = 21;    // max stance blocks (3*21 = 63 <= 64)

__device__ __forceinline__ int wp_base(int J) { return 528 * J - 32 * J * (J - 1); }
__device__ __forceinline__ int wp_stride(int J) { return 66 - 8 * J; }
__device__ __forceinline__ int wp_addr(int row, int col) {
  const int J = col >> 3;
  return wp_base(J) + (col & 7) * wp_stride(J) + row - 8 * J;
}
__host__ __device__ constexpr int tri(int r, int c) { return (r * (r + 1)) / 2 + c; }

__device__ __forceinline__ double wsum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double wmin(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double wmax(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// shared-memory plan of one warp (offsets in doubles, then bytes)
struct WarpPlan {
  int oLP, oU, oGS, oZ, oPRM, oT8, oALIAS, oHU, oRD, oDUA, oDU, oWROW;  // doubles
  int oBW, oCOS, oSIN, oCS, oSS, oE, oAJ, oBJ;                           // setup, inside ALIAS
  int n_doubles;
  int oBLK, oFACE, oBMAP, total_bytes;
};
__host__ __device__ inline WarpPlan make_warp_plan(int H) {
  WarpPlan p;
  int o = 0;
  p.oLP = o; o += WLP;
  p.oU = o; o += WN;
  p.oGS = o; o += WN;
  p.oZ = o; o += 112;
  p.oPRM = o; o += 32;
  p.oT8 = o; o += 16;
  p.oALIAS = o;
  p.oHU = o; p.oRD = o + WN; p.oDUA = o + 2 * WN; p.oDU = o + 3 * WN; p.oWROW = o + 4 * WN;
  const int solve_sz = 4 * WN + 3 * WN;
  int s = o;
  p.oBW = s; s += 36;
  p.oCOS = s; s += H; p.oSIN = s; s += H; p.oCS = s; s += H; p.oSS = s; s += H;
  p.oE = s; s += 12 * H; p.oAJ = s; s += 3 * H; p.oBJ = s; s += 3 * H;
  const int setup_sz = s - o;
  o += solve_sz > setup_sz ? solve_sz : setup_sz;
  p.n_doubles = o;
  int b = o * 8;
  p.oBLK = b; b += 2 * 24;
  p.oFACE = b; b += 4 * 24;
  p.oBMAP = b; b += 2 * 4 * H;
  p.total_bytes = (b + 15) & ~15;
  return p;
}

// ---------------------------------------------------------------- warp-level dense kernels

// y = H x with H full symmetric 64x64 column-major in global memory, x in shared memory.
// Result in registers (y0 = row lane, y1 = row lane+32).
// Hg is written earlier in the same kernel: loads go through L2 (__ldcg), never the read-only path.
__device__ __forceinline__ void wsymv(const double* Hg, const double* x, int lane, double& y0, double& y1) {
  double a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll 8
  for (int c = 0; c < WN; c += 2) {
    const double x0 = x[c], x1 = x[c + 1];
    a0 += __ldcg(Hg + c * WN + lane) * x0;
    b0 += __ldcg(Hg + c * WN + lane + 32) * x0;
    a1 += __ldcg(Hg + (c + 1) * WN + lane) * x1;
    b1 += __ldcg(Hg + (c + 1) * WN + lane + 32) * x1;
  }
  y0 = a0 + a1;
  y1 = b0 + b1;
}

// In-register Cholesky of an 8x8 SPD block (lower, packed tri()) followed by in-place inversion
// of the triangular factor.  On exit D = inv(L) (lower).  Returns false on a non-positive pivot.
__device__ __forceinline__ bool chol8_inv(double (&D)[36]) {
  bool ok = true;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const double d = D[tri(c, c)];
    ok = ok && (d > 0.0);
    const double ri = rsqrt(d > 0.0 ? d : 1.0);
    D[tri(c, c)] = ri;
#pragma unroll
    for (int r = c + 1; r < 8; ++r) D[tri(r, c)] *= ri;
#pragma unroll
    for (int c2 = c + 1; c2 < 8; ++c2)
#pragma unroll
      for (int r = c2; r < 8; ++r) D[tri(r, c2)] -= D[tri(r, c)] * D[tri(c2, c)];
  }
  // inverse of the lower-triangular factor, in place (diagonal already holds 1/L_cc)
#pragma unroll
  for (int c = 0; c < 8; ++c) {
#pragma unroll
    for (int r = c + 1; r < 8; ++r) {
      double s = D[tri(r, c)] * D[tri(c, c)];
#pragma unroll
      for (int k = c + 1; k < r; ++k) s += D[tri(r, k)] * D[tri(k, c)];
      D[tri(r, c)] = -D[tri(r, r)] * s;
    }
  }
  return ok;
}

// Factor M = L L'.  FROM_H: M = Hg + blockdiag(W) with the 3x3 W blocks given row-wise in
// wrow[row][3] (rows >= n carry no W).  !FROM_H: M's lower triangle is already in Lp.
// On exit Lp holds L (strictly lower inside diagonal blocks) and inv(L_JJ)' in the upper
// triangle (incl. diagonal) of every 8x8 diagonal block.
template <bool FROM_H>
__device__ __forceinline__ bool wfactor(const double* Hg, const double* wrow, int n, double* Lp, int lane) {
  bool ok = true;
#pragma unroll 1
  for (int J = 0; J < 8; ++J) {
    const int c0 = 8 * J;
    const int sJ = J >> 2;          // row slot holding the diagonal rows
    const int l0 = c0 & 31;         // first lane of the diagonal rows
    double acc[2][8];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const int row = lane + 32 * s;
      const bool live = row >= c0;
      if (FROM_H) {
        const int b3 = (row / 3) * 3;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          double v = 0.0;
          if (live) {
            v = __ldcg(Hg + (c0 + c) * WN + row);
            const int d = c0 + c - b3;
            if (row < n && d >= 0 && d < 3) v += wrow[3 * row + d];
          }
          acc[s][c] = v;
        }
      } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[s][c] = (live && row >= c0 + c) ? Lp[wp_addr(row, c0 + c)] : 0.0;
      }
    }
    // left-looking update with all previous panels
#pragma unroll 1
    for (int K = 0; K < J; ++K) {
      const int bK = wp_base(K), sK = wp_stride(K);
#pragma unroll 2
      for (int kk = 0; kk < 8; ++kk) {
        const double* col = Lp + bK + kk * sK - 8 * K;  // column 8K+kk, indexed by row
        const double2 p0 = *reinterpret_cast<const double2*>(col + c0);
        const double2 p1 = *reinterpret_cast<const double2*>(col + c0 + 2);
        const double2 p2 = *reinterpret_cast<const double2*>(col + c0 + 4);
        const double2 p3 = *reinterpret_cast<const double2*>(col + c0 + 6);
        const double lc[8] = {p0.x, p0.y, p1.x, p1.y, p2.x, p2.y, p3.x, p3.y};
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const int row = lane + 32 * s;
          if (32 * (s + 1) > c0) {  // slot has rows >= c0 (warp-uniform)
            const double lr = (row >= c0) ? col[row] : 0.0;
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[s][c] -= lr * lc[c];
          }
        }
      }
    }
    // gather the (updated) diagonal block to every lane through its own storage slot
    {
      const int r = lane - l0;
      if (r >= 0 && r < 8) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const double v = sJ == 0 ? acc[0][c] : acc[1][c];
          if (c <= r) Lp[wp_base(J) + c * wp_stride(J) + r] = v;
        }
      }
    }
    __syncwarp();
    double D[36];
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
      for (int r = c; r < 8; ++r) D[tri(r, c)] = Lp[wp_base(J) + c * wp_stride(J) + r];
    __syncwarp();
    ok = chol8_inv(D) && ok;
    // panel solve X = acc * inv(L_JJ)'  and store
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const int row = lane + 32 * s;
      if (32 * (s + 1) > c0) {
        double x[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          double v = 0.0;
#pragma unroll
          for (int c2 = 0; c2 <= c; ++c2) v += acc[s][c2] * D[tri(c, c2)];
          x[c] = v;
        }
        const int r = row - c0;
        if (r >= 0) {
#pragma unroll
          for (int c = 0; c < 8; ++c)
            if (r >= 8 || c < r) Lp[wp_base(J) + c * wp_stride(J) + r] = x[c];
        }
      }
    }
    if (lane == 0) {  // inv(L_JJ)' into the upper triangle (incl. diagonal) of the block
#pragma unroll
      for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int r = 0; r <= c; ++r) Lp[wp_base(J) + c * wp_stride(J) + r] = D[tri(c, r)];
    }
    __syncwarp();
  }
  return ok;
}

// Solve L L' x = b.  b0/b1: this lane's entries (rows lane, lane+32) in registers, in/out.
__device__ __forceinline__ void wsolve(const double* Lp, double* t8, int lane,
                                       double& b0, double& b1) {
  // forward: L y = b
#pragma unroll 1
  for (int J = 0; J < 8; ++J) {
    const int c0 = 8 * J, sJ = J >> 2, l0 = c0 & 31;
    const int bJ = wp_base(J), stJ = wp_stride(J);
    const int r = lane - l0;
    if (r >= 0 && r < 8) t8[r] = sJ == 0 ? b0 : b1;
    __syncwarp();
    double t[8], y[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) t[c] = t8[c];
    // y = inv(L_JJ) t ;  inv(L)[c][c2] is stored at (row c2, col c) of the block
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      double v = 0.0;
#pragma unroll
      for (int c2 = 0; c2 <= c; ++c2) v += Lp[bJ + c * stJ + c2] * t[c2];
      y[c] = v;
    }
    __syncwarp();
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const int row = lane + 32 * s;
      double& b = s == 0 ? b0 : b1;
      const int rr = row - c0;
      if (rr >= 8) {
        double a0 = 0, a1 = 0;
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
          a0 += Lp[bJ + c * stJ + rr] * y[c];
          a1 += Lp[bJ + (c + 1) * stJ + rr] * y[c + 1];
        }
        b -= a0 + a1;
      } else if (rr >= 0) {
        double v = y[0];
#pragma unroll
        for (int c = 1; c < 8; ++c) v = (rr == c) ? y[c] : v;
        b = v;
      }
    }
  }
  // backward: L' x = y
#pragma unroll 1
  for (int J = 7; J >= 0; --J) {
    const int c0 = 8 * J, sJ = J >> 2, l0 = c0 & 31;
    const int bJ = wp_base(J), stJ = wp_stride(J);
    const int r = lane - l0;
    if (r >= 0 && r < 8) t8[r] = sJ == 0 ? b0 : b1;
    __syncwarp();
    double t[8], x[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) t[c] = t8[c];
    // x = inv(L_JJ)' t :  x[c] = sum_{c2 >= c} inv(L)[c2][c] t[c2] ; inv(L)[c2][c] at (row c, col c2)
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      double v = 0.0;
#pragma unroll
      for (int c2 = c; c2 < 8; ++c2) v += Lp[bJ + c2 * stJ + c] * t[c2];
      x[c] = v;
    }
    __syncwarp();
    // rows of this panel take x; earlier rows k < c0 subtract sum_c L[c0+c][k] x[c]
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const int k = lane + 32 * s;
      double& b = s == 0 ? b0 : b1;
      const int rr = k - c0;
      if (rr >= 0 && rr < 8) {
        double v = x[0];
#pragma unroll
        for (int c = 1; c < 8; ++c) v = (rr == c) ? x[c] : v;
        b = v;
      } else if (rr < 0) {
        const double* p = Lp + wp_addr(c0, k);  // 8 consecutive rows c0..c0+7 of column k
        const double2 q0 = *reinterpret_cast<const double2*>(p);
        const double2 q1 = *reinterpret_cast<const double2*>(p + 2);
        const double2 q2 = *reinterpret_cast<const double2*>(p + 4);
        const double2 q3 = *reinterpret_cast<const double2*>(p + 6);
        b -= (q0.x * x[0] + q0.y * x[1]) + (q1.x * x[2] + q1.y * x[3]) + (q2.x * x[4] + q2.y * x[5]) +
             (q3.x * x[6] + q3.y * x[7]);
      }
    }
  }
}

}  // namespace lmpc
