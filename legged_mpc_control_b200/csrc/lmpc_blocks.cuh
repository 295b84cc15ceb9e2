// lmpc_blocks.cuh — friction-pyramid block algebra shared by the solve drivers (one-warp and tile
// paths): slack, slack direction and C_b' w of one stance block, and the slots of the per-problem
// parameter scratch.  Reference rows: ConvexQPSolver.cpp:131-177.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include "lmpc_types.cuh"

namespace lmpc {

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

enum { PRM_Q = 0, PRM_R = 12, PRM_MU = 24, PRM_FZ = 25, PRM_MASS = 26, PRM_SC = 27, PRM_VW = 28 };

}  // namespace lmpc
