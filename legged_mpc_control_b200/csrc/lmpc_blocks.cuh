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

// ---------------------------------------------------------------- face repair of the exact polish
// One stance block after an equality-constrained solve on the guessed face (sx, sy in {-1, 0, 1}: which
// side of the x / y pyramid is active; top: fz = fz_max; apex: f = 0): g = H u + g_lin is the gradient
// (the multipliers follow from it through the shear), f the block's force.  face_step() returns
//   * the face with every indicated ADD applied (a violated row joins the face), and — only if
//     `drop_all` — every indicated DROP as well (a row whose multiplier has the wrong sign leaves);
//   * add: some row of this block is primal infeasible;
//   * dscore / dwhich: the block's most violated multiplier (0 x face, 1 y face, 2 top, 3 leave the apex).
// Rounds 0-2 of both drivers use drop_all (fast, usually 1-2 rounds); later rounds add while any block
// is infeasible and otherwise drop the single most violated multiplier of the whole problem
// (face_single_drop on the block that holds it), the classical rule that cannot cycle.
struct FaceStep { int sx, sy, top, apex; bool add; double dscore; int dwhich; };

__device__ inline FaceStep face_step(int sx, int sy, int top, int apex, double gx, double gy, double gz, double fx,
                                     double fy, double fz, double mu, double fzm, double dual_tol, double feas_tol,
                                     bool drop_all) {
  FaceStep r = {sx, sy, top, apex, false, 0.0, -1};
  if (apex) {
    const double v = mu * (fabs(gx) + fabs(gy)) - gz;
    if (v > dual_tol) { r.dscore = v; r.dwhich = 3; }
    if (drop_all && r.dwhich == 3) {
      r.apex = 0; r.top = 0;
      r.sx = gx > 0 ? -1 : (gx < 0 ? 1 : 0);
      r.sy = gy > 0 ? -1 : (gy < 0 ? 1 : 0);
    }
    return r;
  }
  const double gvz = gz + mu * (sx * gx + sy * gy);
  const double vx = sx != 0 ? -(sx == -1 ? gx : -gx) : 0.0;
  const double vy = sy != 0 ? -(sy == -1 ? gy : -gy) : 0.0;
  const double vt = top ? gvz : 0.0;
  if (vx > dual_tol) { r.dscore = vx; r.dwhich = 0; }
  if (vy > dual_tol && vy > r.dscore) { r.dscore = vy; r.dwhich = 1; }
  if (vt > dual_tol && vt > r.dscore) { r.dscore = vt; r.dwhich = 2; }
  if (fz > fzm + feas_tol && !top) { r.top = 1; r.add = true; }
  if (fz < -feas_tol) { r.apex = 1; r.add = true; }
  else {
    if (sx == 0) { if (-fx - mu * fz > feas_tol) { r.sx = -1; r.add = true; } else if (fx - mu * fz > feas_tol) { r.sx = 1; r.add = true; } }
    if (sy == 0) { if (-fy - mu * fz > feas_tol) { r.sy = -1; r.add = true; } else if (fy - mu * fz > feas_tol) { r.sy = 1; r.add = true; } }
  }
  if (drop_all) {
    if (sx != 0 && r.sx == sx && vx > dual_tol) r.sx = 0;
    if (sy != 0 && r.sy == sy && vy > dual_tol) r.sy = 0;
    if (top && r.top == top && vt > dual_tol) r.top = 0;
  }
  return r;
}
// apply the single drop r.dwhich to r's face (late rounds, on the block whose dscore is the problem's maximum)
__device__ inline void face_single_drop(FaceStep& r, double gx, double gy) {
  if (r.dwhich == 0) r.sx = 0;
  else if (r.dwhich == 1) r.sy = 0;
  else if (r.dwhich == 2) r.top = 0;
  else if (r.dwhich == 3) {
    r.apex = 0; r.top = 0;
    r.sx = gx > 0 ? -1 : (gx < 0 ? 1 : 0);
    r.sy = gy > 0 ? -1 : (gy < 0 ? 1 : 0);
  }
}

enum { PRM_Q = 0, PRM_R = 12, PRM_MU = 24, PRM_FZ = 25, PRM_MASS = 26, PRM_SC = 27, PRM_VW = 28 };

}  // namespace lmpc
