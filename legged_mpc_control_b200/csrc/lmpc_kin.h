// lmpc_kin.h — A1/Go1 leg kinematics for the batched low-level stage downstream of the MPC
// (SURVEY.md §8f-3): forward kinematics, Jacobian, analytic inverse kinematics and a pivoted 3x3
// solve, as __host__ __device__ functions.  Reference (paths relative to
// /root/reference/src/legged_ctrl/): src/utils/A1Kinematics.cpp — fk :8-12 + :42-80 (Matlab
// generated), jac :14-18 + :82-140, analytic inv_kin :321-446; callers
// src/interfaces/BaseInterface.cpp:280-300 (feedback) and :451-500 (tau_ctrl_update).
//
// Geometry (rho_fix = {ox, oy, d, lt, lc}: hip offset x/y, abduction link, thigh, calf;
// rho_opt = {cx, cy, cz}: calibration offsets of the foot, zero in the reference's controllers):
// with s12 = sin(q1 + q2), c12 = cos(q1 + q2), dd = d + cy, lc' = lc - cz,
//     X = -lt sin q1 + cx c12 - lc' s12          (sagittal-plane forward reach)
//     L =  lt cos q1 + cx s12 + lc' c12          (sagittal-plane extension, positive down)
//     p = [ox + X,  oy + dd cos q0 + L sin q0,  dd sin q0 - L cos q0]
// which is the generated polynomial of A1Kinematics.cpp:66-79 collected by angle sums.  Note
// dX/dq1 = -L and dL/dq1 = X.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define LMPC_HD __host__ __device__
#else
#define LMPC_HD
#endif

namespace lmpc {

struct LegGeom { double ox, oy, d, lt, lc, cx, cy, cz; };

LMPC_HD inline void a1_fk(const double q[3], const LegGeom& g, double p[3]) {
  const double s0 = sin(q[0]), c0 = cos(q[0]), s1 = sin(q[1]), c1 = cos(q[1]);
  const double s12 = sin(q[1] + q[2]), c12 = cos(q[1] + q[2]);
  const double dd = g.d + g.cy, lcc = g.lc - g.cz;
  const double X = -g.lt * s1 + g.cx * c12 - lcc * s12;
  const double L = g.lt * c1 + g.cx * s12 + lcc * c12;
  p[0] = g.ox + X;
  p[1] = g.oy + dd * c0 + L * s0;
  p[2] = dd * s0 - L * c0;
}

// J[r + 3 c] = d p_r / d q_c (column-major, Eigen::Matrix3d::data() order)
LMPC_HD inline void a1_jac(const double q[3], const LegGeom& g, double J[9]) {
  const double s0 = sin(q[0]), c0 = cos(q[0]), s1 = sin(q[1]), c1 = cos(q[1]);
  const double s12 = sin(q[1] + q[2]), c12 = cos(q[1] + q[2]);
  const double dd = g.d + g.cy, lcc = g.lc - g.cz;
  const double X = -g.lt * s1 + g.cx * c12 - lcc * s12;
  const double L = g.lt * c1 + g.cx * s12 + lcc * c12;
  const double X2 = -g.cx * s12 - lcc * c12;  // dX/dq2
  const double L2 = g.cx * c12 - lcc * s12;   // dL/dq2
  J[0] = 0.0;  J[1] = -dd * s0 + L * c0;  J[2] = dd * c0 + L * s0;   // d/dq0
  J[3] = -L;   J[4] = s0 * X;             J[5] = -c0 * X;            // d/dq1 (dX/dq1 = -L, dL/dq1 = X)
  J[6] = X2;   J[7] = s0 * L2;            J[8] = -c0 * L2;           // d/dq2
}

// Analytic inverse kinematics with the reference's branch selection (A1Kinematics.cpp:321-446):
// of the two abduction angles that reach (y, z) the one nearer to the current q0 is taken; the
// knee angle is in [-pi, 0]; the thigh angle is wrapped into [-60 deg, 240 deg].  rho_opt is
// ignored, as in the reference.  A foot inside the abduction circle gives NaN (sqrt of a negative
// number, :336-337), which the caller treats as "keep the current joint angle" (:472-478).
LMPC_HD inline void a1_inv_kin(const double p[3], const double cur_q[3], const LegGeom& g, double q[3]) {
  const double kPi = 3.14159265358979323846;
  const double yf = p[1] - g.oy, xf = p[0] - g.ox, zf = p[2];
  double L = sqrt(zf * zf + yf * yf - g.d * g.d);
  double t1 = 0.0, alt = 0.0;
  if (g.oy > 0) {  // left legs (:349-379)
    const double psi = atan2(L, g.d);
    if (zf > 0) {
      t1 = (yf > 0 ? atan2(zf, yf) : (yf == 0 ? kPi / 2 : kPi - atan2(zf, -yf))) - psi;
      alt = atan2(zf, yf) + psi;
    } else if (zf < 0) {
      t1 = (yf > 0 ? atan2(zf, yf) : (yf == 0 ? -kPi / 2 : -kPi - atan2(zf, -yf))) + psi;
      alt = atan2(zf, yf) - psi;
    } else {
      t1 = psi;
      alt = -psi;
    }
  } else {  // right legs (:380-410); zf == 0 leaves both candidates at 0, as the reference does
    const double psi = atan2(L, -g.d);
    if (zf > 0) {
      t1 = (yf < 0 ? -atan2(zf, -yf) : (yf == 0 ? -kPi / 2 : -kPi + atan2(zf, yf))) + psi;
      alt = -psi - atan2(zf, -yf);
    } else if (zf < 0) {
      t1 = (yf < 0 ? atan2(-zf, -yf) : (yf == 0 ? -kPi / 2 : kPi - atan2(-zf, yf))) - psi;
      alt = atan2(-zf, -yf) + psi;
    }
  }
  if (!(fabs(t1 - cur_q[0]) < fabs(alt - cur_q[0]))) t1 = alt;
  // knee (:413-423)
  const double cb = (g.lt * g.lt + g.lc * g.lc - xf * xf - L * L) / (2 * g.lt * g.lc);
  const double beta = fabs(cb + 1) < 0.001 ? kPi : (fabs(cb - 1) < 0.001 ? 0.0 : acos(cb));
  const double t3 = beta - kPi;
  // thigh (:425-441): the sign of the extension depends on which side of the abduction joint the foot is
  if (zf > g.d * sin(t1)) L = -L;
  double t2 = atan2(-xf, L) + atan2(g.lc * sin(-t3), g.lt + g.lc * cos(-t3));
  if (t2 < -60 * kPi / 180) t2 += 2 * kPi;
  else if (t2 > 240 * kPi / 180) t2 -= 2 * kPi;
  q[0] = t1; q[1] = t2; q[2] = t3;
}

// x = A^-1 b for a column-major 3x3 by Gaussian elimination with partial pivoting — what
// Eigen's PartialPivLU does for jac.lu().solve (BaseInterface.cpp:346,480).  A singular matrix yields
// inf/NaN, which the caller maps to its fallback exactly like the reference's isnan tests.
LMPC_HD inline void solve3_pivoted(const double A[9], const double b[3], double x[3]) {
  double M[3][4];
  for (int r = 0; r < 3; ++r) { M[r][0] = A[r]; M[r][1] = A[r + 3]; M[r][2] = A[r + 6]; M[r][3] = b[r]; }
  for (int c = 0; c < 3; ++c) {
    int piv = c;
    for (int r = c + 1; r < 3; ++r)
      if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
    if (piv != c)
      for (int k = 0; k < 4; ++k) { const double t = M[c][k]; M[c][k] = M[piv][k]; M[piv][k] = t; }
    for (int r = c + 1; r < 3; ++r) {
      const double f = M[r][c] / M[c][c];
      for (int k = c; k < 4; ++k) M[r][k] -= f * M[c][k];
    }
  }
  x[2] = M[2][3] / M[2][2];
  x[1] = (M[1][3] - M[1][2] * x[2]) / M[1][1];
  x[0] = (M[0][3] - M[0][1] * x[1] - M[0][2] * x[2]) / M[0][0];
}

}  // namespace lmpc
