/*
 * lmpc_oracle.c — CPU restatement of legged_ctrl's convex-MPC QP.  TEST INFRASTRUCTURE ONLY
 * (see lmpc_oracle.h: parity unpinned; never linked into the product).
 *
 * References are to /root/reference/src/legged_ctrl/:
 *   src/mpc_ctrl/convex_mpc/ConvexQPSolver.cpp   (the QP)
 *   src/utils/LeggedContactFSM.cpp                (contact schedule)
 *   src/utils/Utils.cpp:89-95                     (skew)
 * Plain C99, no third-party code.  Dense, unoptimised on purpose: clarity over speed.
 */
#include "lmpc_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define NS 12 /* state dim  (MPC_STATE_DIM_SPARSE, LeggedParams.h:15) */
#define NU 12 /* input dim  (DIM_GRF, LeggedParams.h:21) */
#define ORC_INF 1e30

/* ------------------------------------------------------------------ small helpers */

static void mat3_mul(const double* A, const double* B, double* C) { /* column-major 3x3 */
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[i + 3 * k] * B[k + 3 * j];
      C[i + 3 * j] = s;
    }
}

static void mat3_transpose(const double* A, double* At) {
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) At[i + 3 * j] = A[j + 3 * i];
}

/* general 3x3 inverse by cofactors (what Eigen's fixed-size inverse() does for 3x3) */
static void mat3_inverse(const double* A, double* Ai) {
  double a = A[0], b = A[3], c = A[6];
  double d = A[1], e = A[4], f = A[7];
  double g = A[2], h = A[5], i = A[8];
  double c00 = e * i - f * h, c01 = -(d * i - f * g), c02 = d * h - e * g;
  double c10 = -(b * i - c * h), c11 = a * i - c * g, c12 = -(a * h - b * g);
  double c20 = b * f - c * e, c21 = -(a * f - c * d), c22 = a * e - b * d;
  double det = a * c00 + b * c01 + c * c02;
  double id = 1.0 / det;
  /* inverse = adj / det, adj = cofactor^T ; store column-major */
  Ai[0] = c00 * id; Ai[3] = c10 * id; Ai[6] = c20 * id;
  Ai[1] = c01 * id; Ai[4] = c11 * id; Ai[7] = c21 * id;
  Ai[2] = c02 * id; Ai[5] = c12 * id; Ai[8] = c22 * id;
}

/* Utils::skew (Utils.cpp:89-95), column-major */
static void skew3(const double* v, double* S) {
  S[0] = 0;     S[3] = -v[2]; S[6] = v[1];
  S[1] = v[2];  S[4] = 0;     S[7] = -v[0];
  S[2] = -v[1]; S[5] = v[0];  S[8] = 0;
}

/* ------------------------------------------------------------------ model build */

/* Continuous matrices.  Ac per update_A_matrix (ConvexQPSolver.cpp:214-225), Bc per
 * update_B_matrix (:198-208).  12x12 column-major each. */
static void continuous_matrices(const lmpc_params* p, const double rot[9], const double foot[12],
                                double yaw_ref, double* Ac, double* Bc) {
  memset(Ac, 0, sizeof(double) * NS * NS);
  memset(Bc, 0, sizeof(double) * NS * NU);
  double cy = cos(yaw_ref), sy = sin(yaw_ref);
  /* ang_vel_to_rpy_rate << c, s, 0, -s, c, 0, 0, 0, 1   placed at block (0,6) */
  Ac[0 + NS * 6] = cy;  Ac[0 + NS * 7] = sy;  Ac[0 + NS * 8] = 0;
  Ac[1 + NS * 6] = -sy; Ac[1 + NS * 7] = cy;  Ac[1 + NS * 8] = 0;
  Ac[2 + NS * 6] = 0;   Ac[2 + NS * 7] = 0;   Ac[2 + NS * 8] = 1;
  /* identity at block (3,9) */
  for (int i = 0; i < 3; ++i) Ac[(3 + i) + NS * (9 + i)] = 1.0;

  /* a1_trunk_inertia_world = R * I * R^T ; inverse taken once per leg in the reference */
  double Rt[9], RI[9], Iw[9], Iwi[9];
  mat3_transpose(rot, Rt);
  mat3_mul(rot, p->inertia, RI);
  mat3_mul(RI, Rt, Iw);
  mat3_inverse(Iw, Iwi);
  for (int leg = 0; leg < LMPC_NUM_LEG; ++leg) {
    double S[9], M[9];
    skew3(foot + 3 * leg, S);
    mat3_mul(Iwi, S, M);
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) Bc[(6 + i) + NS * (3 * leg + j)] = M[i + 3 * j];
    for (int i = 0; i < 3; ++i) Bc[(9 + i) + NS * (3 * leg + i)] = 1.0 / p->mass;
  }
}

void orc_step_matrices(const lmpc_params* p, const double rot[9], const double foot[12],
                       double yaw_ref, double* Ad, double* Bd, double* cd) {
  double Ac[NS * NS], Bc[NS * NU], c[NS];
  continuous_matrices(p, rot, foot, yaw_ref, Ac, Bc);
  memset(c, 0, sizeof(c));
  c[11] = -p->gravity; /* d(v_z)/dt = ... - g  (ConvexQPSolver.cpp:174-176, 296) */
  const double dt = p->dt;
  if (p->discretisation == LMPC_DISC_EULER) {
    /* Ad = I + Ac*dt (:227), Bd = Bc*dt (:211), gravity 9.8*dt (:175) */
    for (int k = 0; k < NS * NS; ++k) Ad[k] = Ac[k] * dt;
    for (int i = 0; i < NS; ++i) Ad[i + NS * i] += 1.0;
    for (int k = 0; k < NS * NU; ++k) Bd[k] = Bc[k] * dt;
    for (int i = 0; i < NS; ++i) cd[i] = c[i] * dt;
    return;
  }
  /* ZOH: exp(M dt), M = [[Ac, Bc, c],[0,0,0],[0,0,0]] (25x25), by Taylor series.  The series
   * terminates (Ac^2 = 0) but is summed generically so no closed form is assumed here. */
  enum { N = NS + NU + 1 };
  static const int TERMS = 24;
  double M[N * N], T[N * N], E[N * N], Tn[N * N];
  memset(M, 0, sizeof(M));
  for (int j = 0; j < NS; ++j)
    for (int i = 0; i < NS; ++i) M[i + N * j] = Ac[i + NS * j] * dt;
  for (int j = 0; j < NU; ++j)
    for (int i = 0; i < NS; ++i) M[i + N * (NS + j)] = Bc[i + NS * j] * dt;
  for (int i = 0; i < NS; ++i) M[i + N * (NS + NU)] = c[i] * dt;
  memset(E, 0, sizeof(E));
  memset(T, 0, sizeof(T));
  for (int i = 0; i < N; ++i) E[i + N * i] = T[i + N * i] = 1.0;
  for (int t = 1; t <= TERMS; ++t) {
    for (int j = 0; j < N; ++j)
      for (int i = 0; i < N; ++i) {
        double s = 0;
        for (int k = 0; k < N; ++k) s += T[i + N * k] * M[k + N * j];
        Tn[i + N * j] = s / t;
      }
    memcpy(T, Tn, sizeof(T));
    for (int k = 0; k < N * N; ++k) E[k] += T[k];
  }
  for (int j = 0; j < NS; ++j)
    for (int i = 0; i < NS; ++i) Ad[i + NS * j] = E[i + N * j];
  for (int j = 0; j < NU; ++j)
    for (int i = 0; i < NS; ++i) Bd[i + NS * j] = E[i + N * (NS + j)];
  for (int i = 0; i < NS; ++i) cd[i] = E[i + N * (NS + NU)];
}

void orc_reference_trajectory(const lmpc_params* p, const double x0[12], const double rot[9],
                              const double cmd[9], double* xref) {
  /* state.ctrl.root_lin_vel_d_world = root_rot_mat * root_lin_vel_d_rel (ConvexQPSolver.cpp:260) */
  double vw[3];
  for (int i = 0; i < 3; ++i)
    vw[i] = rot[i] * cmd[3] + rot[i + 3] * cmd[4] + rot[i + 6] * cmd[5];
  for (int i = 0; i < p->horizon; ++i) {
    const double k = (double)(i + p->xref_index_offset); /* evaluated as (rate * dt) * i, like :267-269 */
    double* xr = xref + NS * i; /* :264-276 */
    xr[0] = cmd[0];                 /* root_euler_d[0] */
    xr[1] = cmd[1];                 /* root_euler_d[1] */
    xr[2] = x0[2] + cmd[8] * p->dt * k; /* fbk yaw + yaw rate_d * dt * i */
    xr[3] = x0[3] + vw[0] * p->dt * k;
    xr[4] = x0[4] + vw[1] * p->dt * k;
    xr[5] = cmd[2];                 /* root_pos_d[2] */
    xr[6] = cmd[6];
    xr[7] = cmd[7];
    xr[8] = cmd[8];
    xr[9] = vw[0];
    xr[10] = vw[1];
    xr[11] = 0.0;
  }
}

static double cost_scale(const lmpc_params* p) {
  return p->cost_scaling == LMPC_COST_NS_TWO ? 2.0 : 1.0;
}

void orc_sparse_qp(const lmpc_params* p, const double x0[12], const double rot[9],
                   const double foot[12], const double cmd[9], const uint8_t* contact,
                   double* Pdiag, double* q, double* A, double* l, double* u) {
  const int H = p->horizon;
  const int nz = 24 * H, ndyn = 12 * H, nfric = 16 * H, nrow = 32 * H;
  const double cs = cost_scale(p);
  memset(A, 0, sizeof(double) * (size_t)nrow * nz);
  memset(q, 0, sizeof(double) * nz);
  /* hessian diag alternates R (12) and Q (12)  (:33-49) */
  for (int k = 0; k < H; ++k)
    for (int i = 0; i < 12; ++i) {
      Pdiag[24 * k + i] = cs * p->r[i];
      Pdiag[24 * k + 12 + i] = cs * p->q[i];
    }
  double* xref = (double*)malloc(sizeof(double) * NS * H);
  orc_reference_trajectory(p, x0, rot, cmd, xref);
  for (int i = 0; i < H; ++i) {
    double Ad[NS * NS], Bd[NS * NU], cd[NS];
    /* linearised about the reference yaw of step i (:279) ; B identical every step (:280-283) */
    orc_step_matrices(p, rot, foot, xref[NS * i + 2], Ad, Bd, cd);
    const int r0 = 12 * i;
    /* Bd at (12i, 24i)  (:105, :286-292) */
    for (int c = 0; c < NU; ++c)
      for (int r = 0; r < NS; ++r) A[(r0 + r) + (size_t)nrow * (24 * i + c)] = Bd[r + NS * c];
    /* -I at (12i, 24i+12)  (:106, :119) */
    for (int r = 0; r < NS; ++r) A[(r0 + r) + (size_t)nrow * (24 * i + 12 + r)] = -1.0;
    if (i == 0) {
      /* l = u = -Ad*x0, +9.8*dt on row 11  (:294-297) ; general affine term: -cd */
      for (int r = 0; r < NS; ++r) {
        double s = 0;
        for (int c = 0; c < NS; ++c) s += Ad[r + NS * c] * x0[c];
        l[r0 + r] = u[r0 + r] = -s - cd[r];
      }
    } else {
      /* Ad at (12i, 24(i-1)+12)  (:104, :299-305) ; rhs 9.8*dt on row 11 (:175-176) */
      for (int c = 0; c < NS; ++c)
        for (int r = 0; r < NS; ++r)
          A[(r0 + r) + (size_t)nrow * (24 * (i - 1) + 12 + c)] = Ad[r + NS * c];
      for (int r = 0; r < NS; ++r) l[r0 + r] = u[r0 + r] = -cd[r];
    }
    /* gradient segment = -Q * x_ref(i) at offset 24i+12  (:308) */
    for (int r = 0; r < NS; ++r) q[24 * i + 12 + r] = -cs * p->q[r] * xref[NS * i + r];
    /* friction rows (:131-158) */
    for (int j = 0; j < LMPC_NUM_LEG; ++j) {
      const int fr = ndyn + 16 * i + 4 * j;
      const int cx = 24 * i + 3 * j, cy = cx + 1, cz = cx + 2;
      A[(fr + 0) + (size_t)nrow * cx] = 1;  A[(fr + 0) + (size_t)nrow * cz] = p->mu;
      A[(fr + 1) + (size_t)nrow * cx] = 1;  A[(fr + 1) + (size_t)nrow * cz] = -p->mu;
      A[(fr + 2) + (size_t)nrow * cy] = 1;  A[(fr + 2) + (size_t)nrow * cz] = p->mu;
      A[(fr + 3) + (size_t)nrow * cy] = 1;  A[(fr + 3) + (size_t)nrow * cz] = -p->mu;
      l[fr + 0] = 0;        u[fr + 0] = ORC_INF;
      l[fr + 1] = -ORC_INF; u[fr + 1] = 0;
      l[fr + 2] = 0;        u[fr + 2] = ORC_INF;
      l[fr + 3] = -ORC_INF; u[fr + 3] = 0;
      /* f_z bound rows (:161-172, :329-346) */
      const int br = ndyn + nfric + 4 * i + j;
      A[br + (size_t)nrow * cz] = 1;
      l[br] = 0;
      u[br] = (contact[4 * i + j] ? 1.0 : 0.0) * p->fz_max;
    }
  }
  free(xref);
}

/* `contact` (may be NULL): when given, the columns of swing legs are skipped — their entries of Hm
 * and g stay untouched and are never read by the swing-eliminated solve.  The entries that ARE
 * computed go through the same operations in the same order either way (bit-identical); skipping
 * only keeps the timed CPU baseline (bench.py) from paying for columns the reduced QP discards. */
static void condensed_qp_impl(const lmpc_params* p, const double x0[12], const double rot[9],
                              const double foot[12], const double cmd[9], const uint8_t* contact,
                              double* Hm, double* g, double* e_out) {
  const int H = p->horizon, n = 12 * H;
#define COL_ON(c) (!contact || contact[4 * ((c) / 12) + ((c) % 12) / 3])
  const double cs = cost_scale(p);
  double* xref = (double*)malloc(sizeof(double) * NS * H);
  double* Ad = (double*)malloc(sizeof(double) * NS * NS * H);
  double* Bd = (double*)malloc(sizeof(double) * NS * NU * H);
  double* cd = (double*)malloc(sizeof(double) * NS * H);
  double* Bq = (double*)calloc((size_t)n * n, sizeof(double)); /* B_qp, col-major n x n */
  double* e = (double*)malloc(sizeof(double) * n);
  orc_reference_trajectory(p, x0, rot, cmd, xref);
  for (int k = 0; k < H; ++k)
    orc_step_matrices(p, rot, foot, xref[NS * k + 2], Ad + NS * NS * k, Bd + NS * NU * k,
                      cd + NS * k);
  /* free response x^f_{k+1} = Ad_k x^f_k + cd_k ;  e_k = x^f_{k+1} - x_ref(k) */
  double xf[NS], xn[NS];
  memcpy(xf, x0, sizeof(xf));
  for (int k = 0; k < H; ++k) {
    const double* A = Ad + NS * NS * k;
    for (int r = 0; r < NS; ++r) {
      double s = cd[NS * k + r];
      for (int c = 0; c < NS; ++c) s += A[r + NS * c] * xf[c];
      xn[r] = s;
    }
    memcpy(xf, xn, sizeof(xf));
    for (int r = 0; r < NS; ++r) e[NS * k + r] = xf[r] - xref[NS * k + r];
  }
  /* B_qp[k][j] = Ad_k ... Ad_{j+1} Bd_j  (rows of x_{k+1}, columns of u_j), k >= j */
  for (int j = 0; j < H; ++j) {
    for (int c = 0; c < NU; ++c) {
      if (!COL_ON(12 * j + c)) continue;
      for (int r = 0; r < NS; ++r) Bq[(12 * j + r) + (size_t)n * (12 * j + c)] = Bd[NS * NU * j + r + NS * c];
    }
    for (int k = j + 1; k < H; ++k) {
      const double* A = Ad + NS * NS * k;
      for (int c = 0; c < NU; ++c)
        for (int r = 0; r < NS && COL_ON(12 * j + c); ++r) {
          double s = 0;
          for (int t = 0; t < NS; ++t) s += A[r + NS * t] * Bq[(12 * (k - 1) + t) + (size_t)n * (12 * j + c)];
          Bq[(12 * k + r) + (size_t)n * (12 * j + c)] = s;
        }
    }
  }
  /* Hm = cs * (B_qp' Qbar B_qp + Rbar),  g = cs * B_qp' Qbar e */
  for (int c2 = 0; c2 < n; ++c2) {
    if (!COL_ON(c2)) continue;
    for (int c1 = c2; c1 < n; ++c1) {
      if (!COL_ON(c1)) continue;
      int kmin = (c1 / 12 > c2 / 12 ? c1 / 12 : c2 / 12);
      double s = 0;
      for (int row = 12 * kmin; row < n; ++row)
        s += Bq[row + (size_t)n * c1] * p->q[row % 12] * Bq[row + (size_t)n * c2];
      if (c1 == c2) s += p->r[c1 % 12];
      Hm[c1 + (size_t)n * c2] = Hm[c2 + (size_t)n * c1] = cs * s;
    }
    double s = 0;
    for (int row = 12 * (c2 / 12); row < n; ++row) s += Bq[row + (size_t)n * c2] * p->q[row % 12] * e[row];
    g[c2] = cs * s;
  }
  if (e_out) memcpy(e_out, e, sizeof(double) * n);
  free(xref); free(Ad); free(Bd); free(cd); free(Bq); free(e);
#undef COL_ON
}

void orc_condensed_qp(const lmpc_params* p, const double x0[12], const double rot[9],
                      const double foot[12], const double cmd[9], double* Hm, double* g,
                      double* e_out) {
  condensed_qp_impl(p, x0, rot, foot, cmd, NULL, Hm, g, e_out);
}

/* ------------------------------------------------------------------ Goldfarb-Idnani */
/* min 1/2 x'Gx + g'x  s.t.  sum_k cv[k] x[ci[k]] >= b[i] for rows i (CSR: rp, ci, cv).
 * Dual active-set method of Goldfarb & Idnani (Math. Prog. 27, 1983), written from the
 * paper: J = L^{-T} Q, R upper triangular with  J1' N = R.  The algorithm class qpOASES
 * implements (online active set); exact up to rounding. */

typedef struct {
  int n, m;
  double *J, *R, *d, *z, *r, *u;
  int *A;      /* active row indices, length q */
  int *where;  /* where[row] = position in A or -1 */
} gi_ws;

static gi_ws* gi_ws_new(int n, int m) {
  gi_ws* w = (gi_ws*)malloc(sizeof(gi_ws));
  w->n = n; w->m = m;
  w->J = (double*)malloc(sizeof(double) * (size_t)n * n);
  w->R = (double*)malloc(sizeof(double) * (size_t)n * n);
  w->d = (double*)malloc(sizeof(double) * n);
  w->z = (double*)malloc(sizeof(double) * n);
  w->r = (double*)malloc(sizeof(double) * (n + 1));
  w->u = (double*)malloc(sizeof(double) * (n + 1));
  w->A = (int*)malloc(sizeof(int) * (n + 1));
  w->where = (int*)malloc(sizeof(int) * (m > 0 ? m : 1));
  return w;
}

static void gi_ws_free(gi_ws* w) {
  free(w->J); free(w->R); free(w->d); free(w->z); free(w->r); free(w->u); free(w->A);
  free(w->where); free(w);
}

static int chol_lower(int n, double* G) { /* in place, column-major, lower */
  for (int j = 0; j < n; ++j) {
    double s = G[j + (size_t)n * j];
    for (int k = 0; k < j; ++k) s -= G[j + (size_t)n * k] * G[j + (size_t)n * k];
    if (!(s > 0.0)) return 1;
    double ljj = sqrt(s);
    G[j + (size_t)n * j] = ljj;
    for (int i = j + 1; i < n; ++i) {
      double t = G[i + (size_t)n * j];
      for (int k = 0; k < j; ++k) t -= G[i + (size_t)n * k] * G[j + (size_t)n * k];
      G[i + (size_t)n * j] = t / ljj;
    }
  }
  return 0;
}

static void rot_cols(int n, double* J, int c1, int c2, double c, double s) {
  double* a = J + (size_t)n * c1;
  double* b = J + (size_t)n * c2;
  for (int i = 0; i < n; ++i) {
    double t1 = c * a[i] + s * b[i];
    double t2 = -s * a[i] + c * b[i];
    a[i] = t1; b[i] = t2;
  }
}

static void gi_drop(gi_ws* w, int* q_io, int l) {
  const int n = w->n;
  int q = *q_io;
  w->where[w->A[l]] = -1;
  for (int c = l + 1; c < q; ++c) {
    memcpy(w->R + (size_t)n * (c - 1), w->R + (size_t)n * c, sizeof(double) * (c + 1));
    w->A[c - 1] = w->A[c];
    w->where[w->A[c - 1]] = c - 1;
    w->u[c - 1] = w->u[c];
  }
  w->u[q - 1] = w->u[q]; /* multiplier of the incoming constraint rides at the end */
  q -= 1;
  for (int k = l; k < q; ++k) {
    double a = w->R[k + (size_t)n * k], b = w->R[(k + 1) + (size_t)n * k];
    if (b == 0.0) continue;
    double h = hypot(a, b), c = a / h, s = b / h;
    for (int cc = k; cc < q; ++cc) {
      double t1 = c * w->R[k + (size_t)n * cc] + s * w->R[(k + 1) + (size_t)n * cc];
      double t2 = -s * w->R[k + (size_t)n * cc] + c * w->R[(k + 1) + (size_t)n * cc];
      w->R[k + (size_t)n * cc] = t1;
      w->R[(k + 1) + (size_t)n * cc] = t2;
    }
    rot_cols(n, w->J, k, k + 1, c, s);
  }
  *q_io = q;
}

static int gi_core(int n, double* G, const double* g, int m, const int* rp, const int* ci,
                   const double* cv, const double* b, double* x, double* lambda, int* iters,
                   gi_ws* w) {
  const double TOL_VIOL = 1e-10; /* primal violation that still counts as satisfied [units of b] */
  const double TOL_DEP = 1e-10;  /* |d2| <= TOL_DEP*|d|  -> incoming row depends on the active set */
  if (iters) *iters = 0;
  if (chol_lower(n, G)) return 3;
  /* J = L^{-T}: solve L' J = I (J upper triangular) */
  memset(w->J, 0, sizeof(double) * (size_t)n * n);
  for (int c = 0; c < n; ++c) {
    double* Jc = w->J + (size_t)n * c;
    for (int i = c; i >= 0; --i) {
      double s = (i == c) ? 1.0 : 0.0;
      for (int k = i + 1; k <= c; ++k) s -= G[k + (size_t)n * i] * Jc[k];
      Jc[i] = s / G[i + (size_t)n * i];
    }
  }
  /* x = -G^{-1} g */
  for (int i = 0; i < n; ++i) {
    double s = -g[i];
    for (int k = 0; k < i; ++k) s -= G[i + (size_t)n * k] * x[k];
    x[i] = s / G[i + (size_t)n * i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = x[i];
    for (int k = i + 1; k < n; ++k) s -= G[k + (size_t)n * i] * x[k];
    x[i] = s / G[i + (size_t)n * i];
  }
  for (int i = 0; i < m; ++i) w->where[i] = -1;
  int q = 0, it = 0;
  const int itmax = 50 * (m + n) + 100;
  for (;;) {
    /* step 1: most violated row */
    int pidx = -1;
    double smin = -TOL_VIOL;
    for (int i = 0; i < m; ++i) {
      if (w->where[i] >= 0) continue;
      double s = -b[i];
      for (int k = rp[i]; k < rp[i + 1]; ++k) s += cv[k] * x[ci[k]];
      if (s < smin) { smin = s; pidx = i; }
    }
    if (pidx < 0) break;
    double sp = smin;
    w->u[q] = 0.0;
    for (;;) {
      if (++it > itmax) { if (iters) *iters = it; return 1; }
      /* step 2a: d = J' n+,  z = J2 d2,  r = R^{-1} d1 */
      double dn2 = 0, d2n2 = 0;
      for (int c = 0; c < n; ++c) {
        const double* Jc = w->J + (size_t)n * c;
        double s = 0;
        for (int k = rp[pidx]; k < rp[pidx + 1]; ++k) s += cv[k] * Jc[ci[k]];
        w->d[c] = s;
        dn2 += s * s;
        if (c >= q) d2n2 += s * s;
      }
      int dependent = !(d2n2 > TOL_DEP * TOL_DEP * dn2);
      if (!dependent) {
        memset(w->z, 0, sizeof(double) * n);
        for (int c = q; c < n; ++c) {
          const double* Jc = w->J + (size_t)n * c;
          double dc = w->d[c];
          for (int i = 0; i < n; ++i) w->z[i] += Jc[i] * dc;
        }
      }
      for (int i = q - 1; i >= 0; --i) {
        double s = w->d[i];
        for (int k = i + 1; k < q; ++k) s -= w->R[i + (size_t)n * k] * w->r[k];
        w->r[i] = s / w->R[i + (size_t)n * i];
      }
      /* step 2b: step lengths */
      double t1 = ORC_INF; int l = -1;
      for (int j = 0; j < q; ++j)
        if (w->r[j] > 0.0) {
          double t = w->u[j] / w->r[j];
          if (t < t1) { t1 = t; l = j; }
        }
      double t2 = dependent ? ORC_INF : -sp / d2n2;
      if (t1 >= ORC_INF && t2 >= ORC_INF) { if (iters) *iters = it; return 2; } /* infeasible */
      if (t2 >= ORC_INF) {
        /* step in dual space only, drop blocking row l */
        for (int j = 0; j < q; ++j) w->u[j] -= t1 * w->r[j];
        w->u[q] += t1;
        gi_drop(w, &q, l);
        continue;
      }
      double t = t1 < t2 ? t1 : t2;
      for (int i = 0; i < n; ++i) x[i] += t * w->z[i];
      for (int j = 0; j < q; ++j) w->u[j] -= t * w->r[j];
      w->u[q] += t;
      if (t2 <= t1) {
        /* full step: add row p.  Rotate J2 so that d2 -> (|d2|, 0, ..., 0) */
        for (int k = n - 1; k > q; --k) {
          double a = w->d[k - 1], bb = w->d[k];
          if (bb == 0.0) continue;
          double h = hypot(a, bb), c = a / h, s = bb / h;
          w->d[k - 1] = h; w->d[k] = 0.0;
          rot_cols(n, w->J, k - 1, k, c, s);
        }
        memcpy(w->R + (size_t)n * q, w->d, sizeof(double) * (q + 1));
        w->A[q] = pidx; w->where[pidx] = q;
        q += 1;
        break;
      }
      /* partial step: drop blocking row l, recompute the violation of p */
      gi_drop(w, &q, l);
      sp = -b[pidx];
      for (int k = rp[pidx]; k < rp[pidx + 1]; ++k) sp += cv[k] * x[ci[k]];
    }
  }
  if (lambda) {
    for (int i = 0; i < m; ++i) lambda[i] = 0.0;
    for (int j = 0; j < q; ++j) lambda[w->A[j]] = w->u[j];
  }
  if (iters) *iters = it;
  return 0;
}

int orc_gi_dense(int n, int m, double* G, const double* g, const double* C, const double* b,
                 double* x, double* lambda, int* iters) {
  int* rp = (int*)malloc(sizeof(int) * (m + 1));
  int* ci = (int*)malloc(sizeof(int) * (size_t)(m > 0 ? m : 1) * n);
  double* cv = (double*)malloc(sizeof(double) * (size_t)(m > 0 ? m : 1) * n);
  int nnz = 0;
  for (int i = 0; i < m; ++i) {
    rp[i] = nnz;
    for (int j = 0; j < n; ++j)
      if (C[i + (size_t)m * j] != 0.0) { ci[nnz] = j; cv[nnz] = C[i + (size_t)m * j]; ++nnz; }
  }
  rp[m] = nnz;
  gi_ws* w = gi_ws_new(n, m);
  int rc = gi_core(n, G, g, m, rp, ci, cv, b, x, lambda, iters, w);
  gi_ws_free(w);
  free(rp); free(ci); free(cv);
  return rc;
}

/* ------------------------------------------------------------------ one problem */

/* bit index of a one-sided row, in the layout documented in include/lmpc.h */
static int bit_fric(int H, int k, int leg, int which) { (void)H; return 16 * k + 4 * leg + which; }
static int bit_lo(int H, int k, int leg) { return 16 * H + 4 * k + leg; }
static int bit_hi(int H, int k, int leg) { return 20 * H + 4 * k + leg; }

int orc_solve(const lmpc_params* p, const double x0[12], const double rot[9],
              const double foot[12], const double cmd[9], const uint8_t* contact,
              int eliminate_swing, double* U, double* grf, uint32_t* active, double* lambda,
              int* iters, double* kkt) {
  const int H = p->horizon, n = 12 * H, mfull = 24 * H;
  const double tol_act = p->active_tol > 0 ? p->active_tol : 1e-6;
  double* Hm = (double*)malloc(sizeof(double) * (size_t)n * n);
  double* g = (double*)malloc(sizeof(double) * n);
  condensed_qp_impl(p, x0, rot, foot, cmd, eliminate_swing ? contact : NULL, Hm, g, NULL);

  /* variable map */
  int* var = (int*)malloc(sizeof(int) * n);   /* reduced -> full */
  int* inv = (int*)malloc(sizeof(int) * n);   /* full -> reduced or -1 */
  int nr = 0;
  for (int k = 0; k < H; ++k)
    for (int leg = 0; leg < 4; ++leg)
      for (int a = 0; a < 3; ++a) {
        int full = 12 * k + 3 * leg + a;
        if (!eliminate_swing || contact[4 * k + leg]) { inv[full] = nr; var[nr++] = full; }
        else inv[full] = -1;
      }
  /* rows:  C u >= b  (one-sided), in bitmask order; rowbit[i] = bit index */
  int* rp = (int*)malloc(sizeof(int) * (mfull + 1));
  int* ci = (int*)malloc(sizeof(int) * 2 * mfull);
  double* cv = (double*)malloc(sizeof(double) * 2 * mfull);
  double* b = (double*)malloc(sizeof(double) * mfull);
  int* rowbit = (int*)malloc(sizeof(int) * mfull);
  int m = 0, nnz = 0;
  for (int k = 0; k < H; ++k)
    for (int leg = 0; leg < 4; ++leg) {
      int fx = inv[12 * k + 3 * leg], fy = inv[12 * k + 3 * leg + 1], fz = inv[12 * k + 3 * leg + 2];
      if (fx < 0) continue;
      double ub = (contact[4 * k + leg] ? 1.0 : 0.0) * p->fz_max;
      /* 0 <= fx + mu fz */
      rp[m] = nnz; ci[nnz] = fx; cv[nnz++] = 1;  ci[nnz] = fz; cv[nnz++] = p->mu; b[m] = 0; rowbit[m++] = bit_fric(H, k, leg, 0);
      /* fx - mu fz <= 0 */
      rp[m] = nnz; ci[nnz] = fx; cv[nnz++] = -1; ci[nnz] = fz; cv[nnz++] = p->mu; b[m] = 0; rowbit[m++] = bit_fric(H, k, leg, 1);
      rp[m] = nnz; ci[nnz] = fy; cv[nnz++] = 1;  ci[nnz] = fz; cv[nnz++] = p->mu; b[m] = 0; rowbit[m++] = bit_fric(H, k, leg, 2);
      rp[m] = nnz; ci[nnz] = fy; cv[nnz++] = -1; ci[nnz] = fz; cv[nnz++] = p->mu; b[m] = 0; rowbit[m++] = bit_fric(H, k, leg, 3);
      /* 0 <= fz */
      rp[m] = nnz; ci[nnz] = fz; cv[nnz++] = 1;  b[m] = 0; rowbit[m++] = bit_lo(H, k, leg);
      /* fz <= ub */
      rp[m] = nnz; ci[nnz] = fz; cv[nnz++] = -1; b[m] = -ub; rowbit[m++] = bit_hi(H, k, leg);
    }
  rp[m] = nnz;

  /* reduced Hessian / gradient */
  double* Gr = (double*)malloc(sizeof(double) * (size_t)(nr > 0 ? nr : 1) * (nr > 0 ? nr : 1));
  double* gr = (double*)malloc(sizeof(double) * (nr > 0 ? nr : 1));
  double* xr = (double*)calloc(nr > 0 ? nr : 1, sizeof(double));
  double* lam = (double*)calloc(mfull, sizeof(double));
  for (int c = 0; c < nr; ++c) {
    gr[c] = g[var[c]];
    for (int r = 0; r < nr; ++r) Gr[r + (size_t)nr * c] = Hm[var[r] + (size_t)n * var[c]];
  }
  int rc = 0, it = 0;
  if (nr > 0) {
    gi_ws* w = gi_ws_new(nr, m);
    rc = gi_core(nr, Gr, gr, m, rp, ci, cv, b, xr, lam, &it, w);
    gi_ws_free(w);
  }
  int finite = 1;
  for (int i = 0; i < nr; ++i) if (!isfinite(xr[i])) finite = 0;
  if (rc >= 2 || !finite) { rc = 2; memset(xr, 0, sizeof(double) * (nr > 0 ? nr : 1)); }

  for (int i = 0; i < n; ++i) U[i] = 0.0;
  for (int i = 0; i < nr; ++i) U[var[i]] = xr[i];
  if (grf) memcpy(grf, U, sizeof(double) * 12); /* solution.segment<12>(0)  (:320-322) */
  if (iters) *iters = it;

  /* primal-active bitmask over the FULL row set (swing legs included), by tolerance */
  if (active) {
    memset(active, 0, sizeof(uint32_t) * LMPC_ACTIVE_WORDS(H));
    for (int k = 0; k < H; ++k)
      for (int leg = 0; leg < 4; ++leg) {
        const double* f = U + 12 * k + 3 * leg;
        double ub = (contact[4 * k + leg] ? 1.0 : 0.0) * p->fz_max;
        double res[6] = { f[0] + p->mu * f[2], -(f[0] - p->mu * f[2]), f[1] + p->mu * f[2],
                          -(f[1] - p->mu * f[2]), f[2], ub - f[2] };
        int bits[6] = { bit_fric(H, k, leg, 0), bit_fric(H, k, leg, 1), bit_fric(H, k, leg, 2),
                        bit_fric(H, k, leg, 3), bit_lo(H, k, leg), bit_hi(H, k, leg) };
        for (int t = 0; t < 6; ++t)
          if (fabs(res[t]) <= tol_act) active[bits[t] >> 5] |= (1u << (bits[t] & 31));
      }
  }
  if (lambda) {
    memset(lambda, 0, sizeof(double) * mfull);
    for (int i = 0; i < m; ++i) lambda[rowbit[i]] = lam[i];
  }
  if (kkt) {
    /* KKT certificate on the reduced problem (exactly the full one when !eliminate_swing) */
    double stat = 0, gmax = 1, viol = 0, lmin = 0, comp = 0;
    for (int r = 0; r < nr; ++r) {
      double s = g[var[r]];
      for (int c = 0; c < nr; ++c) s += Hm[var[r] + (size_t)n * var[c]] * xr[c];
      for (int i = 0; i < m; ++i)
        for (int k = rp[i]; k < rp[i + 1]; ++k)
          if (ci[k] == r) s -= cv[k] * lam[i];
      if (fabs(s) > stat) stat = fabs(s);
      if (fabs(g[var[r]]) > gmax) gmax = fabs(g[var[r]]);
    }
    for (int i = 0; i < m; ++i) {
      double s = -b[i];
      for (int k = rp[i]; k < rp[i + 1]; ++k) s += cv[k] * xr[ci[k]];
      if (-s > viol) viol = -s;
      if (lam[i] < lmin) lmin = lam[i];
      if (fabs(lam[i] * s) > comp) comp = fabs(lam[i] * s);
    }
    kkt[0] = stat / gmax; kkt[1] = viol; kkt[2] = lmin; kkt[3] = comp;
  }
  free(Hm); free(g); free(var); free(inv); free(rp); free(ci); free(cv); free(b); free(rowbit);
  free(Gr); free(gr); free(xr); free(lam);
  return rc;
}

/* ------------------------------------------------------------------ batch driver */

typedef struct {
  const lmpc_params* p;
  const lmpc_batch_in* in;
  const lmpc_batch_out* out;
  double* kkt;
  int begin, end, eliminate, failures;
} batch_job;

static void problem_params(const lmpc_params* base, const lmpc_batch_in* in, int i, lmpc_params* pp) {
  *pp = *base;
  if (in->q) memcpy(pp->q, in->q + 12 * (size_t)i, sizeof(double) * 12);
  if (in->r) memcpy(pp->r, in->r + 12 * (size_t)i, sizeof(double) * 12);
  if (in->mu) pp->mu = in->mu[i];
  if (in->fz_max) pp->fz_max = in->fz_max[i];
  if (in->mass) pp->mass = in->mass[i];
  if (in->inertia) memcpy(pp->inertia, in->inertia + 9 * (size_t)i, sizeof(double) * 9);
}

static void* batch_worker(void* arg) {
  batch_job* job = (batch_job*)arg;
  const int H = job->p->horizon, W = LMPC_ACTIVE_WORDS(H);
  double* U = (double*)malloc(sizeof(double) * 12 * H);
  for (int i = job->begin; i < job->end; ++i) {
    lmpc_params pp;
    problem_params(job->p, job->in, i, &pp);
    double grf[12];
    int it = 0;
    int rc = orc_solve(&pp, job->in->x0 + 12 * (size_t)i, job->in->rot + 9 * (size_t)i,
                       job->in->foot_pos_abs + 12 * (size_t)i, job->in->cmd + 9 * (size_t)i,
                       job->in->contact + 4 * (size_t)H * i, job->eliminate, U, grf,
                       job->out->active_set ? job->out->active_set + (size_t)W * i : NULL, NULL,
                       &it, job->kkt ? job->kkt + 4 * (size_t)i : NULL);
    memcpy(job->out->grf + 12 * (size_t)i, grf, sizeof(grf));
    if (job->out->status) job->out->status[i] = rc == 0 ? LMPC_PROB_OK : (rc == 1 ? LMPC_PROB_MAX_ITER : LMPC_PROB_NON_FINITE);
    if (job->out->iters) job->out->iters[i] = it;
    if (job->out->U) memcpy(job->out->U + 12 * (size_t)H * i, U, sizeof(double) * 12 * H);
    if (rc) job->failures++;
  }
  free(U);
  return NULL;
}

int orc_solve_batch(const lmpc_params* p, int batch, const lmpc_batch_in* in,
                    const lmpc_batch_out* out, double* kkt, int eliminate_swing, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > batch) nthreads = batch > 0 ? batch : 1;
  batch_job* jobs = (batch_job*)calloc(nthreads, sizeof(batch_job));
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
  for (int t = 0; t < nthreads; ++t) {
    jobs[t].p = p; jobs[t].in = in; jobs[t].out = out; jobs[t].kkt = kkt;
    jobs[t].begin = (int)((long long)batch * t / nthreads);
    jobs[t].end = (int)((long long)batch * (t + 1) / nthreads);
    jobs[t].eliminate = eliminate_swing;
    if (nthreads == 1) batch_worker(&jobs[t]);
    else pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
  }
  int fails = 0;
  for (int t = 0; t < nthreads; ++t) {
    if (nthreads > 1) pthread_join(th[t], NULL);
    fails += jobs[t].failures;
  }
  free(jobs); free(th);
  return fails;
}

/* ------------------------------------------------------------------ gait schedule */

/* gait tables: per leg, up to 3 (state, switch_time) entries.  1 = STANCE, 0 = SWING
 * (enum LeggedContactState, LeggedContactFSM.h:11-15). */
typedef struct { int n; int state[3]; double sw[3]; } gait_leg;

static void gait_table(int gait, int leg, gait_leg* t) {
  switch (gait) {
    case LMPC_GAIT_TROT: /* set_default_gait_pattern, LeggedContactFSM.cpp:93-114 */
      t->n = 2; t->sw[0] = 0.5; t->sw[1] = 1.0;
      if (leg == 0 || leg == 3) { t->state[0] = 1; t->state[1] = 0; }
      else { t->state[0] = 0; t->state[1] = 1; }
      break;
    case LMPC_GAIT_TROT_WITH_STAND: /* :116-157 */
      if (leg == 0 || leg == 3) { t->n = 2; t->state[0] = 1; t->state[1] = 0; t->sw[0] = 0.6; t->sw[1] = 1.0; }
      else { t->n = 3; t->state[0] = 1; t->state[1] = 0; t->state[2] = 1; t->sw[0] = 0.1; t->sw[1] = 0.5; t->sw[2] = 1.0; }
      break;
    case LMPC_GAIT_CRAWL: /* :158-199 */
      if (leg == 0) { t->n = 2; t->state[0] = 0; t->state[1] = 1; t->sw[0] = 0.25; t->sw[1] = 1.0; }
      else if (leg == 1) { t->n = 3; t->state[0] = 1; t->state[1] = 0; t->state[2] = 1; t->sw[0] = 0.25; t->sw[1] = 0.5; t->sw[2] = 1.0; }
      else if (leg == 2) { t->n = 3; t->state[0] = 1; t->state[1] = 0; t->state[2] = 1; t->sw[0] = 0.5; t->sw[1] = 0.75; t->sw[2] = 1.0; }
      else { t->n = 2; t->state[0] = 1; t->state[1] = 0; t->sw[0] = 0.75; t->sw[1] = 1.0; }
      break;
    case LMPC_GAIT_STAND: /* :201-212 */
      t->n = 1; t->state[0] = 1; t->sw[0] = 1.0;
      break;
    case LMPC_GAIT_BOUND: /* addition: front pair then rear pair */
      t->n = 2; t->sw[0] = 0.5; t->sw[1] = 1.0;
      if (leg == 0 || leg == 1) { t->state[0] = 1; t->state[1] = 0; }
      else { t->state[0] = 0; t->state[1] = 1; }
      break;
    default: /* LMPC_GAIT_PRONK, addition: all legs in phase */
      t->n = 2; t->sw[0] = 0.5; t->sw[1] = 1.0; t->state[0] = 1; t->state[1] = 0;
      break;
  }
}

int orc_gait_predict_contact(int gait, int leg, double phase, double gait_speed, double delta_t) {
  gait_leg t;
  gait_table(gait, leg, &t);
  double ph = phase + gait_speed * delta_t;      /* LeggedContactFSM.cpp:281 */
  while (ph > 1.0) ph -= 1.0;                    /* :283-285 */
  for (int i = 0; i < t.n; ++i)                  /* :287-291 */
    if (ph <= t.sw[i]) return t.state[i];
  return 1;                                      /* :293 */
}
