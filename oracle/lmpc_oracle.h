/*
 * lmpc_oracle.h — CPU oracle for the convex-MPC hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (liblmpc.so) never links or calls it.
 *
 * PINNED TO THE REFERENCE'S OWN CODE (construction) + exact solve.  The reference holds no test, fixture
 * or golden vector for this path and its solve runs in OSQP + osqp-eigen, neither
 * vendored nor version-pinned (.devcontainer/Dockerfile:55,69), an ADMM method stopped at eps_abs=1e-3 /
 * eps_rel=1e-4 (ConvexQPSolver.cpp:183-184).  What IS pinned: oracle/_ref compiles the reference's
 * unmodified ConvexQPSolver.cpp, ConvexMpc.cpp, LeggedContactFSM.cpp, LeggedState.cpp, Utils.cpp against
 * header shims (oracle/shim) whose OsqpEigen stub captures the (P, q, A, l, u) handed to OSQP;
 * tests/test_ref_pin.py shows orc_sparse_qp equals that capture bit for bit (<= 4 eps on the I_w^-1 [r]x
 * entries) on 1024 states x 4 gaits, orc_gait_predict_contact equals predict_contact_state, and the exact
 * minimiser of the captured QP equals this oracle's answer to 1e-9.  The solve itself is an exact
 * Goldfarb-Idnani dual active-set method with a KKT certificate: parity is defined against the exact
 * minimiser of the reference's QP, which OSQP only approximates.
 */
#ifndef LMPC_ORACLE_H_
#define LMPC_ORACLE_H_

#include <stdint.h>
#include "../include/lmpc.h" /* lmpc_params and the batch layouts only; no code is shared */

#ifdef __cplusplus
extern "C" {
#endif

/* Continuous SRBD matrices and their discretisation for one step.
 * Follows ConvexQPSolver::update_A_matrix (ConvexQPSolver.cpp:214-228) and update_B_matrix
 * (:198-212), Utils::skew (utils/Utils.cpp:89-95).  Ad, Bd: 12x12 column-major; cd: affine
 * gravity term (12) so that x+ = Ad x + Bd u + cd.  EULER reproduces the reference exactly;
 * ZOH is exp([[Ac,Bc,c],[0,0,0]] dt) evaluated by its Taylor series. */
void orc_step_matrices(const lmpc_params* p, const double rot[9], const double foot[12],
                       double yaw_ref, double* Ad, double* Bd, double* cd);

/* x_ref(i), i = 0..H-1, as ConvexQPSolver.cpp:264-276 (with p->xref_index_offset added to i).
 * xref: [H][12]. */
void orc_reference_trajectory(const lmpc_params* p, const double x0[12], const double rot[9],
                              const double cmd[9], double* xref);

/* The reference's sparse QP, assembled densely:  min 1/2 z'diag(Pdiag)z + q'z, l <= A z <= u,
 * z = [u_0; x_1; ...; u_{H-1}; x_H] (24H), rows = 12H dynamics + 16H friction + 4H bounds
 * (ConvexQPSolver.cpp:16-196, 254-346).  A is (32H x 24H) column-major.  +-inf are +-1e30. */
void orc_sparse_qp(const lmpc_params* p, const double x0[12], const double rot[9],
                   const double foot[12], const double cmd[9], const uint8_t* contact,
                   double* Pdiag, double* q, double* A, double* l, double* u);

/* Condensed dense QP over U = [u_0..u_{H-1}] (n = 12H): Hm (n x n col-major), g (n), such that
 * cost = 1/2 U'Hm U + g'U (+const) in the scaling selected by p->cost_scaling.  Also returns
 * the free response minus reference e[H][12] = A_qp x0 + G - X_ref (may be NULL). */
void orc_condensed_qp(const lmpc_params* p, const double x0[12], const double rot[9],
                      const double foot[12], const double cmd[9], double* Hm, double* g,
                      double* e);

/* Exact solve of one problem.  eliminate_swing != 0 removes the variables of leg-steps with
 * contact == 0 (their bounds force f = 0) before the dual active-set solve; 0 keeps all 12H
 * variables and 24H one-sided rows, exactly the reference's constraint set.
 * Outputs (any may be NULL except U): U[12H], grf[12], active[LMPC_ACTIVE_WORDS(H)],
 * lambda[24H] (multipliers in the row order of the active-set bitmask), iters,
 * kkt[4] = {stationarity inf-norm / max(1,|g|inf), max primal violation, min multiplier,
 * max |lambda_i * residual_i|}.
 * Returns 0 on success, 1 if the solver hit its iteration cap, 2 if infeasible/non-finite
 * (U and grf are zeroed then, like ConvexQPSolver.cpp:321-326). */
int orc_solve(const lmpc_params* p, const double x0[12], const double rot[9],
              const double foot[12], const double cmd[9], const uint8_t* contact,
              int eliminate_swing, double* U, double* grf, uint32_t* active, double* lambda,
              int* iters, double* kkt);

/* Batch driver: one problem per thread over `nthreads` pthreads (static block partition).
 * in/out use the ABI's SoA layouts with host pointers.  kkt may be NULL or [B][4].
 * Returns the number of problems whose status != 0. */
int orc_solve_batch(const lmpc_params* p, int batch, const lmpc_batch_in* in,
                    const lmpc_batch_out* out, double* kkt, int eliminate_swing, int nthreads);

/* Generic dense Goldfarb-Idnani solver:  min 1/2 x'Gx + g'x  s.t.  C x >= b  (C: m x n
 * column-major).  G is overwritten by its Cholesky factor.  lambda[m] gets the multipliers.
 * Returns 0 ok, 1 iteration cap, 2 infeasible, 3 G not positive definite. */
int orc_gi_dense(int n, int m, double* G, const double* g, const double* C, const double* b,
                 double* x, double* lambda, int* iters);

/* LeggedContactFSM::predict_contact_state restated (LeggedContactFSM.cpp:280-294 with the
 * gait tables of :93-212; BOUND and PRONK are additions, see include/lmpc.h). */
int orc_gait_predict_contact(int gait, int leg, double phase, double gait_speed, double delta_t);

#ifdef __cplusplus
}
#endif
#endif
