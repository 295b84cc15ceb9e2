// lmpc_types.cuh — kernel-side parameter blocks shared by the CUDA translation units.
#pragma once
#include <stdint.h>

namespace lmpc {

// Device copy of the fields of lmpc_params the kernels need (include/lmpc.h).
struct KParams {
  int H;          // horizon
  int zoh;        // 1: zero-order hold, 0: forward Euler (reference)
  int xoff;       // xref_index_offset
  int max_iter;   // interior-point iteration cap
  double dt, mu, fz_max, mass, gravity;
  double cs;      // cost scaling (1: reference 1/2 z'Pz, 2: north-star)
  double act_tol; // primal activity tolerance [N]
  double tol0;    // complementarity target of the first interior-point stage (then: exact polish)
  double rdtol0;  // relative dual-residual target of the first stage
  double fz0;     // initial fz as a fraction of fz_max (<= 0: m g / #stance legs of the step)
  double inertia[9];
  double q[12];
  double r[12];
};

// Batch pointers (all device pointers).
struct KBatch {
  const double* x0;
  const double* rot;
  const double* foot;
  const double* cmd;
  const uint8_t* contact;
  const double* q;       // optional per-problem overrides (nullptr -> KParams)
  const double* r;
  const double* mu;
  const double* fzmax;
  const double* mass;
  const double* inertia;
  double* grf;
  int* status;           // may be nullptr
  int* iters;            // may be nullptr
  uint32_t* active;      // may be nullptr
  double* U;             // may be nullptr
  const int* order;      // optional problem permutation (size-class sorted), nullptr -> identity
  int B;
};

// Per-launch statistics accumulated with atomics (bench evidence).
struct KStats {
  unsigned long long problems;
  unsigned long long sum_n;        // sum of reduced dimensions
  unsigned long long sum_fact;     // sum of factorizations (IPM iterations + polish rounds)
  unsigned long long sum_polish;   // polish rounds only
  unsigned long long not_ok;       // problems with status >= 2
  unsigned long long ipm_only;     // problems with status == 1
  double flops;                    // executed factorization+solve flops (n^3/3 + 4n^2 per factorization etc.)
  int n_max;                       // max reduced dimension seen by the classify pass
};

}  // namespace lmpc
