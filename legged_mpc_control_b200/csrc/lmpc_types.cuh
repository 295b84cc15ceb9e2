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
  double step_frac; // fraction of the step to the boundary (0.995)
  double z0_scale;  // initial complementarity relative to |grad|_inf * mean slack (see lmpc_api.cu fill_kparams)
  double dual_rel;  // multiplier sign tolerance of the exact polish, relative to the gradient scale
  int sigma_pow;    // exponent of the Mehrotra centering rule (3)
  int try_interior; // experiment: cold problems first try the polish from all-interior faces
  int class0_on_tiles; // experiment: run size class 0 on solve_tile_kernel<8,1> (rolled code) instead of solve_warp_kernel
  int l2_persist;   // experiment: persisting-L2 access window over the one-warp kernel's Hessian workspace
  int warp_occ_cap; // experiment: cap on resident warps per SM of the one-warp kernel (0 = none)
  int tma_prefetch; // experiment: stage the NEXT problem's x0/foot records by TMA bulk copy while this one solves
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
  // Warm start (OSQP warm-start analogue, ConvexQPSolver.cpp:185): per problem 4H face bytes
  // (one per step x leg, see face_encode) + 4 header bytes (byte 4H: record valid).  nullptr -> cold.
  uint8_t* warm;
  int warm_shift;        // horizon steps elapsed since the stored solve (1 when control dt == MPC dt)
  // Problems a launch could not solve (status >= 2) are appended here instead of being counted as failures; the
  // host enqueues a second launch of the same kernel over this list with the conservative interior-point start
  // (KParams::z0_scale = 1).  nullptr -> this launch is final.
  int* retry_list;
  int* retry_count;
};

__host__ __device__ inline int warm_stride(int H) { return 4 * H + 4; }

// One byte per stance block: bit 7 valid, bits 0-1 sx+1, bits 2-3 sy+1, bit 4 top, bit 5 apex.
__host__ __device__ inline unsigned char face_encode(const signed char* f) {
  return (unsigned char)(0x80 | (f[0] + 1) | ((f[1] + 1) << 2) | ((f[2] & 1) << 4) | ((f[3] & 1) << 5));
}
// Face guess of block (step k, leg) from the record stored `shift` steps ago; interior if unknown.
__host__ __device__ inline void warm_face_load(const uint8_t* rec, int H, int shift, int k, int leg, signed char* f) {
  int kk = k + shift;
  if (kk > H - 1) kk = H - 1;
  const unsigned char c = rec[4 * kk + leg];
  if (c & 0x80) {
    f[0] = (signed char)((c & 3) - 1); f[1] = (signed char)(((c >> 2) & 3) - 1);
    f[2] = (signed char)((c >> 4) & 1); f[3] = (signed char)((c >> 5) & 1);
  } else {
    f[0] = 0; f[1] = 0; f[2] = 0; f[3] = 0;
  }
}

// Per-launch statistics accumulated with atomics (bench evidence).
struct KStats {
  unsigned long long problems;
  unsigned long long sum_n;        // sum of reduced dimensions
  unsigned long long sum_fact;     // sum of factorizations (IPM iterations + polish rounds)
  unsigned long long sum_polish;   // polish rounds only
  unsigned long long not_ok;       // problems with status >= 2
  unsigned long long ipm_only;     // problems with status == 1
  unsigned long long warm_hits;    // problems solved from the stored faces alone (no interior-point iteration)
  unsigned long long handed;       // problems the first pass handed to the retry pass (KBatch::retry_list)
  double flops;                    // executed factorization+solve flops (n^3/3 + 4n^2 per factorization etc.)
  int n_max;                       // max reduced dimension seen by the classify pass
};

}  // namespace lmpc
