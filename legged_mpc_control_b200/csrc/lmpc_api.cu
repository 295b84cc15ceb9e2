// lmpc_api.cu — kernels + C ABI of liblmpc.so (include/lmpc.h).  sm_100a only; there is no
// CPU path behind this ABI: every entry point that computes needs a CUDA device.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/lmpc.h"
#include "lmpc_solver.cuh"
#include "lmpc_warp_solve.cuh"
#include "lmpc_tile_solve.cuh"
#include "lmpc_rollout.cuh"

namespace lmpc {

// ---------------------------------------------------------------- kernels

// Size classes by reduced dimension n (3 per stance leg-step):
//   0: n <= 63            one warp per QP, 8 x 8 tiles        (solve_warp_kernel)
//   1: 64 <= n <= 120     one warp per QP, 15 x 15 tiles      (solve_tile_kernel<15>)
//   2: 121 <= n <= 192    one warp per QP, 24 x 24 tiles      (solve_tile_kernel<24>)
//   4: 193 <= n <= 288    4 warps per QP, 36 x 36 tiles       (solve_tile_kernel<36>), factor in the global slot
//   3: n > 288            4 warps per QP, 48 x 48 tiles       (solve_tile_kernel<48>); LMPC_CLS3=cta: scalar CTA kernel (solve_kernel)
// Register cap of the one-warp kernel, measured on B200 (headline workload, solves/s):
//   __launch_bounds__(32, 9): ptxas settles on 168 registers, 600 B of spills, 9 warps/SM   3.75 M
//   __maxnreg__(224):         290 B of spills, 9 warps/SM                                   3.90 M
//   __maxnreg__(255):         230 B of spills, 8 warps/SM                                   4.19 M
// The kernel is latency-bound inside each warp, so fewer spills beat the ninth warp.
#ifndef LMPC_WARP_MAXNREG
#define LMPC_WARP_MAXNREG 255
#endif
constexpr int kNumClasses = 5;
constexpr int kCtrCount = 8;  // counter block: work counters at [c], class sizes at [kCtrCount + c]
constexpr int kTileA = 15, kTileB = 24, kTileC = 48;  // 8 x 48 = 384 = 12 x LMPC_MAX_HORIZON
constexpr int kTileD = 36;  // class 4: 192 < n <= 288 (crawl at H = 30, stand at H = 20), same kernel as class 3
__host__ __device__ inline int size_class(int n) {
  return n <= 3 * WNBMAX ? 0 : (n <= 8 * kTileA ? 1 : (n <= 8 * kTileB ? 2 : (n <= 8 * kTileD ? 4 : 3)));
}

// Compacts the problems of every size class into its own index list (order + cls * B) with
// warp-aggregated atomics; counts[cls] = size of the class; n_max of class 3 sizes its shared memory.
__global__ void classify_kernel(const uint8_t* __restrict__ contact, int H, int B,
                                KStats* __restrict__ stats, int* __restrict__ order, int stride,
                                int* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int n = -1;
  if (i < B) {
    const uint8_t* c = contact + (size_t)(4 * H) * i;
    n = 0;
    for (int t = 0; t < 4 * H; ++t) n += c[t] ? 3 : 0;
  }
  const int cls = n >= 0 ? size_class(n) : -1;
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int k = 0; k < kNumClasses; ++k) {
    const unsigned m = __ballot_sync(0xffffffffu, cls == k);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(&counts[k], __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (cls == k) order[(size_t)k * stride + base + __popc(m & lt)] = i;
  }
  int nB = cls == 3 ? n : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nB = max(nB, __shfl_xor_sync(0xffffffffu, nB, o));
  if (lane == 0 && nB > 0) atomicMax(&stats->n_max, nB);
}

__device__ inline void flush_warp_stats(const WarpStats& ws, KStats* stats) {
  if ((threadIdx.x & 31) == 0 && ws.problems) {
    atomicAdd(&stats->problems, ws.problems);
    atomicAdd(&stats->sum_n, ws.sum_n);
    atomicAdd(&stats->sum_fact, ws.sum_fact);
    atomicAdd(&stats->sum_polish, ws.sum_polish);
    if (ws.not_ok) atomicAdd(&stats->not_ok, ws.not_ok);
    if (ws.ipm_only) atomicAdd(&stats->ipm_only, ws.ipm_only);
    if (ws.warm_hits) atomicAdd(&stats->warm_hits, ws.warm_hits);
    atomicAdd(&stats->flops, ws.flops);
  }
}

// Persistent solve for 64 <= n <= 8 NT, one QP per CTA of NW warps: fragment-ordered factor in shared
// memory (61 KB at NT = 15, 154 KB at NT = 24), Hessian slot in the L2 workspace.  Warp 0 runs the
// whole solve; warps 1 .. NW-1 only share the tile columns of every Cholesky factorization
// (tile_helper_loop), which is where n >= 120 spends its time.
template <int NT, int NW>
__global__ void __maxnreg__(255) solve_tile_kernel(const __grid_constant__ KParams P,
                                                             const __grid_constant__ KBatch Bt, KStats* stats,
                                                             double* gws, const int* __restrict__ count,
                                                             int* work_counter) {
  extern __shared__ __align__(16) unsigned char smem[];
  const TilePlan pl = make_tile_plan<NT>(P.H);
  const size_t kSlot = tile_slot_doubles<NT>();
  double* Hg = gws + (size_t)blockIdx.x * kSlot;
  if (NW > 1 && threadIdx.x >= 32) {
    tile_helper_loop<NT, NW>(P, pl, smem, Hg);
    return;
  }
  const int lane = threadIdx.x;
  const int total = *count;
  WarpStats ws;
  for (;;) {
    int w = 0;
    if (lane == 0) w = atomicAdd(work_counter, 1);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= total) break;
    solve_one_tile<NT, NW>(P, Bt, ws, pl, smem, Hg, Bt.order[w]);
  }
  flush_warp_stats(ws, stats);
  if (NW > 1) {
    if (lane == 0) *tile_job() = 0;
    __syncthreads();
  }
}

// Persistent one-warp-per-QP solve (class A).  One warp per CTA: no block barrier exists in this
// kernel.  Problems are handed out by an atomic counter; the scaled Hessian of the problem in
// flight lives in this CTA's 32 KB slot of the L2-resident workspace `gws`.
__global__ void __maxnreg__(LMPC_WARP_MAXNREG) solve_warp_kernel(const __grid_constant__ KParams P,
                                                        const __grid_constant__ KBatch Bt,
                                                        KStats* stats, double* gws,
                                                        const int* __restrict__ count,
                                                        int* work_counter) {
  extern __shared__ __align__(16) unsigned char smem[];
  const WarpPlan pl = make_warp_plan(P.H);
  double* Hg = gws + (size_t)blockIdx.x * (WN * WN);
  const int lane = threadIdx.x;
  const int total = *count;
  WarpStats ws;
  for (;;) {
    int w = 0;
    if (lane == 0) w = atomicAdd(work_counter, 1);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= total) break;
    solve_one_warp(P, Bt, ws, pl, smem, Hg, Bt.order[w]);
  }
  flush_warp_stats(ws, stats);
}

// Persistent solve: one CTA per QP at a time, problems handed out by an atomic counter.
template <int NT>
__global__ void __launch_bounds__(NT) solve_kernel(const __grid_constant__ KParams P,
                                                   const __grid_constant__ KBatch Bt,
                                                   KStats* stats, int ncap, int mats_in_smem,
                                                   double* gws, const int* __restrict__ count,
                                                   int* work_counter) {
  extern __shared__ __align__(16) unsigned char smem[];
  const SmemPlan pl = make_plan(P.H, ncap, mats_in_smem != 0);
  double* gH = nullptr;
  double* gM = nullptr;
  if (!mats_in_smem) {
    const size_t ps = (size_t)packed_size(ncap);
    gH = gws + (size_t)blockIdx.x * 2 * ps;
    gM = gH + ps;
  }
  int* si = reinterpret_cast<int*>(smem + pl.oINT);
  for (;;) {
    if (threadIdx.x == 0) si[INT_WORK] = atomicAdd(work_counter, 1);
    __syncthreads();
    const int w = si[INT_WORK];
    __syncthreads();
    if (w >= *count) break;
    const int pi = Bt.order[w];
    solve_one<NT>(P, Bt, stats, pl, smem, gH, gM, pi);
  }
}

}  // namespace lmpc

// ---------------------------------------------------------------- host side

using lmpc::KBatch;
using lmpc::KParams;
using lmpc::KStats;

static thread_local std::string g_err = "";
static void set_err(const std::string& s) { g_err = s; }
extern "C" const char* lmpc_last_error_string(void) { return g_err.c_str(); }
extern "C" int32_t lmpc_abi_version(void) { return LMPC_ABI_VERSION; }

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      set_err(std::string(#call) + ": " + cudaGetErrorString(e_));                            \
      return LMPC_ERR_CUDA;                                                                   \
    }                                                                                         \
  } while (0)

static constexpr int kPackMax = 64;  // batches up to this size take the packed host path
static constexpr int kThreads = 64;
static constexpr int kThreadsSpill = 256;
// Warps per QP of the tile classes, measured (C3 / reference preset, solves/s): NT=15: 598k / 647k / 519k at
// 1 / 2 / 4 warps (4 warps cost the third CTA per SM); NT=24: 90k / 106k / 146k (one CTA per SM anyway).
static constexpr int kTileWarpsA = 2, kTileWarpsB = 4;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct lmpc_handle {
  lmpc_params params;
  KParams kp;
  int device = 0;
  int max_batch = 0;
  int num_sms = 0;
  int smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // device staging for host-pointer calls
  DevBuf d_x0, d_rot, d_foot, d_cmd, d_contact, d_q, d_r, d_mu, d_fz, d_mass, d_inertia;
  DevBuf d_grf, d_status, d_iters, d_active, d_U;
  DevBuf d_stats, d_counter, d_ws, d_wsA, d_wsT1, d_wsT2, d_wsT3, d_order, d_warm;
  bool warm_enabled = false;
  int warm_shift = 1;
  // small-batch host path: one packed H2D and one packed D2H (a single robot's tick is latency-bound)
  DevBuf d_pin, d_pout;
  unsigned char* h_pin = nullptr;
  unsigned char* h_pout = nullptr;
  // closed-loop rollout state (lmpc_rollout)
  DevBuf d_rx, d_rfw, d_rph, d_rgait, d_rspeed, d_rtot, d_xlog, d_glog;
  cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;
  cudaEvent_t ev_cls = nullptr;
  int* h_cls = nullptr;  // pinned: n_max, |A|, |B|
  KStats last_stats;
  float last_ms = 0.f;
  int64_t launches = 0;
  bool pending = false;
  cudaStream_t pending_stream = nullptr;
};

static lmpc_status validate_params(const lmpc_params* p) {
  if (!p) { set_err("params is NULL"); return LMPC_ERR_INVALID_ARG; }
  if (p->horizon < 1 || p->horizon > LMPC_MAX_HORIZON) { set_err("horizon out of range [1, 32]"); return LMPC_ERR_INVALID_ARG; }
  if (!(p->dt > 0) || !(p->mu > 0) || !(p->fz_max > 0) || !(p->mass > 0)) { set_err("dt, mu, fz_max, mass must be > 0"); return LMPC_ERR_INVALID_ARG; }
  if (p->discretisation != LMPC_DISC_EULER && p->discretisation != LMPC_DISC_ZOH) { set_err("unknown discretisation"); return LMPC_ERR_INVALID_ARG; }
  if (p->cost_scaling != LMPC_COST_REF_HALF && p->cost_scaling != LMPC_COST_NS_TWO) { set_err("unknown cost_scaling"); return LMPC_ERR_INVALID_ARG; }
  if (p->precision != LMPC_PREC_F64) { set_err("precision: only LMPC_PREC_F64 is implemented in this build"); return LMPC_ERR_INVALID_ARG; }
  if (p->solver != LMPC_SOLVER_IPM && p->solver != LMPC_SOLVER_ACTIVE_SET) { set_err("unknown solver"); return LMPC_ERR_INVALID_ARG; }
  for (int i = 0; i < 12; ++i) {
    if (!(p->r[i] > 0)) { set_err("r weights must be > 0 (the condensed Hessian must be positive definite)"); return LMPC_ERR_INVALID_ARG; }
    if (!(p->q[i] >= 0)) { set_err("q weights must be >= 0"); return LMPC_ERR_INVALID_ARG; }
  }
  return LMPC_OK;
}

static void fill_kparams(const lmpc_params& p, KParams* k) {
  k->H = p.horizon;
  k->zoh = p.discretisation == LMPC_DISC_ZOH ? 1 : 0;
  k->xoff = p.xref_index_offset;
  k->max_iter = p.max_iter > 0 ? p.max_iter : 40;
  k->dt = p.dt; k->mu = p.mu; k->fz_max = p.fz_max; k->mass = p.mass; k->gravity = p.gravity;
  k->cs = p.cost_scaling == LMPC_COST_NS_TWO ? 2.0 : 1.0;
  k->act_tol = p.active_tol > 0 ? p.active_tol : 1e-6;
  // solver tuning (environment overrides are for experiments only)
  k->tol0 = 1e-7; k->rdtol0 = 1e-9; k->fz0 = 0.0;
  if (const char* e = getenv("LMPC_TOL0")) k->tol0 = atof(e);
  if (const char* e = getenv("LMPC_RDTOL0")) k->rdtol0 = atof(e);
  if (const char* e = getenv("LMPC_FZ0")) k->fz0 = atof(e);
  // ACTIVE_SET: every solve starts with the exact active-set iteration (stage 4b) — from the stored
  // faces when the problem has a warm-start record, from the all-interior face set otherwise — and
  // the interior point is only the fallback when its repair rounds do not verify.
  k->step_frac = 0.999; k->sigma_pow = 3;  // swept on the headline workload (tools/gpu_sweep2.sh): flat within 1 %, 0.999 leaves the fewest interior-point-only answers
  if (const char* e = getenv("LMPC_STEP_FRAC")) k->step_frac = atof(e);
  if (const char* e = getenv("LMPC_SIGMA_POW")) k->sigma_pow = atoi(e);
  k->try_interior = p.solver == LMPC_SOLVER_ACTIVE_SET ? 1 : 0;
  if (const char* e = getenv("LMPC_TRY_INTERIOR")) k->try_interior = atoi(e);
  for (int i = 0; i < 9; ++i) k->inertia[i] = p.inertia[i];
  for (int i = 0; i < 12; ++i) { k->q[i] = p.q[i]; k->r[i] = p.r[i]; }
}

extern "C" lmpc_status lmpc_set_warm_start(lmpc_handle* h, int32_t enable, int32_t horizon_shift);
extern "C" void lmpc_destroy(lmpc_handle* h);

extern "C" lmpc_status lmpc_create(const lmpc_params* params, int32_t max_batch, int32_t device,
                                   lmpc_handle** out) {
  if (!out) { set_err("out is NULL"); return LMPC_ERR_INVALID_ARG; }
  *out = nullptr;
  lmpc_status st = validate_params(params);
  if (st != LMPC_OK) return st;
  if (max_batch < 1) { set_err("max_batch must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_err(std::string("no CUDA device: this engine has no CPU fallback (") + cudaGetErrorString(e) + ")");
    return LMPC_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= ndev) { set_err("device ordinal out of range"); return LMPC_ERR_INVALID_ARG; }
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    set_err("device is not sm_100-class; liblmpc.so is built for sm_100a only");
    return LMPC_ERR_NO_DEVICE;
  }
  lmpc_handle* h = new lmpc_handle();
  h->params = *params;
  fill_kparams(*params, &h->kp);
  h->device = device;
  h->max_batch = max_batch;
  h->num_sms = prop.multiProcessorCount;
  h->smem_optin = (int)prop.sharedMemPerBlockOptin;
  CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&h->ev0));
  CK(cudaEventCreate(&h->ev1));
  CK(cudaEventCreateWithFlags(&h->ev_cls, cudaEventDisableTiming));
  CK(cudaMallocHost((void**)&h->h_cls, 4 * sizeof(int)));
  CK(h->d_stats.ensure(sizeof(KStats)));
  CK(h->d_counter.ensure(2 * lmpc::kCtrCount * sizeof(int)));  // work counters, class sizes
  CK(h->d_order.ensure(lmpc::kNumClasses * (size_t)max_batch * sizeof(int)));
  memset(&h->last_stats, 0, sizeof(KStats));
  *out = h;
  if (params->solver == LMPC_SOLVER_ACTIVE_SET) {  // the face records are this solver's state
    lmpc_status ws = lmpc_set_warm_start(h, 1, 1);
    if (ws != LMPC_OK) { lmpc_destroy(h); *out = nullptr; return ws; }
  }
  return LMPC_OK;
}

extern "C" void lmpc_destroy(lmpc_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  DevBuf* bufs[] = {&h->d_x0, &h->d_rot, &h->d_foot, &h->d_cmd, &h->d_contact, &h->d_q, &h->d_r, &h->d_mu,
                    &h->d_fz, &h->d_mass, &h->d_inertia, &h->d_grf, &h->d_status, &h->d_iters, &h->d_active,
                    &h->d_U, &h->d_stats, &h->d_counter, &h->d_ws, &h->d_wsA, &h->d_wsT1, &h->d_wsT2, &h->d_wsT3, &h->d_order, &h->d_warm,
                    &h->d_pin, &h->d_pout, &h->d_rx, &h->d_rfw, &h->d_rph, &h->d_rgait, &h->d_rspeed, &h->d_rtot, &h->d_xlog, &h->d_glog};
  for (DevBuf* b : bufs) b->release();
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev_cls) cudaEventDestroy(h->ev_cls);
  if (h->ev_r0) cudaEventDestroy(h->ev_r0);
  if (h->ev_r1) cudaEventDestroy(h->ev_r1);
  if (h->h_cls) cudaFreeHost(h->h_cls);
  if (h->h_pin) cudaFreeHost(h->h_pin);
  if (h->h_pout) cudaFreeHost(h->h_pout);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

template <int NT, int NW>
static lmpc_status launch_tile_nw(lmpc_handle* h, const KBatch& kb, cudaStream_t s, DevBuf* wsbuf, int* order, int* count,
                                  int* work) {
  const lmpc::TilePlan tp = lmpc::make_tile_plan<NT>(h->kp.H);
  auto kern = lmpc::solve_tile_kernel<NT, NW>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tp.total_bytes));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * NW, tp.total_bytes));
  if (occ < 1) { set_err("tile solve kernel does not fit on an SM"); return LMPC_ERR_CUDA; }
  int grid = h->num_sms * occ;
  if (grid > kb.B) grid = kb.B;
  const size_t slot = lmpc::tile_slot_doubles<NT>();
  CK(wsbuf->ensure((size_t)grid * slot * sizeof(double)));
  KBatch k = kb;
  k.order = order;
  kern<<<grid, 32 * NW, tp.total_bytes, s>>>(h->kp, k, (KStats*)h->d_stats.p, (double*)wsbuf->p, count, work);
  h->launches++;
  CK(cudaGetLastError());
  return LMPC_OK;
}
// Warps per QP of the tile classes (LMPC_TILE_NW overrides for experiments: 1, 2 or 4).
template <int NT>
static lmpc_status launch_tile(lmpc_handle* h, const KBatch& kb, cudaStream_t s, DevBuf* wsbuf, int* order, int* count,
                               int* work) {
  int nw = NT == lmpc::kTileA ? kTileWarpsA : kTileWarpsB;
  if (NT > lmpc::kTileB) return launch_tile_nw<NT, 4>(h, kb, s, wsbuf, order, count, work);
  if (kb.B <= h->num_sms) nw = 4;  // fewer problems than SMs: latency counts, occupancy does not
  if (const char* e = getenv("LMPC_TILE_NW")) nw = atoi(e);
  if (nw == 1) return launch_tile_nw<NT, 1>(h, kb, s, wsbuf, order, count, work);
  if (nw == 2) return launch_tile_nw<NT, 2>(h, kb, s, wsbuf, order, count, work);
  return launch_tile_nw<NT, 4>(h, kb, s, wsbuf, order, count, work);
}

// Launch classify + the per-size-class solves on `s` for the device-resident batch `kb`.  The
// one-warp-per-QP kernels are enqueued right behind the classification (they read their class
// size on the device); the host only waits for the 16-byte read-back that sizes the CTA-per-QP
// launch of the largest class, while the others run.
static lmpc_status launch_solve(lmpc_handle* h, KBatch kb, cudaStream_t s) {
  const int H = h->kp.H;
  kb.warm = h->warm_enabled ? (uint8_t*)h->d_warm.p : nullptr;
  kb.warm_shift = h->warm_shift;
  KStats* dstats = (KStats*)h->d_stats.p;
  int* ctr = (int*)h->d_counter.p;  // [c] work counters, [kCtrCount + c] class sizes
  int* cnt = ctr + lmpc::kCtrCount;
  int* order = (int*)h->d_order.p;  // class c at order + c * max_batch
  const int stride = h->max_batch;
  CK(cudaMemsetAsync(dstats, 0, sizeof(KStats), s));
  CK(cudaMemsetAsync(ctr, 0, 2 * lmpc::kCtrCount * sizeof(int), s));
  CK(cudaEventRecord(h->ev0, s));
  lmpc::classify_kernel<<<(kb.B + 255) / 256, 256, 0, s>>>(kb.contact, H, kb.B, dstats, order, stride, cnt);
  h->launches++;
  const bool big = 12 * H > 8 * lmpc::kTileB;
  static const bool cls3_cta = getenv("LMPC_CLS3") && !strcmp(getenv("LMPC_CLS3"), "cta");  // the scalar CTA kernel, kept for comparison
  const bool may3 = big && cls3_cta;
  if (may3) {
    CK(cudaMemcpyAsync(h->h_cls, &dstats->n_max, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(h->h_cls + 1, cnt + 3, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaEventRecord(h->ev_cls, s));
  }
  // ---- class 0: one warp per QP, 8 x 8 tiles
  {
    const lmpc::WarpPlan wp = lmpc::make_warp_plan(H);
    auto kernA = lmpc::solve_warp_kernel;
    CK(cudaFuncSetAttribute(kernA, cudaFuncAttributeMaxDynamicSharedMemorySize, wp.total_bytes));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernA, 32, wp.total_bytes));
    if (occ < 1) { set_err("warp solve kernel does not fit on an SM"); return LMPC_ERR_CUDA; }
    int grid = h->num_sms * occ;
    if (grid > kb.B) grid = kb.B;
    CK(h->d_wsA.ensure((size_t)grid * lmpc::WN * lmpc::WN * sizeof(double)));
    KBatch ka = kb;
    ka.order = order;
    kernA<<<grid, 32, wp.total_bytes, s>>>(h->kp, ka, dstats, (double*)h->d_wsA.p, cnt + 0, ctr + 0);
    h->launches++;
    CK(cudaGetLastError());
  }
  // ---- classes 1, 2: one warp per QP, 15 x 15 / 24 x 24 tiles (only if the horizon can reach them)
  if (12 * H > 3 * lmpc::WNBMAX) {
    lmpc_status st = launch_tile<lmpc::kTileA>(h, kb, s, &h->d_wsT1, order + stride, cnt + 1, ctr + 1);
    if (st != LMPC_OK) return st;
  }
  if (12 * H > 8 * lmpc::kTileA) {
    lmpc_status st = launch_tile<lmpc::kTileB>(h, kb, s, &h->d_wsT2, order + 2 * (size_t)stride, cnt + 2, ctr + 2);
    if (st != LMPC_OK) return st;
  }
  if (big && !cls3_cta) {  // class 3 (n > 192): 48 x 48 tiles, 4 warps per QP, factor in the global workspace slot
    lmpc_status st = launch_tile<lmpc::kTileC>(h, kb, s, &h->d_ws, order + 3 * (size_t)stride, cnt + 3, ctr + 3);
    if (st != LMPC_OK) return st;
  }
  if (big) {  // class 4 (192 < n <= 288): 36 x 36 tiles, otherwise as class 3
    lmpc_status st = launch_tile<lmpc::kTileD>(h, kb, s, &h->d_wsT3, order + 4 * (size_t)stride, cnt + 4, ctr + 4);
    if (st != LMPC_OK) return st;
  }
  if (may3) {
    CK(cudaEventSynchronize(h->ev_cls));
    const int n_max = h->h_cls[0], countB = h->h_cls[1];
    if (countB > 0) {
      // ---- class 3: one CTA per QP; choose where the two packed matrices live
      int ncap = n_max > 0 ? n_max : 3;
      lmpc::SmemPlan pl = lmpc::make_plan(H, ncap, true);
      int in_smem = 1;
      if (pl.total_bytes > h->smem_optin) {
        in_smem = 0;
        pl = lmpc::make_plan(H, ncap, false);
        if (pl.total_bytes > h->smem_optin) { set_err("shared-memory plan does not fit even with spilled matrices"); return LMPC_ERR_INVALID_ARG; }
      }
      // matrices in L2: latency bound, so more rows in flight per column (256 threads) pays
      int nthr = in_smem ? kThreads : kThreadsSpill;
      if (const char* e = getenv("LMPC_CLS3_THREADS")) nthr = atoi(e) == 64 ? 64 : (atoi(e) == 128 ? 128 : (atoi(e) == 512 ? 512 : 256));
      auto kern = nthr == 64 ? lmpc::solve_kernel<64> : (nthr == 128 ? lmpc::solve_kernel<128> : (nthr == 512 ? lmpc::solve_kernel<512> : lmpc::solve_kernel<256>));
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total_bytes));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, nthr, pl.total_bytes));
      if (occ < 1) { set_err("solve kernel does not fit on an SM"); return LMPC_ERR_CUDA; }
      int grid = h->num_sms * occ;
      if (grid > countB) grid = countB;
      double* gws = nullptr;
      if (!in_smem) {
        size_t bytes = (size_t)grid * 2 * lmpc::packed_size(ncap) * sizeof(double);
        CK(h->d_ws.ensure(bytes));
        gws = (double*)h->d_ws.p;
      }
      KBatch kbB = kb;
      kbB.order = order + 3 * (size_t)stride;
      kern<<<grid, nthr, pl.total_bytes, s>>>(h->kp, kbB, dstats, ncap, in_smem, gws, cnt + 3, ctr + 3);
      h->launches++;
      CK(cudaGetLastError());
    }
  }
  CK(cudaEventRecord(h->ev1, s));
  return LMPC_OK;
}

static lmpc_status finish_solve(lmpc_handle* h, cudaStream_t s) {
  CK(cudaMemcpyAsync(&h->last_stats, h->d_stats.p, sizeof(KStats), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  if (getenv("LMPC_DEBUG_STATS")) {
    const KStats& k = h->last_stats;
    const double np = k.problems ? (double)k.problems : 1.0;
    fprintf(stderr, "[lmpc] problems %llu mean_n %.2f fact %.3f polish %.3f ipm_only %llu not_ok %llu kernel %.3f ms\n", k.problems,
            k.sum_n / np, k.sum_fact / np, k.sum_polish / np, k.ipm_only, k.not_ok, h->last_ms);
  }
  if (h->last_stats.not_ok > 0) {
    set_err(std::to_string((unsigned long long)h->last_stats.not_ok) + " problem(s) did not converge (grf zeroed)");
    return LMPC_ERR_NOT_CONVERGED;
  }
  return LMPC_OK;
}

static lmpc_status check_io(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in, const lmpc_batch_out* out) {
  if (!h || !in || !out) { set_err("NULL handle/in/out"); return LMPC_ERR_INVALID_ARG; }
  if (batch < 1 || batch > h->max_batch) { set_err("batch must be in [1, max_batch]"); return LMPC_ERR_INVALID_ARG; }
  if (!in->x0 || !in->rot || !in->foot_pos_abs || !in->cmd || !in->contact) { set_err("x0, rot, foot_pos_abs, cmd, contact are required"); return LMPC_ERR_INVALID_ARG; }
  if (!out->grf) { set_err("out->grf is required"); return LMPC_ERR_INVALID_ARG; }
  if ((in->on_device != 0) != (out->on_device != 0)) { set_err("in and out must both be host or both be device pointers"); return LMPC_ERR_INVALID_ARG; }
  return LMPC_OK;
}

static KBatch device_batch(const lmpc_batch_in* in, const lmpc_batch_out* out, int batch) {
  KBatch kb;
  kb.x0 = in->x0; kb.rot = in->rot; kb.foot = in->foot_pos_abs; kb.cmd = in->cmd; kb.contact = in->contact;
  kb.q = in->q; kb.r = in->r; kb.mu = in->mu; kb.fzmax = in->fz_max; kb.mass = in->mass; kb.inertia = in->inertia;
  kb.grf = out->grf; kb.status = out->status; kb.iters = out->iters; kb.active = out->active_set; kb.U = out->U;
  kb.order = nullptr;
  kb.B = batch;
  kb.warm = nullptr; kb.warm_shift = 0;
  return kb;
}

extern "C" lmpc_status lmpc_solve_async(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                                        const lmpc_batch_out* out, void* cuda_stream) {
  lmpc_status st = check_io(h, batch, in, out);
  if (st != LMPC_OK) return st;
  if (!in->on_device) { set_err("lmpc_solve_async needs device pointers"); return LMPC_ERR_INVALID_ARG; }
  CK(cudaSetDevice(h->device));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  st = launch_solve(h, device_batch(in, out, batch), s);
  if (st != LMPC_OK) return st;
  h->pending = true;
  h->pending_stream = s;
  return LMPC_OK;
}

extern "C" lmpc_status lmpc_wait(lmpc_handle* h) {
  if (!h) { set_err("NULL handle"); return LMPC_ERR_INVALID_ARG; }
  if (!h->pending) return LMPC_OK;
  CK(cudaSetDevice(h->device));
  h->pending = false;
  return finish_solve(h, h->pending_stream);
}

extern "C" lmpc_status lmpc_solve(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                                  const lmpc_batch_out* out) {
  lmpc_status st = check_io(h, batch, in, out);
  if (st != LMPC_OK) return st;
  CK(cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  if (in->on_device) {
    st = launch_solve(h, device_batch(in, out, batch), s);
    if (st != LMPC_OK) return st;
    return finish_solve(h, s);
  }
  // ---- small batch, plain inputs: pack -> one H2D -> kernels -> one D2H -> unpack (saves ~8 driver
  //      calls; the reference's own use is ONE robot per tick)
  if (batch <= kPackMax && !in->q && !in->r && !in->mu && !in->fz_max && !in->mass && !in->inertia && !out->U) {
    const int Hh = h->kp.H, Wd = LMPC_ACTIVE_WORDS(Hh);
    const size_t nb = (size_t)batch;
    const size_t o_rot = nb * 96, o_foot = o_rot + nb * 72, o_cmd = o_foot + nb * 96, o_con = o_cmd + nb * 72;
    const size_t in_bytes = o_con + ((nb * 4 * Hh + 7) & ~(size_t)7);
    const size_t p_grf = (sizeof(KStats) + 15) & ~(size_t)15, p_st = p_grf + nb * 96, p_it = p_st + nb * 4, p_act = p_it + nb * 4;
    const size_t out_bytes = p_act + nb * 4 * Wd;
    if (!h->h_pin) {
      const size_t cap_in = (size_t)kPackMax * (96 + 72 + 96 + 72 + 4 * LMPC_MAX_HORIZON) + 64;
      const size_t cap_out = 64 + sizeof(KStats) + (size_t)kPackMax * (96 + 8 + 4 * LMPC_ACTIVE_WORDS(LMPC_MAX_HORIZON));
      CK(cudaMallocHost((void**)&h->h_pin, cap_in));
      CK(cudaMallocHost((void**)&h->h_pout, cap_out));
      CK(h->d_pin.ensure(cap_in));
      CK(h->d_pout.ensure(cap_out));
    }
    memcpy(h->h_pin, in->x0, nb * 96);
    memcpy(h->h_pin + o_rot, in->rot, nb * 72);
    memcpy(h->h_pin + o_foot, in->foot_pos_abs, nb * 96);
    memcpy(h->h_pin + o_cmd, in->cmd, nb * 72);
    memcpy(h->h_pin + o_con, in->contact, nb * 4 * Hh);
    CK(cudaMemcpyAsync(h->d_pin.p, h->h_pin, in_bytes, cudaMemcpyHostToDevice, s));
    unsigned char* di = (unsigned char*)h->d_pin.p;
    unsigned char* dout = (unsigned char*)h->d_pout.p;
    KBatch kb;
    memset(&kb, 0, sizeof(kb));
    kb.x0 = (const double*)di; kb.rot = (const double*)(di + o_rot); kb.foot = (const double*)(di + o_foot);
    kb.cmd = (const double*)(di + o_cmd); kb.contact = di + o_con;
    kb.grf = (double*)(dout + p_grf); kb.status = (int*)(dout + p_st); kb.iters = (int*)(dout + p_it);
    kb.active = (uint32_t*)(dout + p_act); kb.B = batch;
    void* saved = h->d_stats.p;  // the statistics block rides at the head of the packed output
    h->d_stats.p = dout;
    st = launch_solve(h, kb, s);
    h->d_stats.p = saved;
    if (st != LMPC_OK) return st;
    CK(cudaMemcpyAsync(h->h_pout, dout, out_bytes, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    memcpy(&h->last_stats, h->h_pout, sizeof(KStats));
    memcpy(out->grf, h->h_pout + p_grf, nb * 96);
    if (out->status) memcpy(out->status, h->h_pout + p_st, nb * 4);
    if (out->iters) memcpy(out->iters, h->h_pout + p_it, nb * 4);
    if (out->active_set) memcpy(out->active_set, h->h_pout + p_act, nb * 4 * Wd);
    CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    if (h->last_stats.not_ok > 0) {
      set_err(std::to_string((unsigned long long)h->last_stats.not_ok) + " problem(s) did not converge (grf zeroed)");
      return LMPC_ERR_NOT_CONVERGED;
    }
    return LMPC_OK;
  }
  // host pointers: stage through device buffers owned by the handle
  const int H = h->kp.H;
  const size_t B = (size_t)batch;
  const int W = LMPC_ACTIVE_WORDS(H);
  struct In { DevBuf* d; const void* src; size_t bytes; } ins[] = {
      {&h->d_x0, in->x0, B * 12 * 8}, {&h->d_rot, in->rot, B * 9 * 8}, {&h->d_foot, in->foot_pos_abs, B * 12 * 8},
      {&h->d_cmd, in->cmd, B * 9 * 8}, {&h->d_contact, in->contact, B * 4 * H}, {&h->d_q, in->q, B * 12 * 8},
      {&h->d_r, in->r, B * 12 * 8}, {&h->d_mu, in->mu, B * 8}, {&h->d_fz, in->fz_max, B * 8},
      {&h->d_mass, in->mass, B * 8}, {&h->d_inertia, in->inertia, B * 9 * 8}};
  for (auto& i : ins) {
    if (!i.src) continue;
    CK(i.d->ensure(i.bytes));
    CK(cudaMemcpyAsync(i.d->p, i.src, i.bytes, cudaMemcpyHostToDevice, s));
  }
  CK(h->d_grf.ensure(B * 12 * 8));
  if (out->status) CK(h->d_status.ensure(B * 4));
  if (out->iters) CK(h->d_iters.ensure(B * 4));
  if (out->active_set) CK(h->d_active.ensure(B * W * 4));
  if (out->U) CK(h->d_U.ensure(B * 12 * H * 8));
  KBatch kb;
  kb.x0 = (const double*)h->d_x0.p; kb.rot = (const double*)h->d_rot.p; kb.foot = (const double*)h->d_foot.p;
  kb.cmd = (const double*)h->d_cmd.p; kb.contact = (const uint8_t*)h->d_contact.p;
  kb.q = in->q ? (const double*)h->d_q.p : nullptr; kb.r = in->r ? (const double*)h->d_r.p : nullptr;
  kb.mu = in->mu ? (const double*)h->d_mu.p : nullptr; kb.fzmax = in->fz_max ? (const double*)h->d_fz.p : nullptr;
  kb.mass = in->mass ? (const double*)h->d_mass.p : nullptr; kb.inertia = in->inertia ? (const double*)h->d_inertia.p : nullptr;
  kb.grf = (double*)h->d_grf.p; kb.status = out->status ? (int*)h->d_status.p : nullptr;
  kb.iters = out->iters ? (int*)h->d_iters.p : nullptr; kb.active = out->active_set ? (uint32_t*)h->d_active.p : nullptr;
  kb.U = out->U ? (double*)h->d_U.p : nullptr; kb.order = nullptr; kb.B = batch;
  kb.warm = nullptr; kb.warm_shift = 0;
  st = launch_solve(h, kb, s);
  if (st != LMPC_OK) return st;
  CK(cudaMemcpyAsync(out->grf, kb.grf, B * 12 * 8, cudaMemcpyDeviceToHost, s));
  if (out->status) CK(cudaMemcpyAsync(out->status, kb.status, B * 4, cudaMemcpyDeviceToHost, s));
  if (out->iters) CK(cudaMemcpyAsync(out->iters, kb.iters, B * 4, cudaMemcpyDeviceToHost, s));
  if (out->active_set) CK(cudaMemcpyAsync(out->active_set, kb.active, B * W * 4, cudaMemcpyDeviceToHost, s));
  if (out->U) CK(cudaMemcpyAsync(out->U, kb.U, B * 12 * H * 8, cudaMemcpyDeviceToHost, s));
  return finish_solve(h, s);
}

extern "C" lmpc_status lmpc_set_warm_start(lmpc_handle* h, int32_t enable, int32_t horizon_shift) {
  if (!h) { set_err("NULL handle"); return LMPC_ERR_INVALID_ARG; }
  if (horizon_shift < 0 || horizon_shift > LMPC_MAX_HORIZON) { set_err("horizon_shift out of range"); return LMPC_ERR_INVALID_ARG; }
  CK(cudaSetDevice(h->device));
  if (enable && !h->d_warm.p) {
    const size_t bytes = (size_t)h->max_batch * lmpc::warm_stride(h->kp.H);
    CK(h->d_warm.ensure(bytes));
    CK(cudaMemsetAsync(h->d_warm.p, 0, bytes, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  h->warm_enabled = enable != 0;
  h->warm_shift = horizon_shift;
  return LMPC_OK;
}

extern "C" lmpc_status lmpc_reset_warm_start(lmpc_handle* h) {
  if (!h) { set_err("NULL handle"); return LMPC_ERR_INVALID_ARG; }
  if (!h->d_warm.p) return LMPC_OK;  // nothing is cached on the device
  CK(cudaSetDevice(h->device));
  CK(cudaMemsetAsync(h->d_warm.p, 0, (size_t)h->max_batch * lmpc::warm_stride(h->kp.H), h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return LMPC_OK;
}

extern "C" int64_t lmpc_last_warm_hits(const lmpc_handle* h) { return h ? (int64_t)h->last_stats.warm_hits : 0; }

// Page-locked host memory for callers without a CUDA runtime of their own (the ctypes mirror):
// host-pointer lmpc_solve calls copy at full PCIe rate only from/to pinned buffers.
extern "C" void* lmpc_alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { set_err("cudaMallocHost failed"); return nullptr; }
  return p;
}
extern "C" void lmpc_free_pinned(void* p) { if (p) cudaFreeHost(p); }

extern "C" int64_t lmpc_launch_count(const lmpc_handle* h) { return h ? h->launches : 0; }
extern "C" float lmpc_last_kernel_ms(const lmpc_handle* h) { return h ? h->last_ms : 0.f; }

extern "C" lmpc_status lmpc_last_stats(const lmpc_handle* h, double* mean_n_eff, double* mean_fact,
                                       double* flops_executed) {
  if (!h) { set_err("NULL handle"); return LMPC_ERR_INVALID_ARG; }
  const double np = h->last_stats.problems > 0 ? (double)h->last_stats.problems : 1.0;
  if (mean_n_eff) *mean_n_eff = (double)h->last_stats.sum_n / np;
  if (mean_fact) *mean_fact = (double)h->last_stats.sum_fact / np;
  if (flops_executed) *flops_executed = h->last_stats.flops;
  return LMPC_OK;
}

// ---------------------------------------------------------------- closed-loop rollout (config 5)

extern "C" lmpc_status lmpc_rollout(lmpc_handle* h, int32_t batch, const lmpc_rollout_cfg* cfg,
                                    const lmpc_rollout_state* st, lmpc_rollout_result* res) {
  if (!h || !cfg || !st || !res) { set_err("NULL handle/cfg/state/result"); return LMPC_ERR_INVALID_ARG; }
  if (batch < 1 || batch > h->max_batch) { set_err("batch must be in [1, max_batch]"); return LMPC_ERR_INVALID_ARG; }
  if (cfg->steps < 1) { set_err("steps must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  if (!st->x || !st->foot_world || !st->phase || !st->gait || !st->gait_speed || !st->cmd) {
    set_err("x, foot_world, phase, gait, gait_speed, cmd are required"); return LMPC_ERR_INVALID_ARG;
  }
  CK(cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  const int H = h->kp.H;
  const size_t B = (size_t)batch, S = (size_t)cfg->steps;
  const bool dev = st->on_device != 0;
  if (!h->ev_r0) { CK(cudaEventCreate(&h->ev_r0)); CK(cudaEventCreate(&h->ev_r1)); }
  if (cfg->warm_start) {
    lmpc_status w = lmpc_set_warm_start(h, 1, 1);
    if (w != LMPC_OK) return w;
    if (cfg->warm_start == 1) { w = lmpc_reset_warm_start(h); if (w != LMPC_OK) return w; }  // 2: keep the faces of the previous call
  } else {
    h->warm_enabled = false;
  }
  // solver input/output arrays owned by the handle
  CK(h->d_x0.ensure(B * 12 * 8)); CK(h->d_rot.ensure(B * 9 * 8)); CK(h->d_foot.ensure(B * 12 * 8));
  CK(h->d_contact.ensure(B * 4 * H)); CK(h->d_grf.ensure(B * 12 * 8));
  CK(h->d_rtot.ensure(sizeof(lmpc::RolloutTotals)));
  CK(cudaMemsetAsync(h->d_rtot.p, 0, sizeof(lmpc::RolloutTotals), s));
  lmpc::RolloutDev D;
  memset(&D, 0, sizeof(D));
  const double* d_cmd = st->cmd;
  if (dev) {
    D.x = st->x; D.foot_world = st->foot_world; D.phase = st->phase; D.gait = st->gait; D.gait_speed = st->gait_speed;
    D.x_log = res->x_log; D.grf_log = res->grf_log;
  } else {
    CK(h->d_rx.ensure(B * 12 * 8)); CK(h->d_rfw.ensure(B * 12 * 8)); CK(h->d_rph.ensure(B * 4 * 8));
    CK(h->d_rgait.ensure(B * 4)); CK(h->d_rspeed.ensure(B * 8)); CK(h->d_cmd.ensure(B * 9 * 8));
    CK(cudaMemcpyAsync(h->d_rx.p, st->x, B * 12 * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_rfw.p, st->foot_world, B * 12 * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_rph.p, st->phase, B * 4 * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_rgait.p, st->gait, B * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_rspeed.p, st->gait_speed, B * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_cmd.p, st->cmd, B * 9 * 8, cudaMemcpyHostToDevice, s));
    D.x = (double*)h->d_rx.p; D.foot_world = (double*)h->d_rfw.p; D.phase = (double*)h->d_rph.p;
    D.gait = (const int*)h->d_rgait.p; D.gait_speed = (const double*)h->d_rspeed.p;
    d_cmd = (const double*)h->d_cmd.p;
    if (res->x_log) { CK(h->d_xlog.ensure(S * B * 12 * 8)); D.x_log = (double*)h->d_xlog.p; }
    if (res->grf_log) { CK(h->d_glog.ensure(S * B * 12 * 8)); D.grf_log = (double*)h->d_glog.p; }
  }
  D.cmd = d_cmd;
  D.x0 = (double*)h->d_x0.p; D.rot = (double*)h->d_rot.p; D.foot_abs = (double*)h->d_foot.p;
  D.contact = (uint8_t*)h->d_contact.p; D.grf = (const double*)h->d_grf.p; D.B = batch;
  lmpc::RolloutCfgDev C;
  for (int i = 0; i < 12; ++i) C.foot_rel[i] = cfg->default_foot_pos_rel[i];
  C.dlim_x = cfg->foot_delta_x_limit > 0 ? cfg->foot_delta_x_limit : 0.8;  // FOOT_DELTA_X_LIMIT, LeggedParams.h:29
  C.dlim_y = cfg->foot_delta_y_limit > 0 ? cfg->foot_delta_y_limit : 0.8;  // FOOT_DELTA_Y_LIMIT, LeggedParams.h:30
  C.ground_z = cfg->ground_z;
  KBatch kb;
  memset(&kb, 0, sizeof(kb));
  kb.x0 = D.x0; kb.rot = D.rot; kb.foot = D.foot_abs; kb.cmd = d_cmd; kb.contact = D.contact;
  kb.grf = (double*)h->d_grf.p; kb.B = batch;
  const int tpb = 128, nblk = (batch + tpb - 1) / tpb;
  CK(cudaEventRecord(h->ev_r0, s));
  for (int step = 0; step < cfg->steps; ++step) {
    lmpc::rollout_prepare_kernel<<<nblk, tpb, 0, s>>>(h->kp, D, C);
    h->launches++;
    lmpc_status r = launch_solve(h, kb, s);
    if (r != LMPC_OK) return r;
    lmpc::rollout_step_kernel<<<nblk, tpb, 0, s>>>(h->kp, D, nullptr, nullptr, (const KStats*)h->d_stats.p,
                                                   (lmpc::RolloutTotals*)h->d_rtot.p, step);
    h->launches++;
  }
  CK(cudaGetLastError());
  CK(cudaEventRecord(h->ev_r1, s));
  lmpc::RolloutTotals tot;
  CK(cudaMemcpyAsync(&tot, h->d_rtot.p, sizeof(tot), cudaMemcpyDeviceToHost, s));
  if (res->grf) CK(cudaMemcpyAsync(res->grf, h->d_grf.p, B * 12 * 8, dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
  if (!dev) {
    CK(cudaMemcpyAsync(st->x, D.x, B * 12 * 8, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(st->foot_world, D.foot_world, B * 12 * 8, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(st->phase, D.phase, B * 4 * 8, cudaMemcpyDeviceToHost, s));
    if (res->x_log) CK(cudaMemcpyAsync(res->x_log, D.x_log, S * B * 12 * 8, cudaMemcpyDeviceToHost, s));
    if (res->grf_log) CK(cudaMemcpyAsync(res->grf_log, D.grf_log, S * B * 12 * 8, cudaMemcpyDeviceToHost, s));
  }
  CK(cudaStreamSynchronize(s));
  CK(cudaEventElapsedTime(&res->device_ms, h->ev_r0, h->ev_r1));
  res->solves = (int64_t)tot.solves; res->factorizations = (int64_t)tot.sum_fact; res->polish_rounds = (int64_t)tot.sum_polish;
  res->not_converged = (int64_t)tot.not_ok; res->warm_hits = (int64_t)tot.warm_hits; res->flops_executed = tot.flops;
  res->mean_n_eff = tot.solves ? (double)tot.sum_n / (double)tot.solves : 0.0;
  if (tot.not_ok > 0) {
    set_err(std::to_string((unsigned long long)tot.not_ok) + " solve(s) of the rollout did not converge (grf zeroed for that tick)");
    return LMPC_ERR_NOT_CONVERGED;
  }
  return LMPC_OK;
}
