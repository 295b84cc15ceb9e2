// lmpc_api.cu — kernels + C ABI of liblmpc.so (include/lmpc.h).  sm_100a only; there is no
// CPU path behind this ABI: every entry point that computes needs a CUDA device.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "lmpc_internal.h"
#include "lmpc_rollout.cuh"

namespace lmpc {

// ---------------------------------------------------------------- kernels

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

}  // namespace lmpc

// ---------------------------------------------------------------- host side

using lmpc::KBatch;
using lmpc::KParams;
using lmpc::KStats;

static thread_local std::string g_err = "";
namespace lmpc { void set_err(const std::string& s) { g_err = s; } }
using lmpc::set_err;
using lmpc::DevBuf;
extern "C" const char* lmpc_last_error_string(void) { return g_err.c_str(); }
extern "C" int32_t lmpc_abi_version(void) { return LMPC_ABI_VERSION; }

static constexpr int kPackMax = 64;  // batches up to this size take the packed host path

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
  k->tol0 = 1e-7; k->rdtol0 = 1e-7; k->fz0 = 0.0;  // hand-over to the polish: swept with the small-complementarity start (tools/gpu_r2_tol0.sh): rdtol0 1e-9 -> 1e-7 is +0.7 % (H=10) ... +4 % (H=30 trot)
  if (const char* e = getenv("LMPC_TOL0")) k->tol0 = atof(e);
  if (const char* e = getenv("LMPC_RDTOL0")) k->rdtol0 = atof(e);
  if (const char* e = getenv("LMPC_FZ0")) k->fz0 = atof(e);
  // ACTIVE_SET: every solve starts with the exact active-set iteration (stage 4b) — from the stored
  // faces when the problem has a warm-start record, from the all-interior face set otherwise — and
  // the interior point is only the fallback when its repair rounds do not verify.
  k->step_frac = 0.999; k->sigma_pow = 3;  // swept on the headline workload (tools/gpu_sweep2.sh): flat within 1 %, 0.999 leaves the fewest interior-point-only answers
  if (const char* e = getenv("LMPC_STEP_FRAC")) k->step_frac = atof(e);
  k->z0_scale = 1e-4;  // swept on every workload (tools/gpu_r2_z0.sh, profiles/r02_v12_z0_sweep.txt): 1 -> 1e-4 saves 2.5-3 factorizations per cold solve
  if (const char* e = getenv("LMPC_Z0_SCALE")) k->z0_scale = atof(e);
  if (const char* e = getenv("LMPC_SIGMA_POW")) k->sigma_pow = atoi(e);
  k->dual_rel = 1e-11;  // 1e-9 let multipliers of -1e-10 pass: amplified by cond(H) ~ 1e7 that moved late-horizon forces by 6e-6 relative (tests/test_gpu_parity_fullsize.py)
  if (const char* e = getenv("LMPC_DUAL_REL")) k->dual_rel = atof(e);
  k->try_interior = p.solver == LMPC_SOLVER_ACTIVE_SET ? 1 : 0;
  if (const char* e = getenv("LMPC_TRY_INTERIOR")) k->try_interior = atoi(e);
  k->class0_on_tiles = 0;
  if (const char* e = getenv("LMPC_CLASS0_TILES")) k->class0_on_tiles = atoi(e);
  k->l2_persist = 0;
  if (const char* e = getenv("LMPC_L2_PERSIST")) k->l2_persist = atoi(e);
  k->warp_occ_cap = 0;
  if (const char* e = getenv("LMPC_WARP_OCC")) k->warp_occ_cap = atoi(e);
  k->tma_prefetch = 0;
  if (const char* e = getenv("LMPC_TMA_PREFETCH")) k->tma_prefetch = atoi(e);
  for (int i = 0; i < 9; ++i) k->inertia[i] = p.inertia[i];
  for (int i = 0; i < 12; ++i) { k->q[i] = p.q[i]; k->r[i] = p.r[i]; }
}

extern "C" lmpc_status lmpc_set_warm_start(lmpc_handle* h, int32_t enable, int32_t horizon_shift);
extern "C" void lmpc_destroy(lmpc_handle* h);

static lmpc_status create_resources(lmpc_handle* h) {
  CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&h->ev0));
  CK(cudaEventCreate(&h->ev1));
  CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
  for (int c = 1; c < lmpc::kNumClasses; ++c) {
    CK(cudaStreamCreateWithFlags(&h->side[c], cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->ev_join[c], cudaEventDisableTiming));
  }
  CK(h->d_stats.ensure(sizeof(KStats)));
  CK(h->d_counter.ensure(4 * lmpc::kCtrCount * sizeof(int)));  // work counters, class sizes, and the same two for the retry launches
  CK(h->d_retry.ensure(lmpc::kNumClasses * (size_t)h->max_batch * sizeof(int)));
  CK(h->d_order.ensure(lmpc::kNumClasses * (size_t)h->max_batch * sizeof(int)));
  return LMPC_OK;
}

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
  if (device < 0 || device >= ndev || device >= lmpc::kMaxDevices) { set_err("device ordinal out of range"); return LMPC_ERR_INVALID_ARG; }
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
  h->kp_retry = h->kp;
  h->kp_retry.z0_scale = 1.0;
  h->device = device;
  h->max_batch = max_batch;
  h->num_sms = prop.multiProcessorCount;
  h->smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (const char* e = getenv("LMPC_TILE_NW")) h->tune.tile_nw = atoi(e);
  h->tune.debug_stats = getenv("LMPC_DEBUG_STATS") != nullptr;
  if (const char* e = getenv("LMPC_HOST_LANES")) h->tune.host_lanes = atoi(e);
  if (const char* e = getenv("LMPC_FORK_CLASSES")) h->tune.fork_classes = atoi(e) != 0;
  memset(&h->last_stats, 0, sizeof(KStats));
  if (h->kp.l2_persist) {
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)prop.persistingL2CacheMaxSize);
    if (h->tune.debug_stats)
      fprintf(stderr, "[lmpc] persisting L2: max %d MB, window max %d MB\n", prop.persistingL2CacheMaxSize >> 20, prop.accessPolicyMaxWindowSize >> 20);
  }
  st = create_resources(h);
  if (st == LMPC_OK && params->solver == LMPC_SOLVER_ACTIVE_SET) st = lmpc_set_warm_start(h, 1, 1);  // the face records are this solver's state
  if (st != LMPC_OK) {  // nothing leaks: the handle, its stream, events and buffers go through lmpc_destroy
    const std::string keep = g_err;
    lmpc_destroy(h);
    g_err = keep;
    return st;
  }
  *out = h;
  return LMPC_OK;
}

extern "C" void lmpc_destroy_sharded(lmpc_sharded* s);
extern "C" void lmpc_destroy(lmpc_handle* h) {
  if (!h) return;
  if (h->lanes) lmpc_destroy_sharded(h->lanes);
  h->lanes = nullptr;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  DevBuf* bufs[] = {&h->d_x0, &h->d_rot, &h->d_foot, &h->d_cmd, &h->d_contact, &h->d_q, &h->d_r, &h->d_mu,
                    &h->d_fz, &h->d_mass, &h->d_inertia, &h->d_grf, &h->d_status, &h->d_iters, &h->d_active,
                    &h->d_U, &h->d_stats, &h->d_counter, &h->d_ws, &h->d_wsA, &h->d_wsT1, &h->d_wsT2, &h->d_wsT3, &h->d_order, &h->d_warm, &h->d_retry,
                    &h->d_pin, &h->d_pout, &h->d_rx, &h->d_rfw, &h->d_rph, &h->d_rgait, &h->d_rspeed, &h->d_rtot, &h->d_xlog, &h->d_glog, &h->d_rfsm, &h->d_aux};
  for (DevBuf* b : bufs) b->release();
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev_r0) cudaEventDestroy(h->ev_r0);
  if (h->ev_r1) cudaEventDestroy(h->ev_r1);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  for (int c = 1; c < lmpc::kNumClasses; ++c) {
    if (h->side[c]) { cudaStreamSynchronize(h->side[c]); cudaStreamDestroy(h->side[c]); }
    if (h->ev_join[c]) cudaEventDestroy(h->ev_join[c]);
  }
  if (h->h_pin) cudaFreeHost(h->h_pin);
  if (h->h_pout) cudaFreeHost(h->h_pout);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

// Launch classify + the per-size-class solves on `s` for the device-resident batch `kb`.  Every
// solve kernel is enqueued right behind the classification and reads its class size on the device:
// no host synchronisation.  A class is launched only if the horizon can produce it (n <= 12 H).
// class_mask: bit c set = size class c may be non-empty (host-pointer calls know the contact rows and skip
// the launches of the classes no problem falls into; device-pointer calls pass ~0u).
// mode 0: both passes enqueued back to back (no host synchronisation: asynchronous calls, the closed loop);
//      1: the first pass only, failures are counted (the packed small-batch path looks at the count on the host and
//         calls again with mode 2);
//      2: one pass from the conservative start (also what a warm-started engine uses: its interior point only runs
//         for the few problems whose stored faces do not verify);
//      3: the first pass, failures handed to the per-class lists (KStats::handed says whether there are any);
//      4: the retry pass over those lists — no reset, no classification: continues the launch of mode 3.
// Blocking calls use 3, read the statistics, and enqueue 4 only if something was handed over: the retry launches
// are empty almost always, and an empty launch behind the persistent grids still cost ~75 us each
// (warm-started 65 536-problem batch: 4.20 -> 4.35 ms with two of them).
static lmpc_status launch_solve(lmpc_handle* h, KBatch kb, cudaStream_t s, unsigned class_mask = ~0u, int mode = 0) {
  const int H = h->kp.H;
  kb.warm = h->warm_enabled ? (uint8_t*)h->d_warm.p : nullptr;
  kb.warm_shift = h->warm_shift;
  KStats* dstats = (KStats*)h->d_stats.p;
  int* ctr = (int*)h->d_counter.p;  // [c] work counters, [kCtrCount + c] class sizes, then the retry launches' pair
  int* cnt = ctr + lmpc::kCtrCount;
  int* rctr = ctr + 2 * lmpc::kCtrCount;
  int* rcnt = ctr + 3 * lmpc::kCtrCount;
  int* order = (int*)h->d_order.p;  // class c at order + c * max_batch
  int* rlist = (int*)h->d_retry.p;  // same layout: what class c's first launch could not solve
  const size_t stride = (size_t)h->max_batch;
  if (mode == 0 && (h->warm_enabled || h->kp.z0_scale >= 1.0)) mode = 2;
  if (mode != 4) {
    CK(cudaMemsetAsync(dstats, 0, sizeof(KStats), s));
    CK(cudaMemsetAsync(ctr, 0, 4 * lmpc::kCtrCount * sizeof(int), s));
    CK(cudaEventRecord(h->ev0, s));
    lmpc::classify_kernel<<<(kb.B + 255) / 256, 256, 0, s>>>(kb.contact, H, kb.B, dstats, order, (int)stride, cnt);
    h->launches++;
  }
  const bool fork = h->tune.fork_classes;
  if (fork) CK(cudaEventRecord(h->ev_fork, s));  // everything the class kernels read is enqueued on `s` before this point
  const int nmax = 12 * H;  // largest reduced dimension this horizon can produce
  // Pass 0 solves a class from the small initial complementarity (kp.z0_scale); the few problems it cannot solve
  // are collected per class and pass 1 runs the same kernel over that list with the conservative start.  Both
  // passes are enqueued back to back on the class's stream (the list lengths live on the device: no host
  // synchronisation).  An empty launch behind a persistent grid costs ~75 us of stream time (not the kernel: it runs
  // 3.5 us; neither its grid nor the order of the launches changes that), which is why blocking calls enqueue pass 1
  // only on demand (modes 3 / 4) and why the classes run on forked streams.
  const int pass_lo = mode == 4 ? 1 : 0, pass_hi = (mode == 0 || mode == 4) ? 2 : 1;
  const bool lists = mode == 0 || mode == 3;  // does the first pass hand its failures over?
  auto run_class = [&](int cls, auto&& launcher) -> lmpc_status {
    const bool side = fork && cls > 0;
    cudaStream_t cs = side ? h->side[cls] : s;
    if (side) CK(cudaStreamWaitEvent(cs, h->ev_fork, 0));
    for (int pass = pass_lo; pass < pass_hi; ++pass) {
      const bool first = pass == 0;
      const lmpc::LaunchCtx c{h->device, (first && mode != 2) ? &h->kp : &h->kp_retry, dstats, h->num_sms, cs, &h->launches};
      KBatch k = kb;
      k.retry_list = (first && lists) ? rlist + (size_t)cls * stride : nullptr;
      k.retry_count = (first && lists) ? rcnt + cls : nullptr;
      if (!first && k.B > 256) k.B = 256;  // (B only caps the grid, and above the SM count it does not change the kernel variant) a handful of problems
      const lmpc_status st = launcher(c, k, (first ? order : rlist) + (size_t)cls * stride, (first ? cnt : rcnt) + cls, (first ? ctr : rctr) + cls);
      if (st != LMPC_OK) return st;
    }
    if (side) {  // join: the call's stream (and its end event) waits for this class
      CK(cudaEventRecord(h->ev_join[cls], cs));
      CK(cudaStreamWaitEvent(s, h->ev_join[cls], 0));
    }
    return LMPC_OK;
  };
  using lmpc::LaunchCtx;
  lmpc_status st = LMPC_OK;
  if (class_mask & 1u)
    st = run_class(0, [&](const LaunchCtx& c, const KBatch& k, const int* lst, const int* len, int* wrk) {
      return h->kp.class0_on_tiles ? lmpc::launch_tile8(c, k, &h->d_wsA, lst, len, wrk, 1) : lmpc::launch_warp_class(c, k, &h->d_wsA, lst, len, wrk);
    });
  if (st != LMPC_OK) return st;
  if (nmax > lmpc::kWarpNMax && (class_mask & 2u))
    st = run_class(1, [&](const LaunchCtx& c, const KBatch& k, const int* lst, const int* len, int* wrk) {
      return lmpc::launch_tile15(c, k, &h->d_wsT1, lst, len, wrk, h->tune.tile_nw);
    });
  if (st != LMPC_OK) return st;
  if (nmax > 8 * lmpc::kTileA && (class_mask & 4u))
    st = run_class(2, [&](const LaunchCtx& c, const KBatch& k, const int* lst, const int* len, int* wrk) {
      return lmpc::launch_tile24(c, k, &h->d_wsT2, lst, len, wrk, h->tune.tile_nw);
    });
  if (st != LMPC_OK) return st;
  if (nmax > 8 * lmpc::kTileB && (class_mask & 16u))  // class 4 (192 < n <= 288)
    st = run_class(4, [&](const LaunchCtx& c, const KBatch& k, const int* lst, const int* len, int* wrk) {
      return lmpc::launch_tile36(c, k, &h->d_wsT3, lst, len, wrk, 0);
    });
  if (st != LMPC_OK) return st;
  if (nmax > 8 * lmpc::kTileD && (class_mask & 8u))  // class 3 (n > 288): only horizons above 24 reach it
    st = run_class(3, [&](const LaunchCtx& c, const KBatch& k, const int* lst, const int* len, int* wrk) {
      return lmpc::launch_tile48(c, k, &h->d_ws, lst, len, wrk, 0);
    });
  if (st != LMPC_OK) return st;
  CK(cudaEventRecord(h->ev1, s));
  return LMPC_OK;
}

// Blocking two-pass solve: first pass, look at the statistics, retry pass only if something was handed over.
// Leaves the kernels finished or enqueued on `s`; the caller copies results out and calls finish_solve.
static lmpc_status solve_passes_blocking(lmpc_handle* h, const KBatch& kb, cudaStream_t s, unsigned class_mask = ~0u) {
  if (h->warm_enabled || h->kp.z0_scale >= 1.0) return launch_solve(h, kb, s, class_mask, 2);
  lmpc_status st = launch_solve(h, kb, s, class_mask, 3);
  if (st != LMPC_OK) return st;
  CK(cudaMemcpyAsync(&h->last_stats, h->d_stats.p, sizeof(KStats), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  if (h->last_stats.handed > 0) st = launch_solve(h, kb, s, class_mask, 4);
  return st;
}

static lmpc_status finish_solve(lmpc_handle* h, cudaStream_t s) {
  CK(cudaMemcpyAsync(&h->last_stats, h->d_stats.p, sizeof(KStats), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
  if (h->tune.debug_stats) {
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

// Slices smaller than this are not worth a launch of their own (a few waves of the persistent grid).
static constexpr int kLaneMinSlice = 2048;
static lmpc_status solve_host_lanes(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in, const lmpc_batch_out* out);

extern "C" lmpc_status lmpc_solve(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                                  const lmpc_batch_out* out) {
  lmpc_status st = check_io(h, batch, in, out);
  if (st != LMPC_OK) return st;
  CK(cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  if (in->on_device) {
    st = solve_passes_blocking(h, device_batch(in, out, batch), s);
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
    unsigned mask = 0u;  // a single robot's tick launches classify + ONE solve kernel instead of up to five
    for (size_t i = 0; i < nb; ++i) {
      int n = 0;
      for (int t = 0; t < 4 * Hh; ++t) n += in->contact[i * 4 * Hh + t] ? 3 : 0;
      mask |= 1u << lmpc::size_class(n);
    }
    void* saved = h->d_stats.p;  // the statistics block rides at the head of the packed output
    h->d_stats.p = dout;
    st = launch_solve(h, kb, s, mask, 1);
    if (st != LMPC_OK) { h->d_stats.p = saved; return st; }
    CK(cudaMemcpyAsync(h->h_pout, dout, out_bytes, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    memcpy(&h->last_stats, h->h_pout, sizeof(KStats));
    if (h->last_stats.not_ok > 0 && h->kp.z0_scale < 1.0) {  // rare: once more from the conservative interior-point start
      st = launch_solve(h, kb, s, mask, 2);
      if (st != LMPC_OK) { h->d_stats.p = saved; return st; }
      CK(cudaMemcpyAsync(h->h_pout, dout, out_bytes, cudaMemcpyDeviceToHost, s));
      CK(cudaStreamSynchronize(s));
      memcpy(&h->last_stats, h->h_pout, sizeof(KStats));
    }
    h->d_stats.p = saved;
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
  // ---- large batch, host pointers: slices in flight on this device (copies under the neighbouring slice's solve)
  if (!h->is_lane && h->tune.host_lanes > 1 && !h->warm_enabled && batch >= kLaneMinSlice * h->tune.host_lanes)
    return solve_host_lanes(h, batch, in, out);
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
  st = solve_passes_blocking(h, kb, s);
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
  // portable: pinned for EVERY device's context (lmpc_solve_sharded copies one array from several GPUs)
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { set_err("cudaHostAlloc failed"); return nullptr; }
  return p;
}
extern "C" void lmpc_free_pinned(void* p) { if (p) cudaFreeHost(p); }

extern "C" int64_t lmpc_last_not_converged(const lmpc_handle* h) { return h ? (int64_t)h->last_stats.not_ok : 0; }

// ---------------------------------------------------------------- one process, several GPUs
struct lmpc_sharded {
  lmpc_params params;
  int per = 0;  // capacity of every shard
  std::vector<lmpc_handle*> eng;
  int64_t last_not_ok = 0;
};

extern "C" void lmpc_destroy_sharded(lmpc_sharded* s) {
  if (!s) return;
  for (lmpc_handle* h : s->eng) lmpc_destroy(h);
  delete s;
}

extern "C" lmpc_status lmpc_create_sharded(const lmpc_params* params, int32_t max_batch, const int32_t* devices,
                                           int32_t n_devices, lmpc_sharded** out) {
  if (!out) { set_err("out is NULL"); return LMPC_ERR_INVALID_ARG; }
  *out = nullptr;
  if (!devices || n_devices < 1) { set_err("devices[] with n_devices >= 1 is required"); return LMPC_ERR_INVALID_ARG; }
  if (max_batch < 1) { set_err("max_batch must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  if (!params) { set_err("params is NULL"); return LMPC_ERR_INVALID_ARG; }
  lmpc_sharded* s = new lmpc_sharded();
  s->params = *params;
  s->per = (max_batch + n_devices - 1) / n_devices;  // ceil: floor-partitioned shards never exceed it
  for (int g = 0; g < n_devices; ++g) {
    lmpc_handle* h = nullptr;
    const lmpc_status st = lmpc_create(params, s->per, devices[g], &h);
    if (st != LMPC_OK) {
      const std::string keep = g_err;
      lmpc_destroy_sharded(s);
      g_err = keep;
      return st;
    }
    s->eng.push_back(h);
  }
  *out = s;
  return LMPC_OK;
}

extern "C" int32_t lmpc_sharded_num_devices(const lmpc_sharded* s) { return s ? (int32_t)s->eng.size() : 0; }
extern "C" lmpc_handle* lmpc_sharded_handle(lmpc_sharded* s, int32_t shard) {
  return (s && shard >= 0 && shard < (int32_t)s->eng.size()) ? s->eng[shard] : nullptr;
}
extern "C" int64_t lmpc_sharded_last_not_converged(const lmpc_sharded* s) { return s ? s->last_not_ok : 0; }

extern "C" lmpc_status lmpc_set_warm_start_sharded(lmpc_sharded* s, int32_t enable, int32_t horizon_shift) {
  if (!s) { set_err("NULL handle"); return LMPC_ERR_INVALID_ARG; }
  for (lmpc_handle* h : s->eng) {
    const lmpc_status st = lmpc_set_warm_start(h, enable, horizon_shift);
    if (st != LMPC_OK) return st;
  }
  return LMPC_OK;
}

extern "C" lmpc_status lmpc_solve_sharded(lmpc_sharded* s, int32_t batch, const lmpc_batch_in* in,
                                          const lmpc_batch_out* out) {
  if (!s || !in || !out) { set_err("NULL handle/in/out"); return LMPC_ERR_INVALID_ARG; }
  if (in->on_device || out->on_device) { set_err("lmpc_solve_sharded takes host pointers"); return LMPC_ERR_INVALID_ARG; }
  const int G = (int)s->eng.size();
  // every argument check happens BEFORE the first thread starts: nothing can throw or return with
  // joinable threads behind it
  if (batch < 1 || (long long)batch > (long long)s->per * G) { set_err("batch must be in [1, max_batch]"); return LMPC_ERR_INVALID_ARG; }
  if (!in->x0 || !in->rot || !in->foot_pos_abs || !in->cmd || !in->contact || !out->grf) {
    set_err("x0, rot, foot_pos_abs, cmd, contact and out->grf are required"); return LMPC_ERR_INVALID_ARG;
  }
  const int H = s->params.horizon, W = LMPC_ACTIVE_WORDS(H);
  auto begin = [&](int g) { return (int)(((long long)batch * g) / G); };
  for (int g = 0; g < G; ++g)
    if (begin(g + 1) - begin(g) > s->per) { set_err("shard larger than its engine (internal)"); return LMPC_ERR_INVALID_ARG; }
  std::vector<lmpc_status> rc(G, LMPC_OK);
  std::vector<std::string> err(G);
  std::vector<std::thread> th;
  th.reserve(G);
  for (int g = 0; g < G; ++g) {
    const int b0 = begin(g), b1 = begin(g + 1);
    if (b1 <= b0) continue;
    th.emplace_back([&, g, b0, b1] {
      lmpc_batch_in i = *in;
      lmpc_batch_out o = *out;
      const size_t k = (size_t)b0;
      i.x0 += 12 * k; i.rot += 9 * k; i.foot_pos_abs += 12 * k; i.cmd += 9 * k; i.contact += (size_t)(4 * H) * k;
      if (i.q) i.q += 12 * k;
      if (i.r) i.r += 12 * k;
      if (i.mu) i.mu += k;
      if (i.fz_max) i.fz_max += k;
      if (i.mass) i.mass += k;
      if (i.inertia) i.inertia += 9 * k;
      o.grf += 12 * k;
      if (o.status) o.status += k;
      if (o.iters) o.iters += k;
      if (o.active_set) o.active_set += (size_t)W * k;
      if (o.U) o.U += (size_t)(12 * H) * k;
      rc[g] = lmpc_solve(s->eng[g], b1 - b0, &i, &o);
      if (rc[g] != LMPC_OK) err[g] = lmpc_last_error_string();  // thread-local: copy it out of the worker
    });
  }
  for (auto& t : th) t.join();
  s->last_not_ok = 0;
  lmpc_status worst = LMPC_OK;
  for (int g = 0; g < G; ++g) {
    s->last_not_ok += lmpc_last_not_converged(s->eng[g]);
    if (rc[g] == LMPC_OK) continue;
    if (rc[g] != LMPC_ERR_NOT_CONVERGED || worst == LMPC_OK) { worst = rc[g]; set_err("shard " + std::to_string(g) + ": " + err[g]); }
    if (rc[g] != LMPC_ERR_NOT_CONVERGED) break;
  }
  return worst;
}

// Large host-pointer batch on ONE device: the batch is cut into contiguous slices, every slice owned by an
// engine of its own on h->device (stream, staging buffers, workspace, counters), all driven through
// lmpc_solve_sharded.  Slice k+1's H2D copies and slice k-1's D2H copies run on the copy engines while slice
// k solves, and the persistent grid of slice k+1 takes over the SMs CTA by CTA as slice k's last problems
// finish, so the tail of one slice is covered by the head of the next.  Same answers as the one-shot path
// (every problem is independent and solved by the same kernels).
static lmpc_status solve_host_lanes(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in, const lmpc_batch_out* out) {
  if (!h->lanes) {
    std::vector<int32_t> dev((size_t)h->tune.host_lanes, (int32_t)h->device);
    lmpc_sharded* s = nullptr;
    const lmpc_status st = lmpc_create_sharded(&h->params, h->max_batch, dev.data(), (int32_t)dev.size(), &s);
    if (st != LMPC_OK) return st;
    for (lmpc_handle* e : s->eng) { e->is_lane = true; e->tune = h->tune; e->tune.host_lanes = 0; }
    h->lanes = s;
  }
  lmpc_sharded* s = h->lanes;
  const lmpc_status st = lmpc_solve_sharded(s, batch, in, out);
  KStats tot;
  memset(&tot, 0, sizeof(tot));
  float ms = 0.f;
  for (const lmpc_handle* e : s->eng) {
    const KStats& k = e->last_stats;
    tot.problems += k.problems; tot.sum_n += k.sum_n; tot.sum_fact += k.sum_fact; tot.sum_polish += k.sum_polish;
    tot.not_ok += k.not_ok; tot.ipm_only += k.ipm_only; tot.warm_hits += k.warm_hits; tot.flops += k.flops;
    tot.n_max = k.n_max > tot.n_max ? k.n_max : tot.n_max;
    ms += e->last_ms;
  }
  h->last_stats = tot;
  h->last_ms = ms;  // sum of the slices' device spans (they run back to back)
  if (st == LMPC_ERR_NOT_CONVERGED)
    set_err(std::to_string((unsigned long long)tot.not_ok) + " problem(s) did not converge (grf zeroed)");
  return st;
}

extern "C" int64_t lmpc_launch_count(const lmpc_handle* h) {
  if (!h) return 0;
  int64_t n = h->launches;
  if (h->lanes) for (const lmpc_handle* e : h->lanes->eng) n += e->launches;
  return n;
}
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
  // the rollout's warm-start mode is local to this call: the handle's own setting (lmpc_set_warm_start,
  // or the ACTIVE_SET solver's) is put back on every exit path
  struct WarmGuard {
    lmpc_handle* h; bool en; int shift;
    ~WarmGuard() { h->warm_enabled = en; h->warm_shift = shift; }
  } warm_guard{h, h->warm_enabled, h->warm_shift};
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
  const bool fsm = st->fsm_state || st->fsm_index || st->fsm_start || st->fsm_end;
  if (fsm && !(st->fsm_state && st->fsm_index && st->fsm_start && st->fsm_end)) {
    set_err("fsm_state, fsm_index, fsm_start, fsm_end: give all four or none"); return LMPC_ERR_INVALID_ARG;
  }
  if (dev) {
    D.x = st->x; D.foot_world = st->foot_world; D.phase = st->phase; D.gait = st->gait; D.gait_speed = st->gait_speed;
    D.x_log = res->x_log; D.grf_log = res->grf_log;
    D.fsm_state = st->fsm_state; D.fsm_index = st->fsm_index; D.fsm_start = st->fsm_start; D.fsm_end = st->fsm_end;
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
    if (fsm) {  // [state | index | start | end], 4 B entries each
      CK(h->d_rfsm.ensure(B * 4 * (4 + 4 + 8 + 8)));
      unsigned char* f = (unsigned char*)h->d_rfsm.p;
      D.fsm_state = (int*)f; D.fsm_index = (int*)(f + B * 16); D.fsm_start = (double*)(f + B * 32); D.fsm_end = (double*)(f + B * 64);
      CK(cudaMemcpyAsync(D.fsm_state, st->fsm_state, B * 16, cudaMemcpyHostToDevice, s));
      CK(cudaMemcpyAsync(D.fsm_index, st->fsm_index, B * 16, cudaMemcpyHostToDevice, s));
      CK(cudaMemcpyAsync(D.fsm_start, st->fsm_start, B * 32, cudaMemcpyHostToDevice, s));
      CK(cudaMemcpyAsync(D.fsm_end, st->fsm_end, B * 32, cudaMemcpyHostToDevice, s));
    }
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
    if (fsm) {
      CK(cudaMemcpyAsync(st->fsm_state, D.fsm_state, B * 16, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(st->fsm_index, D.fsm_index, B * 16, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(st->fsm_start, D.fsm_start, B * 32, cudaMemcpyDeviceToHost, s));
      CK(cudaMemcpyAsync(st->fsm_end, D.fsm_end, B * 32, cudaMemcpyDeviceToHost, s));
    }
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
