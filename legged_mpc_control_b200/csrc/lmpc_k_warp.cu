// lmpc_k_warp.cu — size class 0 (n <= 63): persistent one-warp-per-QP solve and its launcher.
#include <string.h>

#include "lmpc_internal.h"
#include "lmpc_warp_solve.cuh"

namespace lmpc {

// Register cap of the one-warp kernel, measured on B200 (headline workload, solves/s):
//   __launch_bounds__(32, 9): ptxas settles on 168 registers, 600 B of spills, 9 warps/SM   3.75 M
//   __maxnreg__(224):         290 B of spills, 9 warps/SM                                   3.90 M
//   __maxnreg__(255):         230 B of spills, 8 warps/SM                                   4.19 M
// The kernel is latency-bound inside each warp, so fewer spills beat the ninth warp.
#ifndef LMPC_WARP_MAXNREG
#define LMPC_WARP_MAXNREG 255
#endif

// One warp per CTA: no block barrier exists in this kernel.  Problems are handed out by an atomic
// counter; the scaled Hessian of the problem in flight lives in this CTA's 32 KB slot of the
// L2-resident workspace `gws`.
template <int MAXNREG>
__global__ void __maxnreg__(MAXNREG) solve_warp_kernel_r(const __grid_constant__ KParams P,
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
  if (P.tma_prefetch) {
    // A/B experiment (north_star: "prediction matrices staged in shared memory via TMA"): the persistent
    // loop runs one work item AHEAD — while problem w solves, lane 0 has already claimed w_next and a
    // 1-D bulk copy (cp.async.bulk + mbarrier complete_tx) brings its x0 and foot_pos_abs records (96 B
    // each, 16-byte aligned; rot/cmd records are 72 B and contact 4H B, not bulk-copyable without
    // over-fetch) into a double-buffered staging area behind the plan.  Measured: see DESIGN.md §4.
    __shared__ __align__(16) double stage[2][24];
    __shared__ __align__(8) unsigned long long bar[2];
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bar[0]), bar1 = (unsigned)__cvta_generic_to_shared(&bar[1]);
    if (lane == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar1));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto claim = [&]() { int w = 0; if (lane == 0) w = atomicAdd(work_counter, 1); return __shfl_sync(0xffffffffu, w, 0); };
    auto prefetch = [&](int w, int slot) {  // lane 0 only
      const unsigned b = slot ? bar1 : bar0;
      const unsigned dst = (unsigned)__cvta_generic_to_shared(&stage[slot][0]);
      const size_t pi = (size_t)Bt.order[w];
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 192;" ::"r"(b) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 96, [%2];" ::"r"(dst),
                   "l"(Bt.x0 + 12 * pi), "r"(b) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 96, [%2];" ::"r"(dst + 96),
                   "l"(Bt.foot + 12 * pi), "r"(b) : "memory");
    };
    int w = claim(), slot = 0;
    unsigned phase[2] = {0u, 0u};
    if (w < total && lane == 0) prefetch(w, 0);
    while (w < total) {
      const int w_next = claim();
      if (w_next < total && lane == 0) prefetch(w_next, slot ^ 1);
      const unsigned b = slot ? bar1 : bar0;
      unsigned done = 0;
      while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(b), "r"(phase[slot]) : "memory");
      phase[slot] ^= 1u;
      solve_one_warp(P, Bt, ws, pl, smem, Hg, Bt.order[w], &stage[slot][0], &stage[slot][12]);
      w = w_next;
      slot ^= 1;
    }
    flush_warp_stats(ws, stats);
    return;
  }
  for (;;) {
    int w = 0;
    if (lane == 0) w = atomicAdd(work_counter, 1);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= total) break;
    solve_one_warp(P, Bt, ws, pl, smem, Hg, Bt.order[w]);
  }
  flush_warp_stats(ws, stats);
}

// Register cap -> resident warps per SM (shared memory allows 10 since the exchange buffers went): 255 -> 8,
// 224 -> 9, 200 -> 10.  Measured this round on the headline workload: 4.12 / 4.13 / 4.14 M solves/s — a third
// warp per scheduler adds nothing (the second one already saturates instruction fetch), so 255.
template <int MAXNREG>
static lmpc_status launch_warp_r(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work) {
  auto solve_warp_kernel = solve_warp_kernel_r<MAXNREG>;
  const WarpPlan wp = make_warp_plan(c.kp->H);
  static int occ_dev[kMaxDevices], bytes_dev[kMaxDevices];  // per device: the attribute lives in the device's context
  int& occ_cached = occ_dev[c.device];
  int& occ_bytes = bytes_dev[c.device];
  if (occ_bytes != wp.total_bytes) {
    CK(cudaFuncSetAttribute(solve_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wp.total_bytes));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_cached, solve_warp_kernel, 32, wp.total_bytes));
    occ_bytes = wp.total_bytes;
  }
  if (occ_cached < 1) { set_err("warp solve kernel does not fit on an SM"); return LMPC_ERR_CUDA; }
  int occ = occ_cached;
  if (c.kp->warp_occ_cap > 0 && occ > c.kp->warp_occ_cap) occ = c.kp->warp_occ_cap;  // experiment: fewer resident warps
  int grid = c.num_sms * occ;
  if (grid > kb.B) grid = kb.B;
  CK(ws->ensure((size_t)grid * WN * WN * sizeof(double)));
  if (c.kp->l2_persist) {  // experiment: pin the Hessian workspace in L2 (persisting access-policy window on the launch stream)
    cudaStreamAttrValue av;
    memset(&av, 0, sizeof(av));
    av.accessPolicyWindow.base_ptr = ws->p;
    av.accessPolicyWindow.num_bytes = (size_t)grid * WN * WN * sizeof(double);
    av.accessPolicyWindow.hitRatio = 1.0f;
    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    CK(cudaStreamSetAttribute(c.stream, cudaStreamAttributeAccessPolicyWindow, &av));
  }
  KBatch k = kb;
  k.order = order;
  solve_warp_kernel<<<grid, 32, wp.total_bytes, c.stream>>>(*c.kp, k, c.stats, (double*)ws->p, count, work);
  ++*c.launches;
  CK(cudaGetLastError());
  return LMPC_OK;
}

lmpc_status launch_warp_class(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work) {
  return launch_warp_r<LMPC_WARP_MAXNREG>(c, kb, ws, order, count, work);
}

}  // namespace lmpc
