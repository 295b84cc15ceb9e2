// lmpc_k_tile.cu — size classes 1-4 (64 <= n <= 384): persistent tile solve, one QP per CTA of NW
// warps, and its launcher.  Compiled once per tile count (-DLMPC_TILE_NT=15|24|36|48) so the four
// classes build in parallel.
#include "lmpc_internal.h"
#include "lmpc_tile_solve.cuh"

#ifndef LMPC_TILE_NT
#error "compile with -DLMPC_TILE_NT=15|24|36|48"
#endif

namespace lmpc {

// Fragment-ordered factor in shared memory (61 KB at NT = 15, 154 KB at NT = 24) or, for NT = 36 / 48,
// in the CTA's global workspace slot; Hessian slot in the L2 workspace.  Warp 0 runs the solve and
// posts jobs to a mailbox; warps 1 .. NW-1 share the tile columns of every Cholesky factorization,
// the Hessian block build, the reduced-matrix build and H*x (tile_helper_loop).
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

template <int NT, int NW>
static lmpc_status launch_tile_nw(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work) {
  const TilePlan tp = make_tile_plan<NT>(c.kp->H);
  auto kern = solve_tile_kernel<NT, NW>;
  static int occ_dev[kMaxDevices], bytes_dev[kMaxDevices];  // per device: the attribute lives in the device's context
  int& occ_cached = occ_dev[c.device];
  int& occ_bytes = bytes_dev[c.device];
  if (occ_bytes != tp.total_bytes) {
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tp.total_bytes));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_cached, kern, 32 * NW, tp.total_bytes));
    occ_bytes = tp.total_bytes;
  }
  if (occ_cached < 1) { set_err("tile solve kernel does not fit on an SM"); return LMPC_ERR_CUDA; }
  int grid = c.num_sms * occ_cached;
  if (grid > kb.B) grid = kb.B;
  CK(ws->ensure((size_t)grid * tile_slot_doubles<NT>() * sizeof(double)));
  KBatch k = kb;
  k.order = order;
  kern<<<grid, 32 * NW, tp.total_bytes, c.stream>>>(*c.kp, k, c.stats, (double*)ws->p, count, work);
  ++*c.launches;
  CK(cudaGetLastError());
  return LMPC_OK;
}

// Warps per QP, measured (C3 / reference preset, solves/s): NT=15: 598k / 647k / 519k at 1 / 2 / 4 warps
// (4 warps cost the third CTA per SM); NT=24: 90k / 106k / 141k / 150k at 1 / 2 / 4 / 8 warps (one CTA per SM anyway);
// NT=36/48: 4.
#if LMPC_TILE_NT == 8
// experiment: the rolled tile code at the one-warp size, one warp per QP — an A/B against
// solve_warp_kernel's straight-line code for the instruction-cache question
lmpc_status launch_tile8(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int) {
  return launch_tile_nw<8, 1>(c, kb, ws, order, count, work);
}
#elif LMPC_TILE_NT == 15
lmpc_status launch_tile15(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw) {
  if (nw == 0) nw = kb.B <= c.num_sms ? 4 : 2;  // fewer problems than SMs: latency counts, occupancy does not
  if (nw == 1) return launch_tile_nw<15, 1>(c, kb, ws, order, count, work);
  if (nw == 2) return launch_tile_nw<15, 2>(c, kb, ws, order, count, work);
  return launch_tile_nw<15, 4>(c, kb, ws, order, count, work);
}
#elif LMPC_TILE_NT == 24
lmpc_status launch_tile24(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw) {
  if (nw == 0) nw = 8;  // one CTA per SM either way (154 KB factor): 8 warps give every scheduler a second warp, +6 % over 4
  if (nw == 1) return launch_tile_nw<24, 1>(c, kb, ws, order, count, work);
  if (nw == 2) return launch_tile_nw<24, 2>(c, kb, ws, order, count, work);
  if (nw == 8) return launch_tile_nw<24, 8>(c, kb, ws, order, count, work);
  return launch_tile_nw<24, 4>(c, kb, ws, order, count, work);
}
#elif LMPC_TILE_NT == 36
lmpc_status launch_tile36(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int) {
  return launch_tile_nw<36, 4>(c, kb, ws, order, count, work);
}
#elif LMPC_TILE_NT == 48
lmpc_status launch_tile48(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int) {
  return launch_tile_nw<48, 4>(c, kb, ws, order, count, work);
}
#else
#error "unsupported LMPC_TILE_NT"
#endif

}  // namespace lmpc
