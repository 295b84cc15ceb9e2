// lmpc_noalloc.cu — lmpc_solve_noalloc: the one-warp-per-QP solve (size class 0, n <= 63) for callers
// that cannot tolerate an allocation on the call path (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The existing path (lmpc_k_warp.cu) takes its per-warp Hessian slots, work counter and statistics
// block from cudaMalloc'ed buffers that grow on demand.  Here all of that lives in ONE file-scope
// device array of fixed size:
//
//   g_noalloc_scratch = [ header, 256 B | kNoAllocSlots Hessian slots of 64 x 64 doubles (32 KB each) ]
//                     = 256 + 148 * 32 768 = 4 849 920 B  (4.63 MiB, held by every GPU that loads
//                       liblmpc.so until the process exits)
//
// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, NOT a device address: the
// device address is taken once per device with cudaGetSymbolAddress (noalloc_setup below) and that
// pointer is what the kernel receives.
//
// Every call shares the array and it keeps the last call's contents.  The header (work counter,
// too-large counter, statistics) must start at zero and is cleared by a cudaMemsetAsync in front of
// every launch; a Hessian slot is fully rewritten by solve_one_warp for every problem before it is
// read, so the slots need no clearing.  Two calls that overlap on the GPU (different streams, or two
// host threads on the same device) race on the header and the slots: callers serialise the calls of
// one device.
#include <mutex>
#include <string>

#include "lmpc_internal.h"
#include "lmpc_warp_solve.cuh"

namespace lmpc {

constexpr int kNoAllocSlots = 148;                       // one resident warp per B200 SM
constexpr size_t kNoAllocHeader = 256;                   // bytes
constexpr size_t kNoAllocSlotBytes = (size_t)WN * WN * sizeof(double);
constexpr size_t kNoAllocBytes = kNoAllocHeader + kNoAllocSlots * kNoAllocSlotBytes;

struct NoAllocHeader {   // first 256 bytes of the scratch; zeroed before every launch
  int work;              // atomic work counter of the persistent loop
  int too_large;         // problems refused because n > 63
  KStats stats;
};
static_assert(sizeof(NoAllocHeader) <= kNoAllocHeader, "header does not fit its 256 bytes");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// Same body as solve_warp_kernel_r (lmpc_k_warp.cu) with three differences: the workspace is the
// static scratch, problems are taken in batch order (no classify pass, no order list), and a problem
// whose reduced dimension exceeds the one-warp limit is refused with LMPC_PROB_TOO_LARGE and zero
// outputs instead of being routed to a tile kernel.
__global__ void __maxnreg__(255) solve_warp_static_kernel(const __grid_constant__ KParams P,
                                                          const __grid_constant__ KBatch Bt,
                                                          unsigned char* scratch) {
  extern __shared__ __align__(16) unsigned char smem[];
  const WarpPlan pl = make_warp_plan(P.H);
  NoAllocHeader* hd = reinterpret_cast<NoAllocHeader*>(scratch);
  double* Hg = reinterpret_cast<double*>(scratch + kNoAllocHeader) + (size_t)blockIdx.x * (WN * WN);
  const int lane = threadIdx.x;
  const int H = P.H;
  WarpStats ws;
  for (;;) {
    int w = 0;
    if (lane == 0) w = atomicAdd(&hd->work, 1);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= Bt.B) break;
    const uint8_t* contact = Bt.contact + (size_t)(4 * H) * w;
    int nb = 0;  // stance leg-steps = reduced dimension / 3
    for (int base = 0; base < 4 * H; base += 32) {
      const int idx = base + lane;
      nb += __popc(__ballot_sync(0xffffffffu, idx < 4 * H && contact[idx] != 0));
    }
    if (nb > WNBMAX) {
      const size_t p = (size_t)w;
      if (lane < 12) Bt.grf[12 * p + lane] = 0.0;
      if (Bt.U) for (int k = lane; k < 12 * H; k += 32) Bt.U[(size_t)(12 * H) * p + k] = 0.0;
      const int words = (24 * H + 31) / 32;
      if (Bt.active) for (int k = lane; k < words; k += 32) Bt.active[(size_t)words * p + k] = 0u;
      if (lane == 0) {
        if (Bt.status) Bt.status[p] = LMPC_PROB_TOO_LARGE;
        if (Bt.iters) Bt.iters[p] = 0;
        atomicAdd(&hd->too_large, 1);
      }
      continue;
    }
    solve_one_warp(P, Bt, ws, pl, smem, Hg, w);
  }
  flush_warp_stats(ws, &hd->stats);
}

// Once per device, with that device current: the device address of the scratch symbol and the kernel's
// dynamic shared-memory attribute.  Neither allocates; the array itself is part of the loaded module.
struct NoAllocDev { unsigned char* scratch = nullptr; int smem_bytes = 0; };
static NoAllocDev g_dev[kMaxDevices];
static std::mutex g_dev_mu;

static lmpc_status noalloc_setup(int device, int smem_bytes, unsigned char** scratch) {
  std::lock_guard<std::mutex> lk(g_dev_mu);
  NoAllocDev& d = g_dev[device];
  if (!d.scratch) {
    unsigned char* ptr = nullptr;
    CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (!ptr) { set_err("cudaGetSymbolAddress returned a null scratch address"); return LMPC_ERR_CUDA; }
    d.scratch = ptr;
  }
  if (d.smem_bytes < smem_bytes) {
    CK(cudaFuncSetAttribute(solve_warp_static_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    d.smem_bytes = smem_bytes;
  }
  *scratch = d.scratch;
  return LMPC_OK;
}

}  // namespace lmpc

using lmpc::set_err;

extern "C" lmpc_status lmpc_solve_noalloc(lmpc_handle* h, int32_t batch, const lmpc_batch_in* in,
                                          const lmpc_batch_out* out, void* cuda_stream, int32_t blocking) {
  if (!h || !in || !out) { set_err("NULL handle/in/out"); return LMPC_ERR_INVALID_ARG; }
  if (batch < 1) { set_err("batch must be >= 1"); return LMPC_ERR_INVALID_ARG; }
  if (batch > LMPC_NOALLOC_MAX_BATCH) {
    set_err("lmpc_solve_noalloc: batch " + std::to_string(batch) + " exceeds LMPC_NOALLOC_MAX_BATCH (" +
            std::to_string(LMPC_NOALLOC_MAX_BATCH) + "); use lmpc_solve / lmpc_solve_async");
    return LMPC_ERR_INVALID_ARG;
  }
  if (!in->on_device || !out->on_device) {
    set_err("lmpc_solve_noalloc needs device pointers (staging host arrays would allocate)");
    return LMPC_ERR_INVALID_ARG;
  }
  if (!in->x0 || !in->rot || !in->foot_pos_abs || !in->cmd || !in->contact) { set_err("x0, rot, foot_pos_abs, cmd, contact are required"); return LMPC_ERR_INVALID_ARG; }
  if (!out->grf) { set_err("out->grf is required"); return LMPC_ERR_INVALID_ARG; }
  if (h->device < 0 || h->device >= lmpc::kMaxDevices) { set_err("device ordinal out of range"); return LMPC_ERR_INVALID_ARG; }
  CK(cudaSetDevice(h->device));
  const lmpc::WarpPlan wp = lmpc::make_warp_plan(h->kp.H);
  unsigned char* scratch = nullptr;
  lmpc_status st = lmpc::noalloc_setup(h->device, wp.total_bytes, &scratch);
  if (st != LMPC_OK) return st;

  lmpc::KBatch kb;
  kb.x0 = in->x0; kb.rot = in->rot; kb.foot = in->foot_pos_abs; kb.cmd = in->cmd; kb.contact = in->contact;
  kb.q = in->q; kb.r = in->r; kb.mu = in->mu; kb.fzmax = in->fz_max; kb.mass = in->mass; kb.inertia = in->inertia;
  kb.grf = out->grf; kb.status = out->status; kb.iters = out->iters; kb.active = out->active_set; kb.U = out->U;
  kb.order = nullptr;
  kb.B = batch;
  kb.warm = nullptr; kb.warm_shift = 0;  // no warm start: its record store is a cudaMalloc'ed buffer of the handle
  kb.retry_list = nullptr; kb.retry_count = nullptr;  // one launch, no retry list (it would be a cudaMalloc'ed buffer): ...

  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  CK(cudaMemsetAsync(scratch, 0, lmpc::kNoAllocHeader, s));  // work counter, too-large counter, statistics
  const int grid = batch < lmpc::kNoAllocSlots ? batch : lmpc::kNoAllocSlots;
  lmpc::solve_warp_static_kernel<<<grid, 32, wp.total_bytes, s>>>(h->kp_retry, kb, scratch);  // ... so the conservative interior-point start
  CK(cudaGetLastError());
  h->launches++;
  if (!blocking) return LMPC_OK;

  lmpc::NoAllocHeader hd;
  CK(cudaMemcpyAsync(&hd, scratch, sizeof(hd), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  if (hd.too_large > 0 || hd.stats.not_ok > 0) {
    set_err(std::to_string(hd.too_large) + " problem(s) above the one-warp limit n <= 63 (LMPC_PROB_TOO_LARGE), " +
            std::to_string((unsigned long long)hd.stats.not_ok) + " did not converge (grf zeroed)");
    return LMPC_ERR_NOT_CONVERGED;
  }
  return LMPC_OK;
}
