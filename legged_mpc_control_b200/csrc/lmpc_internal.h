// lmpc_internal.h — host-side plumbing shared by the translation units of liblmpc.so: error
// reporting, device buffers, size classes and the per-class launchers (each kernel family is
// compiled in its own .cu so the library builds in parallel).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/lmpc.h"
#include "lmpc_types.cuh"

namespace lmpc {

void set_err(const std::string& s);

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      ::lmpc::set_err(std::string(#call) + ": " + cudaGetErrorString(e_));                    \
      return LMPC_ERR_CUDA;                                                                   \
    }                                                                                         \
  } while (0)

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

// Size classes by reduced dimension n (3 per stance leg-step):
//   0: n <= 63            one warp per QP, 8 x 8 tiles        (solve_warp_kernel / solve_warp_f32_kernel)
//   1: 64 <= n <= 120     15 x 15 tiles                       (solve_tile_kernel<15>)
//   2: 121 <= n <= 192    24 x 24 tiles                       (solve_tile_kernel<24>)
//   4: 193 <= n <= 288    36 x 36 tiles, factor in the global slot (solve_tile_kernel<36>)
//   3: n > 288            48 x 48 tiles, factor in the global slot (solve_tile_kernel<48>)
constexpr int kNumClasses = 5;
constexpr int kCtrCount = 8;   // counter block: work counters at [c], class sizes at [kCtrCount + c]; retry launches: work counters at [2 kCtrCount + c], list lengths at [3 kCtrCount + c]
constexpr int kWarpNMax = 63;  // 3 * WNBMAX
constexpr int kTileA = 15, kTileB = 24, kTileC = 48;  // 8 x 48 = 384 = 12 x LMPC_MAX_HORIZON
constexpr int kTileD = 36;     // class 4: 192 < n <= 288 (crawl at H = 30, stand at H = 20)
__host__ __device__ inline int size_class(int n) {
  return n <= kWarpNMax ? 0 : (n <= 8 * kTileA ? 1 : (n <= 8 * kTileB ? 2 : (n <= 8 * kTileD ? 4 : 3)));
}

// What a per-class launcher needs from the handle.
constexpr int kMaxDevices = 64;
struct LaunchCtx {
  int device;         // CUDA ordinal < kMaxDevices
  const KParams* kp;
  KStats* stats;      // device
  int num_sms;
  cudaStream_t stream;
  int64_t* launches;  // host counter (bench evidence)
};

// Each launcher enqueues ONE persistent kernel that solves the problems listed in order[0 .. *count)
// (count and the work counter live on the device; no host synchronisation).  `ws` is the launcher's
// own L2-resident workspace, grown on demand.
lmpc_status launch_warp_class(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work);
// nw: warps per QP (1, 2 or 4; 0 = the measured default of the class)
lmpc_status launch_tile8(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw);
lmpc_status launch_tile15(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw);
lmpc_status launch_tile24(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw);
lmpc_status launch_tile36(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw);
lmpc_status launch_tile48(const LaunchCtx& c, const KBatch& kb, DevBuf* ws, const int* order, const int* count, int* work, int nw);

}  // namespace lmpc

// ---- the engine object behind the C ABI (one per GPU), shared by lmpc_api.cu and lmpc_aux.cu
// Experiment knobs, read ONCE per handle at lmpc_create (never on the solve path).
struct Tuning {
  int tile_nw = 0;        // LMPC_TILE_NW: warps per QP of tile classes 1/2 (0 = measured default)
  bool debug_stats = false;  // LMPC_DEBUG_STATS: print per-solve counters to stderr
  bool fork_classes = true;  // LMPC_FORK_CLASSES=0: all size classes on the call's stream, one after the other
  int host_lanes = 4;     // LMPC_HOST_LANES: slices of a large host-pointer batch in flight (copy/solve overlap); <= 1 = off
};

struct lmpc_handle {
  lmpc_params params;
  lmpc::KParams kp;
  int device = 0;
  int max_batch = 0;
  int num_sms = 0;
  int smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // device staging for host-pointer calls
  lmpc::DevBuf d_x0, d_rot, d_foot, d_cmd, d_contact, d_q, d_r, d_mu, d_fz, d_mass, d_inertia;
  lmpc::DevBuf d_grf, d_status, d_iters, d_active, d_U;
  lmpc::DevBuf d_stats, d_counter, d_ws, d_wsA, d_wsT1, d_wsT2, d_wsT3, d_order, d_warm;
  // Size classes 1.. launch on side streams forked from the call's stream after the classification and joined
  // before its end event: their lists are independent, so an empty (or short) class runs beside the one that has
  // the work instead of adding its launch latency behind it.
  cudaStream_t side[lmpc::kNumClasses] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_fork = nullptr, ev_join[lmpc::kNumClasses] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  lmpc::DevBuf d_retry;     // per size class: the problems the first launch handed to the retry launch (KBatch::retry_list)
  lmpc::KParams kp_retry;   // kp with the conservative interior-point start (z0_scale = 1)
  bool warm_enabled = false;
  int warm_shift = 1;
  // small-batch host path: one packed H2D and one packed D2H (a single robot's tick is latency-bound)
  lmpc::DevBuf d_pin, d_pout;
  unsigned char* h_pin = nullptr;
  unsigned char* h_pout = nullptr;
  // closed-loop rollout state (lmpc_rollout)
  lmpc::DevBuf d_rx, d_rfw, d_rph, d_rgait, d_rspeed, d_rtot, d_xlog, d_glog, d_rfsm;
  cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;
  lmpc::DevBuf d_aux;  // staging of the host-pointer FSM / leg-kinematics / torque-map calls (lmpc_aux.cu)
  Tuning tune;
  lmpc::KStats last_stats;
  float last_ms = 0.f;
  int64_t launches = 0;
  bool pending = false;
  cudaStream_t pending_stream = nullptr;
  // Large host-pointer batches are cut into `tune.host_lanes` slices, each with its own stream, staging
  // buffers and workspace on THIS device (an lmpc_sharded whose shards all name h->device), so that the
  // H2D copy of slice k+1 and the D2H copy of slice k-1 run under the solve of slice k, and the persistent
  // kernel of slice k+1 fills the SMs as the last problems of slice k drain.  Created on first use.
  struct lmpc_sharded* lanes = nullptr;
  bool is_lane = false;  // a slice engine never slices again
};

