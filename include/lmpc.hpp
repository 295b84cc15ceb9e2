// lmpc.hpp — header-only C++17 face of the C ABI in lmpc.h (liblmpc.so).
//
// Mirrors the object the reference's controller owns for this path, legged::ConvexQPSolver
// (/root/reference/src/legged_ctrl/include/mpc_ctrl/convex_mpc/ConvexQPSolver.h:25-38): constructed
// once from the weights, then per tick calc_mpc_reference + update_cons_matrix + compute_grfs
// (src/mpc_ctrl/convex_mpc/ConvexMpc.cpp:70-72).  Inputs are "Eigen-compatible": anything with a
// .data() returning column-major doubles (Eigen::Vector3d, Matrix3d, Matrix<double,3,4>, ...) or a
// plain pointer; Eigen itself is not needed to compile this header.  Errors become exceptions on
// this side of the boundary only (the C ABI never throws).  There is no CPU fallback: construction
// throws lmpc::Error(LMPC_ERR_NO_DEVICE) when no sm_100 GPU is usable.
#ifndef LMPC_HPP_
#define LMPC_HPP_

#include <array>
#include <cstdint>
#include <stdexcept>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "lmpc.h"

namespace lmpc {

class Error : public std::runtime_error {
 public:
  Error(lmpc_status code, const std::string& what) : std::runtime_error(what), code_(code) {}
  lmpc_status code() const noexcept { return code_; }

 private:
  lmpc_status code_;
};

namespace detail {
inline const double* ptr(const double* p) { return p; }
template <class M>
auto ptr(const M& m) -> decltype(static_cast<const double*>(m.data())) { return m.data(); }
inline void check(lmpc_status st, bool allow_not_converged = false) {
  if (st == LMPC_OK || (allow_not_converged && st == LMPC_ERR_NOT_CONVERGED)) return;
  throw Error(st, std::string("lmpc: ") + lmpc_last_error_string());
}
}  // namespace detail

inline lmpc_params preset(const char* name) {
  lmpc_params p;
  if (lmpc_params_preset(name, &p) != LMPC_OK) throw Error(LMPC_ERR_INVALID_ARG, std::string("lmpc: unknown preset ") + name);
  return p;
}

// One robot's inputs for one tick, as the reference reads them from LeggedState
// (ConvexQPSolver.cpp:254-346; LeggedState.h:27-33,52,79-96).
struct RobotTick {
  std::array<double, 12> x0{};        // root_euler, root_pos, root_ang_vel, root_lin_vel
  std::array<double, 9> rot{};        // root_rot_mat, column-major
  std::array<double, 12> foot_pos_abs{};  // 3x4 column-major
  std::array<double, 9> cmd{};        // roll_d, pitch_d, z_d, v_d_rel(3), w_d_rel(3)
  std::array<std::uint8_t, 4 * LMPC_MAX_HORIZON> contact{};  // [H][4]
};

// Batched replacement of legged::ConvexQPSolver: one engine per GPU, single caller (like the
// reference's solver object, which only thread 1 touches: src/main.cpp:130-163).
class ConvexMpcEngine {
 public:
  // ConvexQPSolver(q_weights, r_weights) (ConvexQPSolver.cpp:16): start from a preset (the
  // reference's compile-time constants) and overwrite the weights; q and r are 12 doubles each.
  template <class Q, class R>
  ConvexMpcEngine(const Q& q_weights, const R& r_weights, const char* preset_name = "reference_a1", int max_batch = 1,
                  int device = 0)
      : ConvexMpcEngine(with_weights(preset(preset_name), detail::ptr(q_weights), detail::ptr(r_weights)), max_batch, device) {}

  explicit ConvexMpcEngine(const lmpc_params& p, int max_batch = 1, int device = 0) : params_(p), max_batch_(max_batch) {
    detail::check(lmpc_create(&params_, max_batch, device, &h_));
  }
  ~ConvexMpcEngine() { lmpc_destroy(h_); }
  ConvexMpcEngine(const ConvexMpcEngine&) = delete;
  ConvexMpcEngine& operator=(const ConvexMpcEngine&) = delete;
  ConvexMpcEngine(ConvexMpcEngine&& o) noexcept : params_(o.params_), max_batch_(o.max_batch_), h_(std::exchange(o.h_, nullptr)) {}

  const lmpc_params& params() const noexcept { return params_; }
  int horizon() const noexcept { return params_.horizon; }
  lmpc_handle* handle() noexcept { return h_; }

  // calc_mpc_reference + update_cons_matrix + compute_grfs for ONE robot (ConvexMpc.cpp:70-72).
  // Arguments are the LeggedState fields the reference reads; `contact` is [H][4] (row 0 =
  // plan_contacts, row i = leg_FSM[j].predict_contact_state(i*dt), ConvexQPSolver.cpp:329-346).
  // Returns u_0 as 12 doubles, leg-major xyz == Eigen::Matrix<double,3,4> column-major
  // (ConvexMpc.cpp:73-75).  A failed solve returns zeros, like ConvexQPSolver.cpp:321-326.
  template <class V12, class M3, class M34, class V9>
  std::array<double, 12> compute_grfs(const V12& x0, const M3& root_rot_mat, const M34& foot_pos_abs, const V9& cmd,
                                      const std::uint8_t* contact, int* status = nullptr) {
    lmpc_batch_in in{};
    lmpc_batch_out out{};
    in.x0 = detail::ptr(x0); in.rot = detail::ptr(root_rot_mat); in.foot_pos_abs = detail::ptr(foot_pos_abs);
    in.cmd = detail::ptr(cmd); in.contact = contact;
    std::array<double, 12> grf{};
    std::int32_t st = -1;
    out.grf = grf.data(); out.status = &st;
    detail::check(lmpc_solve(h_, 1, &in, &out), /*allow_not_converged=*/true);
    if (status) *status = st;
    return grf;
  }
  std::array<double, 12> compute_grfs(const RobotTick& t, int* status = nullptr) {
    return compute_grfs(t.x0, t.rot, t.foot_pos_abs, t.cmd, t.contact.data(), status);
  }

  // Whole batch, SoA (see lmpc_batch_in).  Returns the number of problems that did not converge
  // (their grf is zero, ConvexQPSolver.cpp:321-326).
  std::int64_t solve(int batch, const lmpc_batch_in& in, const lmpc_batch_out& out) {
    detail::check(lmpc_solve(h_, batch, &in, &out), true);
    return lmpc_last_not_converged(h_);
  }
  void solve_async(int batch, const lmpc_batch_in& in, const lmpc_batch_out& out, void* cuda_stream = nullptr) {
    detail::check(lmpc_solve_async(h_, batch, &in, &out, cuda_stream));
  }
  void wait() { detail::check(lmpc_wait(h_), true); }

  // OSQP warm-start analogue (ConvexQPSolver.cpp:185)
  void set_warm_start(bool enable, int horizon_shift = 1) { detail::check(lmpc_set_warm_start(h_, enable ? 1 : 0, horizon_shift)); }
  void reset_warm_start() { detail::check(lmpc_reset_warm_start(h_)); }
  std::int64_t last_warm_hits() const { return lmpc_last_warm_hits(h_); }

  // closed-loop batched rollout (BASELINE config 5)
  lmpc_rollout_result rollout(int batch, const lmpc_rollout_cfg& cfg, const lmpc_rollout_state& state,
                              lmpc_rollout_result result = lmpc_rollout_result{}) {
    detail::check(lmpc_rollout(h_, batch, &cfg, &state, &result), true);
    return result;
  }

  // contact[H][4] of one robot from its gait (LeggedContactFSM tables + predict_contact_state)
  std::vector<std::uint8_t> contact_table(int gait, const double phase[4], double gait_speed) const {
    std::vector<std::uint8_t> c(4 * (size_t)params_.horizon);
    const std::int32_t g = gait;
    detail::check(lmpc_gait_contact_table(1, params_.horizon, params_.dt, &g, phase, &gait_speed, c.data()));
    return c;
  }

  std::int64_t launch_count() const { return lmpc_launch_count(h_); }
  float last_kernel_ms() const { return lmpc_last_kernel_ms(h_); }

 private:
  static lmpc_params with_weights(lmpc_params p, const double* q, const double* r) {
    for (int i = 0; i < 12; ++i) { p.q[i] = q[i]; p.r[i] = r[i]; }
    return p;
  }
  lmpc_params params_;
  int max_batch_;
  lmpc_handle* h_ = nullptr;
};

// One process, several GPUs — thin wrapper of lmpc_create_sharded / lmpc_solve_sharded:
// contiguous index ranges [g B/G, (g+1) B/G), one engine + one host thread per device, no collective.
// Host pointers only; warm-start records stay with the device that owns the slice, so keep the batch
// order stable between ticks.  All argument checks run in the C layer before any thread starts.
class ShardedEngine {
 public:
  ShardedEngine(const lmpc_params& p, int max_batch, const std::vector<int>& devices) : params_(p) {
    std::vector<std::int32_t> d(devices.begin(), devices.end());
    detail::check(lmpc_create_sharded(&params_, max_batch, d.data(), (std::int32_t)d.size(), &s_));
  }
  ~ShardedEngine() { lmpc_destroy_sharded(s_); }
  ShardedEngine(const ShardedEngine&) = delete;
  ShardedEngine& operator=(const ShardedEngine&) = delete;
  int shards() const noexcept { return lmpc_sharded_num_devices(s_); }
  lmpc_handle* shard(int g) { return lmpc_sharded_handle(s_, g); }
  // first problem of shard g for a batch of B problems
  static int shard_begin(int B, int G, int g) { return (int)(((long long)B * g) / G); }

  // Returns the number of problems that did not converge (their grf is zero); throws on any hard error.
  std::int64_t solve(int batch, const lmpc_batch_in& in, const lmpc_batch_out& out) {
    detail::check(lmpc_solve_sharded(s_, batch, &in, &out), /*allow_not_converged=*/true);
    return lmpc_sharded_last_not_converged(s_);
  }
  void set_warm_start(bool enable, int horizon_shift = 1) {
    detail::check(lmpc_set_warm_start_sharded(s_, enable ? 1 : 0, horizon_shift));
  }

 private:
  lmpc_params params_;
  lmpc_sharded* s_ = nullptr;
};

}  // namespace lmpc
#endif  // LMPC_HPP_
