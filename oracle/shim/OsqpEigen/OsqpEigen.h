/*
 * OsqpEigen/OsqpEigen.h — CAPTURING STUB of the osqp-eigen interface the reference calls
 * (ConvexQPSolver.cpp:182-194 setup, :315-320 update/solve/getSolution).  TEST INFRASTRUCTURE
 * ONLY.  It does not solve anything itself: it records the (P, q, A, l, u) the unmodified
 * reference code hands to OSQP, and `solve()` asks a hook (installed by oracle/ref_harness.cpp)
 * for the primal solution, so that the rest of the reference tick can run on an exact answer.
 * OSQP itself is neither vendored nor version-pinned by the reference (.devcontainer/Dockerfile:55,69).
 */
#ifndef LMPC_SHIM_OSQPEIGEN_H_
#define LMPC_SHIM_OSQPEIGEN_H_
#include <memory>
#include <vector>
#include "../mini_eigen.h"
#include "../mini_eigen_sparse.h"

namespace OsqpEigen {

const double INFTY = 1e30; /* OSQP_INFTY of osqp >= 0.6 */

struct CapturedQP {
  int n = 0, m = 0;
  Eigen::SparseMatrix<double> P, A;
  Eigen::VectorXd q, l, u;
  bool verbose = true, warm_start = false;
  double eps_abs = 1e-3, eps_rel = 1e-3; /* OSQP defaults */
  int solves = 0, matrix_updates = 0;
};

/* hook(qp, z): fill z (size n) with the primal solution; return false to leave NaNs (the
 * reference then returns zeros, ConvexQPSolver.cpp:321-326). */
typedef bool (*SolveHook)(const CapturedQP& qp, Eigen::VectorXd& z, void* user);
inline SolveHook& solve_hook() { static SolveHook h = nullptr; return h; }
inline void*& solve_hook_user() { static void* u = nullptr; return u; }
/* the most recently solved (or initialised) problem, for the harness to read back */
inline const CapturedQP*& last_qp() { static const CapturedQP* p = nullptr; return p; }

class Settings {
 public:
  explicit Settings(CapturedQP* qp) : qp_(qp) {}
  void setVerbosity(bool v) { qp_->verbose = v; }
  void setAbsoluteTolerance(double e) { qp_->eps_abs = e; }
  void setRelativeTolerance(double e) { qp_->eps_rel = e; }
  void setWarmStart(bool w) { qp_->warm_start = w; }
  void setMaxIteration(int) {}
  void setPolish(bool) {}
 private:
  CapturedQP* qp_;
};

class Data {
 public:
  explicit Data(CapturedQP* qp) : qp_(qp) {}
  void setNumberOfVariables(int n) { qp_->n = n; }
  void setNumberOfConstraints(int m) { qp_->m = m; }
  template <class S> bool setHessianMatrix(const S& P) { qp_->P = P; return true; }
  template <class S> bool setLinearConstraintsMatrix(const S& A) { qp_->A = A; return true; }
  bool setGradient(Eigen::VectorXd& q) { qp_->q = q; return true; }
  bool setLowerBound(Eigen::VectorXd& l) { qp_->l = l; return true; }
  bool setUpperBound(Eigen::VectorXd& u) { qp_->u = u; return true; }
 private:
  CapturedQP* qp_;
};

class Solver {
 public:
  Solver() : qp_(new CapturedQP()), settings_(new Settings(qp_.get())), data_(new Data(qp_.get())) {}
  Solver(Solver&&) = default;
  Solver& operator=(Solver&&) = default;
  const std::unique_ptr<Settings>& settings() const { return settings_; }
  const std::unique_ptr<Data>& data() const { return data_; }
  bool initSolver() { inited_ = true; sol_.resize(qp_->n); sol_.setZero(); last_qp() = qp_.get(); return true; }
  bool isInitialized() const { return inited_; }
  template <class S> bool updateLinearConstraintsMatrix(const S& A) { qp_->A = A; ++qp_->matrix_updates; return true; }
  template <class S> bool updateHessianMatrix(const S& P) { qp_->P = P; return true; }
  bool updateBounds(const Eigen::VectorXd& l, const Eigen::VectorXd& u) { qp_->l = l; qp_->u = u; return true; }
  bool updateGradient(const Eigen::VectorXd& q) { qp_->q = q; return true; }
  bool solve() {
    ++qp_->solves; last_qp() = qp_.get();
    sol_.resize(qp_->n);
    bool ok = solve_hook() ? solve_hook()(*qp_, sol_, solve_hook_user()) : false;
    if (!ok) sol_.setConstant(std::numeric_limits<double>::quiet_NaN());
    return ok;
  }
  const Eigen::VectorXd& getSolution() { return sol_; }
  const CapturedQP& captured() const { return *qp_; }
 private:
  std::unique_ptr<CapturedQP> qp_;
  std::unique_ptr<Settings> settings_;
  std::unique_ptr<Data> data_;
  Eigen::VectorXd sol_;
  bool inited_ = false;
};

}  // namespace OsqpEigen
#endif
