// CG.hpp — the CGSolver::CG solver template on a B200.
//
// Drop-in for the reference header of the same name (reference src/CG.hpp:53-601): same class
// template, same public methods (createCG, setMatrix, setTarget, setInital / setInitial, solve,
// accuracy, extract, extractTo, getDimension, memoryFootprint, calculateExpectedStepCount), same
// exceptions.  The body of solve() is one call into the C-ABI (cgb_cg_solve, cgb200.h): the whole
// iteration runs as fused sm_100a kernels under a CUDA graph with the stop flag on the device,
// where the reference submits 12 device operations and one host wait per iteration (:359-436).
//
// Stop rule, kept exactly (reference :396-404, :436): ABSOLUTE tolerance `improvement` on the norm
// of the recurrence residual, tested against r.r from the start of the iteration (one-iteration
// lag), NaN stops, at most N + 1 loop trips.  solve() with the default improvement = 0 therefore
// iterates until r.r underflows, as the reference does.
//
// Additive API (the reference returns none of it): iterations(), residualNorm(), seconds(),
// stats(), options().  The HiFlow3 overloads (#ifdef HIFLOW in the reference) are not provided:
// that dependency is absent from the reference tree and the macro is disabled there.
#ifndef CGB200_CG_HPP
#define CGB200_CG_HPP

#include <AdaptiveCpp/sycl/sycl.hpp>

#include <cmath>
#include <cstddef>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "LinearAlgebraTypes.hpp"
#include "VectorOperations.hpp"

namespace CGSolver {

namespace asycl = acpp::sycl;

template <typename DT, Debuglevel debug = Debuglevel::None> class CG {
  static_assert(std::is_same<DT, double>::value, "the cgb200 kernels are fp64 (the reference instantiates double only)");

public:
  using Scalar = CGSolver::Scalar<DT>;
  using Matrix = CGSolver::Matrix<DT>;
  using Vector = CGSolver::Vector<DT>;

  // No device memory is touched until a set* call.
  CG(asycl::queue &queue) : _queue(queue), A(queue), x(_queue), b(_queue) {
    cgb_cg_default_opts(&_opts);
    _stats = cgb_cg_stats();
    log("Constructing CG Object");
  }

  // A solver on its own queue (device 0, or $CGB_DEVICE).
  static std::unique_ptr<CG> createCG() {
    asycl::queue q;
    return std::unique_ptr<CG>(new CG(q));
  }

  void setMatrix(std::vector<DT> &data, std::vector<int> &columns, std::vector<int> &rows) {
    log("Setting Matrix");
    A.init(data, columns, rows);
  }
  void setMatrix(Matrix &&M) { A = std::forward<Matrix>(M); }

  int getDimension() const { return static_cast<int>(this->A.N()); }

  void setTarget(std::vector<DT> &_data) {
    b.init(_data);
    executeQueue();
  }
  void setTarget(Vector &&V) { b = std::forward<Vector>(V); }

  // (sic) the reference spells the std::vector overload "setInital" (:215)
  void setInital(std::vector<DT> &_data) {
    x.init(_data);
    executeQueue();
  }
  void setInitial(Vector &&V) { x = std::forward<Vector>(V); }

  void calculateExpectedStepCount(DT /*accuracy*/) {} // empty in the reference as well (:235)

  // Solve A x = b by conjugate gradients, starting from x (zeros unless set).
  void solve(DT improvement = static_cast<DT>(0)) {
    log("Solving System");
    if (this->b.data() == nullptr) throw std::runtime_error("No right hand side to solve for");
    if (this->A.columns().get() == nullptr) throw std::runtime_error("No Matrix given");
    const auto N = A.N();
    if (x.ptr() == nullptr) {
      log("x init empty");
      x.init_empty(N);
    }
    log("Prepared Memory");
    if constexpr (debug == Debuglevel::Verbose) { // the reference's percentage line (:428-433), driven by the per-batch poll
      _progress_mark = -1;
      cgb_ctx_set_progress(_queue.native(), &CG::progress, this);
    }
    try {
      asycl::detail::check(
          cgb_cg_solve(_queue.native(), A.csr(), b.ptr(), x.ptr(), improvement, /*max_iter = N + 1*/ -1, &_opts, &_stats));
    } catch (std::exception &e) {
      if constexpr (debug == Debuglevel::Verbose) cgb_ctx_set_progress(_queue.native(), nullptr, nullptr);
      std::cerr << "Caught Exception " << e.what() << std::endl;
      throw;
    }
    if constexpr (debug == Debuglevel::Verbose) {
      cgb_ctx_set_progress(_queue.native(), nullptr, nullptr);
      if (_progress_mark >= 0) std::clog << "\r\033[2K";
    }
    this->is_solved = true;
    if constexpr (debug == Debuglevel::Verbose) {
      std::clog << "Finished solving: " << _stats.iterations << " iterations, |r| = " << std::sqrt(_stats.rr) << " ("
                << _stats.solve_ms << " ms on the device)" << std::endl;
    }
  }

  // | sum (b - A x)^2 / sum x^2 |  — a ratio of squared norms, relative to x (reference :463-515)
  DT accuracy() {
    log("Calculating accuracy");
    double rr = 0, xx = 0;
    asycl::detail::check(cgb_residual(_queue.native(), A.csr(), x.ptr(), b.ptr(), &rr, &xx));
    return std::abs(rr / xx);
  }

  std::vector<DT> extract() {
    std::vector<DT> ret(A.N());
    this->_queue.copy(this->x.ptr(), ret.data(), A.N()).wait();
    return ret;
  }
  void extractTo(std::vector<DT> &result) {
    result.resize(A.N());
    this->_queue.copy(this->x.ptr(), result.data(), A.N()).wait();
  }

  std::size_t memoryFootprint() const { return cgb_memory_footprint((int64_t)A.N(), (int64_t)A.NNZ()); }

  // ---- additive ---------------------------------------------------------------------------------
  long long iterations() const { return (long long)_stats.iterations; } // loop trips of the last solve
  DT residualNorm() const { return std::sqrt(_stats.rr); }               // recurrence |r| after the last trip
  double seconds() const { return _stats.solve_ms * 1e-3; }              // device time of the last solve
  bool converged() const { return _stats.converged != 0; }
  const cgb_cg_stats &stats() const { return _stats; }
  cgb_cg_opts &options() { return _opts; } // loop mode, batch length, fused p-update, L2 policy, ...
  Matrix &matrix() { return A; }
  Vector &solution() { return x; }

private:
  void log(const char *what) {
    if constexpr (debug == Debuglevel::Verbose) std::clog << what << "\n";
  }
  // Same text as the reference: the trip counter as a percentage of N, refreshed every 100 trips (:428-433).  The
  // device-side loop is polled once per batch, so the line moves at the first poll past each multiple of 100.
  static void progress(int64_t trips, int64_t n, double /*rr*/, void *self) {
    CG *me = static_cast<CG *>(self);
    const long long mark = (long long)(trips / 100);
    if (mark == me->_progress_mark || n <= 0) return;
    me->_progress_mark = mark;
    std::clog << "\r\033[2K";
    std::clog << ((static_cast<double>(trips) / static_cast<double>(n)) * 100) << "%";
    std::flush(std::clog);
  }
  void executeQueue() {
    try {
      this->_queue.wait_and_throw();
    } catch (std::exception &e) {
      std::cerr << "Caught Exception " << e.what() << std::endl;
      throw;
    }
  }

  asycl::queue _queue;
  bool is_solved = false;
  Matrix A;
  Vector x;
  Vector b;
  cgb_cg_opts _opts;
  cgb_cg_stats _stats;
  long long _progress_mark = -1;
};

} // namespace CGSolver

#endif // CGB200_CG_HPP
