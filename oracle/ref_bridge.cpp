// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI bridge around the UNMODIFIED reference sources, compiled where they lie under
// /root/reference (include paths only; nothing is copied) against oracle/sycl_standin.
// Output goes to oracle/_ref/ (see oracle/Makefile). Python tests load it with ctypes to
//   * pin the C restatement (oracle/cg_oracle.c) against the real reference, and
//   * time the reference's CPU path (OpenMP flavour) as the CPU baseline in bench.py.
//
// Reference entry points exercised:
//   CGSolver::CG<double>            src/CG.hpp:53-601   (setMatrix, setTarget, setInital, solve, accuracy, extract)
//   CGSolver::VectorOperations      src/VectorOperations.hpp:39-488 (every public op)
//   read_file                       test/mm_reader.cpp:154-171
#include "CG.hpp"
#include "utils.hpp"

#include <chrono>
#include <cstring>
#include <string>
#include <tuple>
#include <vector>

using DT = double;
namespace sy = acpp::sycl;
using namespace CGSolver;

namespace {
struct MMHandle {
  std::vector<double> data;
  std::vector<int> cols, rows;
};
} // namespace

extern "C" {

int ref_is_openmp() {
#ifdef STANDIN_OPENMP
  return 1;
#else
  return 0;
#endif
}

int ref_num_threads() {
#ifdef STANDIN_OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// Thread count of the OpenMP flavour (no-op for the serial one).
void ref_set_threads(int t) {
#ifdef STANDIN_OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}

// Bounded timing samples: stop the reference loop after `trips` trips (0 = off). See the stand-in.
void ref_set_trip_limit(long trips) { sy::standin_stats().trip_limit = trips; }

// Whole solve through the reference's CG class, exactly as test/Tester.cpp:36-53 drives it.
// iterations = number of AlphaCalculation submissions (one per loop trip, src/CG.hpp:385).
int ref_cg_solve(long n, long nnz, const double *vals, const int *cols, const int *rows, const double *b,
                 const double *x0 /* may be NULL */, double improvement, double *x_out, long *iterations,
                 double *accuracy, double *solve_ms, long *footprint) {
  try {
    std::vector<DT> data(vals, vals + nnz);
    std::vector<int> c(cols, cols + nnz), r(rows, rows + n + 1);
    std::vector<DT> target(b, b + n);
    auto cgp = CG<DT, Debuglevel::None>::createCG();
    auto cg = *cgp;
    cg.setMatrix(data, c, r);
    cg.setTarget(target);
    if (x0) {
      std::vector<DT> init(x0, x0 + n);
      cg.setInital(init);
    }
    sy::standin_stats().clear();
    auto t0 = std::chrono::steady_clock::now();
    cg.solve(improvement);
    auto t1 = std::chrono::steady_clock::now();
    if (iterations) *iterations = sy::standin_stats().count_containing("AlphaCalculation");
    if (solve_ms) *solve_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (x_out) {
      std::vector<DT> x;
      cg.extractTo(x);
      std::memcpy(x_out, x.data(), sizeof(DT) * n);
    }
    if (accuracy) *accuracy = cg.accuracy();
    if (footprint) *footprint = static_cast<long>(cg.memoryFootprint());
    if (cg.getDimension() != n) return 2;
    return 0;
  } catch (std::exception &e) {
    std::fprintf(stderr, "ref_cg_solve: %s\n", e.what());
    return 1;
  }
}

// Persistent flavour for timing loops: the matrix is set once (setMatrix), every run sets b and a
// zero initial guess and times solve() alone, exactly the region test/Tester.cpp:43-46 times.
struct RefCG {
  std::unique_ptr<CG<DT, Debuglevel::None>> cgp;
  std::vector<DT> data;
  std::vector<int> c, r;
  long n;
};
void *ref_cg_create(long n, long nnz, const double *vals, const int *cols, const int *rows) {
  auto *h = new RefCG;
  h->n = n;
  h->data.assign(vals, vals + nnz);
  h->c.assign(cols, cols + nnz);
  h->r.assign(rows, rows + n + 1);
  h->cgp = CG<DT, Debuglevel::None>::createCG();
  h->cgp->setMatrix(h->data, h->c, h->r);
  // the host copies are only needed until the (synchronous) stand-in copy is done
  std::vector<DT>().swap(h->data);
  std::vector<int>().swap(h->c);
  std::vector<int>().swap(h->r);
  return h;
}
int ref_cg_run(void *handle, const double *b, double improvement, long trip_limit, double *x_out, long *iterations,
               double *solve_ms) {
  auto *h = static_cast<RefCG *>(handle);
  try {
    std::vector<DT> target(b, b + h->n), zeros(h->n, 0.0);
    h->cgp->setTarget(target);
    h->cgp->setInital(zeros);
    sy::standin_stats().clear();
    sy::standin_stats().trip_limit = trip_limit;
    auto t0 = std::chrono::steady_clock::now();
    h->cgp->solve(improvement);
    auto t1 = std::chrono::steady_clock::now();
    sy::standin_stats().trip_limit = 0;
    if (iterations) *iterations = sy::standin_stats().count_containing("AlphaCalculation");
    if (solve_ms) *solve_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (x_out) {
      std::vector<DT> x;
      h->cgp->extractTo(x);
      std::memcpy(x_out, x.data(), sizeof(DT) * h->n);
    }
    return 0;
  } catch (std::exception &e) {
    std::fprintf(stderr, "ref_cg_run: %s\n", e.what());
    return 1;
  }
}
void ref_cg_destroy(void *handle) { delete static_cast<RefCG *>(handle); }

// Error behaviour of solve() with missing inputs (src/CG.hpp:266-272): returns the what() text.
int ref_cg_solve_missing(int have_matrix, int have_rhs, char *msg, int msg_len) {
  std::vector<DT> data{2.0}, target{1.0};
  std::vector<int> c{0}, r{0, 1};
  auto cgp = CG<DT, Debuglevel::None>::createCG();
  auto cg = *cgp;
  if (have_matrix) cg.setMatrix(data, c, r);
  if (have_rhs) cg.setTarget(target);
  try {
    cg.solve(1e-12);
  } catch (std::runtime_error &e) {
    std::snprintf(msg, msg_len, "%s", e.what());
    return 1;
  }
  msg[0] = 0;
  return 0;
}

// ---- per-op entry points (VectorOperations) -------------------------------------------------
// which: 0 = dot_product_trivial, 1 = dot_product_optimised, 2 = dot_product (deprecated), 3 = norm
double ref_reduce(int which, long n, const double *x, const double *y, double initial) {
  sy::queue q;
  std::vector<DT> hx(x, x + n), hy(y ? y : x, (y ? y : x) + n);
  Vector<DT> X(q, hx), Y(q, hy);
  Scalar<DT> res(q, initial);
  VectorOperations<DT> ops(q);
  ops.setVectorSize(n);
  switch (which) {
  case 0: ops.dot_product_trivial(X, Y, res); break;
  case 1: ops.dot_product_optimised(X, Y, res.ptr()); break;
  case 2: ops.dot_product(X.ptr(), Y.ptr(), res.ptr()); break;
  default: ops.norm(X, res); break;
  }
  q.wait();
  return *res.ptr();
}

// which: 0 = saxpby (a*x + b*y), 1 = sapbx (x + b*y), 2 = sambx (x - b*y)
// alias: 0 = separate result, 1 = result aliases X, 2 = result aliases Y (src/CG.hpp:390,392,418)
void ref_axpy(int which, int alias, long n, const double *x, const double *y, double a, double b, double *out) {
  sy::queue q;
  std::vector<DT> hx(x, x + n), hy(y, y + n);
  Vector<DT> X(q, hx), Y(q, hy), R(q, static_cast<std::size_t>(n));
  Scalar<DT> A(q, a), B(q, b);
  VectorOperations<DT> ops(q);
  ops.setVectorSize(n);
  Vector<DT> &Res = alias == 1 ? X : (alias == 2 ? Y : R);
  switch (which) {
  case 0: ops.saxpby(X, Y, A.ptr(), B.ptr(), Res); break;
  case 1: ops.sapbx(X, Y, B.ptr(), Res); break;
  default: ops.sambx(X, Y, B.ptr(), Res); break;
  }
  q.wait();
  std::memcpy(out, Res.ptr(), sizeof(DT) * n);
}

void ref_spmv(long n, long nnz, const double *vals, const int *cols, const int *rows, const double *v, double *out) {
  sy::queue q;
  std::vector<DT> data(vals, vals + nnz), hv(v, v + n);
  std::vector<int> c(cols, cols + nnz), r(rows, rows + n + 1);
  Matrix<DT> A(q, data, c, r);
  Vector<DT> V(q, hv), R(q, static_cast<std::size_t>(n));
  VectorOperations<DT> ops(q);
  ops.setVectorSize(n);
  ops.spmv(A, V, R, nnz);
  q.wait();
  std::memcpy(out, R.ptr(), sizeof(DT) * n);
}

// ---- Matrix Market reader (test/mm_reader.cpp:154) ---------------------------------------------
void *ref_mm_open(const char *path, long *n, long *nnz) {
  auto *h = new MMHandle;
  std::tie(h->data, h->cols, h->rows) = read_file(std::string(path));
  *n = static_cast<long>(h->rows.size()) - 1;
  *nnz = static_cast<long>(h->data.size());
  return h;
}
void ref_mm_copy(void *handle, double *vals, int *cols, int *rows) {
  auto *h = static_cast<MMHandle *>(handle);
  std::memcpy(vals, h->data.data(), sizeof(double) * h->data.size());
  std::memcpy(cols, h->cols.data(), sizeof(int) * h->cols.size());
  std::memcpy(rows, h->rows.data(), sizeof(int) * h->rows.size());
}
void ref_mm_close(void *handle) { delete static_cast<MMHandle *>(handle); }

} // extern "C"
