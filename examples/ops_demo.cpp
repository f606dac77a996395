// ops_demo.cpp — exercises the per-op CGSolver API (Matrix / Vector / Scalar / VectorOperations) through
// the drop-in headers, the way code written against the reference's src/VectorOperations.hpp would.
// All inputs are small dyadic rationals, so every result is exact in fp64 whatever the summation
// order; tests/test_cpp_headers.py recomputes them with NumPy and compares bit for bit.
//   ops_demo [n] [m]     n = vector length (default 1000), m = 2D Poisson grid size (default 16)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "CG.hpp"

using namespace CGSolver;

static double sum(const std::vector<double> &v) {
  double s = 0;
  for (double a : v) s += a;
  return s;
}

int main(int argc, char **argv) {
  const std::size_t n = argc > 1 ? std::strtoul(argv[1], nullptr, 10) : 1000;
  const int m = argc > 2 ? std::atoi(argv[2]) : 16;
  asycl::queue q;
  std::vector<double> hx(n), hy(n);
  for (std::size_t i = 0; i < n; i++) {
    hx[i] = ((double)(i % 17) - 8.0) * 0.125;
    hy[i] = ((double)((i * 7) % 13) - 6.0) * 0.25;
  }
  Vector<double> X(q, hx), Y(q, hy), R(q, n);
  Scalar<double> res(q, 2.5), a(q, 0.5), b(q, -2.0);
  VectorOperations<double> ops(q);
  ops.setVectorSize(n);

  ops.dot_product_trivial(X, Y, res).wait(); // accumulates into the 2.5
  std::printf("dot_trivial %.17g\n", res.get());
  ops.dot_product_optimised(X, Y, res.ptr()).wait();
  std::printf("dot_optimised %.17g\n", res.get());
  ops.dot_product(X.ptr(), Y.ptr(), res).wait();
  std::printf("dot_product %.17g\n", res.get());
  Scalar<double> nr(q, 0.0);
  ops.norm(X, nr).wait();
  std::printf("norm %.17g\n", nr.get());

  std::vector<double> out(n);
  ops.saxpby(X, Y, a, b, R);
  q.copy(R.ptr(), out.data(), n).wait();
  std::printf("saxpby %.17g\n", sum(out));
  ops.sapbx(X, Y, b, R);
  q.copy(R.ptr(), out.data(), n).wait();
  std::printf("sapbx %.17g\n", sum(out));
  ops.sambx(X, Y, b, X); // in place, as in the CG loop
  q.copy(X.ptr(), out.data(), n).wait();
  std::printf("sambx_inplace %.17g\n", sum(out));

  // 2D 5-point Laplacian, natural ordering
  std::vector<double> vals;
  std::vector<int> cols, rows(1, 0);
  for (int j = 0; j < m; j++)
    for (int i = 0; i < m; i++) {
      const int r = i + m * j;
      if (j > 0) { cols.push_back(r - m); vals.push_back(-1.0); }
      if (i > 0) { cols.push_back(r - 1); vals.push_back(-1.0); }
      cols.push_back(r); vals.push_back(4.0);
      if (i < m - 1) { cols.push_back(r + 1); vals.push_back(-1.0); }
      if (j < m - 1) { cols.push_back(r + m); vals.push_back(-1.0); }
      rows.push_back((int)cols.size());
    }
  const std::size_t N = (std::size_t)m * m;
  Matrix<double> A(q, vals, cols, rows);
  std::vector<double> hv(N);
  for (std::size_t i = 0; i < N; i++) hv[i] = (double)((i * 5) % 11) - 5.0;
  Vector<double> V(q, hv), AV(q, N);
  ops.setVectorSize(N);
  ops.spmv(A, V, AV, A.NNZ()).wait();
  std::vector<double> hav(N);
  q.copy(AV.ptr(), hav.data(), N).wait();
  double s1 = 0, s2 = 0;
  for (std::size_t i = 0; i < N; i++) {
    s1 += hav[i] * (double)(i % 7);
    s2 += hav[i] * hav[i];
  }
  std::printf("spmv %zu %zu %.17g %.17g\n", (std::size_t)A.N(), (std::size_t)A.NNZ(), s1, s2);

  // whole solve through CG with device objects moved in (setMatrix(Matrix&&) / setTarget(Vector&&))
  std::vector<double> hb(N);
  for (std::size_t i = 0; i < N; i++) hb[i] = (double)(i + 1);
  CG<double> cg(q);
  cg.setMatrix(std::move(A));
  cg.setTarget(Vector<double>(q, hb));
  try {
    CG<double> empty(q);
    empty.solve(1e-8);
  } catch (std::runtime_error &e) {
    std::printf("missing %s\n", e.what());
  }
  cg.solve(1e-6);
  std::printf("solve %d %lld %d %.3e %zu\n", cg.getDimension(), cg.iterations(), (int)cg.converged(), cg.accuracy(),
              cg.memoryFootprint());
  return 0;
}
