// e2e_cpp.cpp — the reference-facing path end to end, timed the way a user of the reference API meets it:
// plain std::vector storage in (pageable memory), std::vector out, through the drop-in headers.
//
//   e2e_cpp [m = 128] [steps = 3] [warmup = 1] [rtol = 1e-8]
//
// Builds the 3D 7-point Poisson matrix m^3 (BASELINE config C2 at m = 128) in std::vectors, then per step does what
// the reference's driver does between reading the file and printing (test/Tester.cpp:36-49):
//   createCG, setMatrix(vectors), setTarget(vector), solve(tol), extract()
// and prints ONE JSON object on stdout: iterations per second over the timed steps (wall clock, steady_clock, the
// reference's Timer) and the bytes that crossed PCIe per step.  bench.py runs it for its `e2e_cpp` block.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "CG.hpp"

using namespace CGSolver;

int main(int argc, char **argv) {
  const long m = argc > 1 ? std::atol(argv[1]) : 128;
  const int steps = argc > 2 ? std::atoi(argv[2]) : 3;
  const int warmup = argc > 3 ? std::atoi(argv[3]) : 1;
  const double rtol = argc > 4 ? std::atof(argv[4]) : 1e-8;
  const long n = m * m * m;
  std::vector<double> data;
  std::vector<int> cols, rows(1, 0);
  data.reserve((size_t)7 * n);
  cols.reserve((size_t)7 * n);
  rows.reserve((size_t)n + 1);
  for (long k = 0; k < m; k++)
    for (long j = 0; j < m; j++)
      for (long i = 0; i < m; i++) {
        const long row = i + m * j + m * m * k;
        auto put = [&](long c, double v) {
          cols.push_back((int)c);
          data.push_back(v);
        };
        if (k > 0) put(row - m * m, -1.0);
        if (j > 0) put(row - m, -1.0);
        if (i > 0) put(row - 1, -1.0);
        put(row, 6.0);
        if (i < m - 1) put(row + 1, -1.0);
        if (j < m - 1) put(row + m, -1.0);
        if (k < m - 1) put(row + m * m, -1.0);
        rows.push_back((int)data.size());
      }
  std::vector<double> target((size_t)n);
  double bb = 0.0;
  for (long i = 0; i < n; i++) {
    target[(size_t)i] = (double)(i + 1);
    bb += target[(size_t)i] * target[(size_t)i];
  }
  const double tol = rtol * std::sqrt(bb);
  const bool trace = std::getenv("E2E_TRACE") != nullptr;
  long long iters = 0, last_iters = 0;
  double seconds = 0.0, device_seconds = 0.0, checksum = 0.0;
  for (int s = 0; s < warmup + steps; s++) {
    const auto t0 = std::chrono::steady_clock::now();
    auto cg = CG<double>::createCG();
    const auto ta = std::chrono::steady_clock::now();
    cg->setMatrix(data, cols, rows);
    const auto tb = std::chrono::steady_clock::now();
    cg->setTarget(target);
    const auto tc = std::chrono::steady_clock::now();
    cg->solve(tol);
    const auto td = std::chrono::steady_clock::now();
    std::vector<double> x = cg->extract();
    const auto t1 = std::chrono::steady_clock::now();
    if (trace) { // E2E_TRACE=1: where a step's wall clock goes
      auto ms = [](auto a, auto b) { return 1e3 * std::chrono::duration<double>(b - a).count(); };
      std::fprintf(stderr, "[e2e_cpp] step %d: createCG %.2f | setMatrix %.2f | setTarget %.2f | solve %.2f (device %.2f) | extract %.2f | total %.2f ms\n",
                   s, ms(t0, ta), ms(ta, tb), ms(tb, tc), ms(tc, td), 1e3 * cg->seconds(), ms(td, t1), ms(t0, t1));
    }
    if (s >= warmup) {
      iters += cg->iterations();
      seconds += std::chrono::duration<double>(t1 - t0).count();
      device_seconds += cg->seconds();
    }
    last_iters = cg->iterations();
    checksum = 0.0;
    for (double v : x) checksum += v * v;
  }
  const double h2d = 12.0 * (double)data.size() + 4.0 * (double)(n + 1) + 8.0 * (double)n, d2h = 8.0 * (double)n;
  std::printf("{\"value\": %.6f, \"unit\": \"iterations/s\", \"steps\": %d, \"warmup\": %d, \"ms_per_step\": %.4f, "
              "\"iterations_per_solve\": %lld, \"device_ms_per_solve\": %.4f, \"h2d_bytes_per_step\": %.0f, "
              "\"d2h_bytes_per_step\": %.0f, \"x_norm2\": %.17g, \"n\": %ld, \"nnz\": %zu}\n",
              (double)iters / seconds, steps, warmup, 1e3 * seconds / steps, last_iters, 1e3 * device_seconds / steps, h2d, d2h,
              std::sqrt(checksum), n, data.size());
  return 0;
}
