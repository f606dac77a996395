// tester.cpp — driver with the protocol of the reference's test/Tester.cpp (:12-57):
//   tester file.mtx [tolerance] [device]
// reads a Matrix Market file, solves A x = b with b = 1..n and x0 = 0, prints
//   dim NNZ elapsed_ms accuracy
// on stdout.  `tolerance` defaults to the reference's 1e-24 (absolute, on the recurrence residual);
// "rel:1e-8" means 1e-8 * ||b||_2.  With "device" the file goes straight to the GPU (Matrix::from_file: parallel
// parse, CSR assembled on the device) instead of through read_file + setMatrix(vectors).  Extra information
// (iterations, final |r|) goes to stderr.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "CG.hpp"
#include "utils.hpp"

using namespace CGSolver;

int main(int argc, char *argv[]) {
  if (argc < 2 || argc > 4) {
    std::cout << "Usage: " << argv[0] << " filename [abs_tol | rel:<rtol>] [device]" << std::endl;
    std::cout << "Prints: dimension  non-zeros  solve time in ms  |b-Ax|^2/|x|^2" << std::endl;
    return 1;
  }
  const bool on_device = argc == 4 && std::strcmp(argv[3], "device") == 0;
  asycl::queue q;
  CG<double> solver(q);
  CG<double> *cg = &solver;
  std::size_t nnz = 0, n = 0;
  if (on_device) {
    auto A = Matrix<double>::from_file(q, argv[1]);
    nnz = A.NNZ();
    n = A.N();
    cg->setMatrix(std::move(A));
  } else {
    auto [data, cols, rows] = read_file(argv[1]);
    nnz = data.size();
    n = rows.size() - 1;
    cg->setMatrix(data, cols, rows);
    q.wait();
  }
  std::vector<double> target(n);
  double bb = 0.0;
  for (std::size_t i = 0; i < n; i++) {
    target[i] = (double)(i + 1);
    bb += target[i] * target[i];
  }
  double tol = 1e-24;
  if (argc >= 3) tol = std::strncmp(argv[2], "rel:", 4) == 0 ? std::atof(argv[2] + 4) * std::sqrt(bb) : std::atof(argv[2]);

  cg->setTarget(target);
  Timer t;
  t.start_measure();
  cg->solve(tol);
  t.stop_measure();
  auto x = cg->extract();
  const double acc = cg->accuracy();
  std::cout << cg->getDimension() << " " << nnz << " " << t.get_duration().count() << " " << acc << std::endl;
  std::cerr << "iterations " << cg->iterations() << " residual_norm " << cg->residualNorm() << " device_seconds "
            << cg->seconds() << " footprint " << cg->memoryFootprint() << std::endl;
  return 0;
}
