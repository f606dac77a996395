// tester.cpp — driver with the protocol of the reference's test/Tester.cpp (:12-57):
//   tester file.mtx [tolerance]
// reads a Matrix Market file, solves A x = b with b = 1..n and x0 = 0, prints
//   dim NNZ elapsed_ms accuracy
// on stdout.  `tolerance` defaults to the reference's 1e-24 (absolute, on the recurrence residual);
// "rel:1e-8" means 1e-8 * ||b||_2.  Extra information (iterations, final |r|) goes to stderr.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "CG.hpp"
#include "utils.hpp"

using namespace CGSolver;

int main(int argc, char *argv[]) {
  if (argc < 2 || argc > 3) {
    std::cout << "Usage: " << argv[0] << " filename [abs_tol | rel:<rtol>]" << std::endl;
    std::cout << "Prints: dimension  non-zeros  solve time in ms  |b-Ax|^2/|x|^2" << std::endl;
    return 1;
  }
  auto [data, cols, rows] = read_file(argv[1]);
  const std::size_t nnz = data.size(), n = rows.size() - 1;
  std::vector<double> target(n);
  double bb = 0.0;
  for (std::size_t i = 0; i < n; i++) {
    target[i] = (double)(i + 1);
    bb += target[i] * target[i];
  }
  double tol = 1e-24;
  if (argc == 3) tol = std::strncmp(argv[2], "rel:", 4) == 0 ? std::atof(argv[2] + 4) * std::sqrt(bb) : std::atof(argv[2]);

  auto cg = CG<double>::createCG();
  cg->setMatrix(data, cols, rows);
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
