// utils.hpp — driver helpers with the names the reference's test/utils.hpp declares (:24-60):
// a steady_clock Timer and read_file(), the Matrix Market -> CSR entry point
// (reference test/mm_reader.cpp:154-171).  read_file is implemented by the library's
// dialect-compatible reader (cgb_mm_read, cgb200.h), so a driver needs no extra source file.
#ifndef CGB200_UTILS_HPP
#define CGB200_UTILS_HPP

#include <chrono>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "cgb200.h"

struct Timer {
  void start_measure() { _t0 = std::chrono::steady_clock::now(); }
  void stop_measure() { _elapsed = std::chrono::steady_clock::now() - _t0; }
  std::chrono::duration<double, std::milli> get_duration() { return _elapsed; }

private:
  std::chrono::steady_clock::time_point _t0;
  std::chrono::duration<double, std::milli> _elapsed{0};
};

// (values, column indices, row offsets), 0-based, rows sorted, columns ascending inside a row.
// Same dialect as the reference reader: the line after the banner is discarded, '%' lines are
// skipped, the size line is not used, every off-diagonal entry is mirrored.
inline std::tuple<std::vector<double>, std::vector<int>, std::vector<int>> read_file(std::string filename) {
  cgb_mm *m = nullptr;
  if (cgb_mm_read(filename.c_str(), &m) != CGB_OK) throw std::runtime_error(cgb_last_error());
  const std::size_t n = (std::size_t)cgb_mm_n(m), nnz = (std::size_t)cgb_mm_nnz(m);
  std::vector<double> data(cgb_mm_vals(m), cgb_mm_vals(m) + nnz);
  std::vector<int> cols(cgb_mm_cols(m), cgb_mm_cols(m) + nnz);
  std::vector<int> rows(cgb_mm_rowptr(m), cgb_mm_rowptr(m) + n + 1);
  cgb_mm_free(m);
  return {std::move(data), std::move(cols), std::move(rows)};
}

#endif // CGB200_UTILS_HPP
