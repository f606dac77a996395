// matrix.cuh — the device CSR handle (replaces CGSolver::Matrix, reference
// src/LinearAlgebraTypes.hpp:57-132) plus the launch plan and the cached CG workspace.
#pragma once
#include "common.cuh"

#include <vector>

namespace cgb {

// Work vectors and scalars of CG::solve (reference allocates them per solve, twice:
// src/CG.hpp:276-301); here they live with the matrix and are reused across solves.
struct Workspace {
  int64_t n = 0;       // rows
  int64_t n_ext = 0;   // rows + halo slots (multi-GPU), == n on one GPU
  double *r = nullptr;
  double *p0 = nullptr; // p (unfused) / ping
  double *p1 = nullptr; // pong (fused p-update only, lazily allocated)
  double *Ap = nullptr;
  double *dinv = nullptr; // 1 / diag(A), built on the first Jacobi-preconditioned solve
  CgState *st = nullptr;
  double *partials = nullptr; // kMaxPartials * 2
  double *hist = nullptr;
  int64_t hist_cap = 0;
  int64_t hist_count = 0;
  CgState *h_st = nullptr; // pinned mirror for polling
  // cached CUDA graphs of one batch of iterations, keyed by the launch configuration
  cudaGraphExec_t graph = nullptr;
  uint64_t graph_key = 0;
  const double *graph_x = nullptr; // the x pointer baked into the graph
  const double *graph_b = nullptr; // the b pointer baked into it (residual replacement only)
  // kernel timing events (time_kernels)
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};

} // namespace cgb

struct cgb_csr {
  cgb_ctx *ctx = nullptr;
  int64_t n = 0;
  int64_t nnz = 0;
  int64_t n_cols = 0; // length the gathered vector must have (n, or n + halo for a local block)
  double *vals = nullptr;
  int32_t *cols = nullptr;
  int32_t *rowptr = nullptr;
  bool owns = false;
  int format = CGB_FMT_AUTO;
  // plan
  int lanes = 8;       // lanes per row of the CSR kernel
  int grid_spmv = 0;   // CTAs for SpMV-shaped kernels
  int grid_vec = 0;    // CTAs for the x/r update (+ r.r) kernel: one resident wave
  int grid_pup = 0;    // CTAs for the p-update / x-flush kernels: one resident wave
  // TMA-staged vector passes (vec_tma.cuh) for the p-update and the r-update when the vectors are long enough
  bool vec_tma = false;
  int grid_pup_tma = 0, grid_vec_tma = 0;
  bool use_tma = true; // TMA-staged row engine (spmv_tma.cuh) instead of direct loads
  int lanes_tma = 1;   // lanes per row of the TMA engine (tile = 256 / lanes_tma rows)
  int grid_tma = 0;
  int grid_sell = 0;   // CTAs of the SELL engine (8 warps each, a warp per slice)
  // row-length statistics (csr_stats, computed once on first use) and what CGB_FMT_AUTO makes of them
  bool stats_done = false;
  int max_row = 0;
  int64_t overflow_tiles = 0; // TMA tiles whose non-zeros do not fit a pipeline stage (served by slow direct loads)
  bool auto_sell = false;
  // SELL-C-sigma shadow copy
  int sell_sigma = 0;
  int64_t sell_slices = 0;
  int64_t sell_padded = 0;
  double *sell_vals = nullptr;
  int32_t *sell_cols = nullptr;
  int32_t *sell_width = nullptr; // column-major steps per slice (sell_slices)
  int64_t *sell_off = nullptr;  // slice offsets (in entries), sell_slices + 1
  int32_t *sell_perm = nullptr; // slot -> row, ~row for rows served from the long-row list (NULL: identity, no long rows)
  int32_t *sell_long = nullptr; // rows longer than kSellLongRow: processed a warp per row from the CSR arrays
  int64_t sell_nlong = 0;
  int64_t *sell_chunk = nullptr; // grid_sell * 8 + 1 slice offsets: warp w streams slices [chunk[w], chunk[w+1]) (equal cost)
  // Column panels (sell.cu: csr_build_panels): A = [A_0 | A_1 | ... ] split by column range so that the slice of
  // the gathered vector one pass touches fits in L2; every panel is a full SELL matrix over the same rows with
  // its own row order, processed by consecutive launches that accumulate into the output vector.
  std::vector<cgb_csr *> panels;
  int panels_req = 0;    // cgb_csr_set_panels: 0 = automatic (by the size of the gathered vector), 1 = none, P = P panels
  bool panels_ok = true; // false for the library's internal sub-matrices (panels themselves, multi-GPU blocks)
  // row-pattern dictionary copy (dict.cu, spmv_dict.cuh): explicit format CGB_FMT_DICT
  uint16_t *dict_ids = nullptr; // pattern of each row
  int32_t *dict_ptr = nullptr, *dict_off = nullptr;
  double *dict_val = nullptr;
  int dict_npat = 0, dict_nent = 0;
  int grid_dict = 0;
  cgb::Workspace ws;
};

namespace cgb {
int csr_plan(cgb_csr *A);
int csr_build_panels(cgb_csr *A, int P, int sigma); // sell.cu
void csr_release_panels(cgb_csr *A);
int csr_build_dict(cgb_csr *A); // dict.cu
void csr_release_dict(cgb_csr *A);
int csr_prepare(cgb_csr *A); // before any SpMV-shaped launch: statistics, CGB_FMT_AUTO decision, lazy SELL build
int workspace_ensure(cgb_csr *A, int64_t n_ext, bool need_p1, int64_t hist_cap);
void workspace_release(cgb_csr *A);
} // namespace cgb
