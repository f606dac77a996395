// csr.cu — device CSR handle, launch plan, plain SpMV, residual (accuracy) and the on-device
// stencil generators.
// Interfaces replaced: CGSolver::Matrix (reference src/LinearAlgebraTypes.hpp:57-132),
// VectorOperations::spmv (src/VectorOperations.hpp:438-466), CG::accuracy (src/CG.hpp:463-515).
#include "launch.cuh"

#include <algorithm>

#include <cub/device/device_scan.cuh>
#include <stdlib.h>

using namespace cgb;

namespace cgb {

int csr_plan(cgb_csr *A) {
  const double mean = A->n > 0 ? (double)A->nnz / (double)A->n : 1.0;
  int lanes = 2;
  while (lanes < 32 && (double)lanes < mean) lanes <<= 1;
  A->lanes = lanes;
  const int sms = A->ctx->sm_count;
  A->grid_spmv = sms * 8;
  // SELL engine: every warp owns one cost-balanced range of slices; many more CTAs than fit at once (24 per SM,
  // measured on C5: 8 -> 1.17 ms, 16 -> 1.11, 24 -> 1.09), so the hardware scheduler evens out what the static cost
  // model misses.  Never more CTAs than reduction partials.
  int sell_per_sm = 24;
  if (const char *e = getenv("CGB_SELL_CTAS")) sell_per_sm = atoi(e) > 0 ? atoi(e) : sell_per_sm;
  A->grid_sell = sms * sell_per_sm;
  if (A->grid_sell > kMaxPartials) A->grid_sell = kMaxPartials;
  int occ_d = 8; // dictionary engine: a thread per row, grid-stride, one resident wave
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_d, cg_spmv_kernel<1, kEngDict, false>, kBlock, 0);
  if (const char *e = getenv("CGB_DICT_CTAS")) occ_d = atoi(e) > 0 ? atoi(e) : occ_d;
  A->grid_dict = sms * (occ_d > 0 ? occ_d : 1);
  if ((A->n + kBlock - 1) / kBlock < A->grid_dict) A->grid_dict = (int)std::max<int64_t>(1, (A->n + kBlock - 1) / kBlock);
  if ((A->n + 255) / 256 < A->grid_sell) A->grid_sell = (int)((A->n + 255) / 256 > 0 ? (A->n + 255) / 256 : 1);
  // streaming vector kernels: exactly one resident wave (grid-stride loops), so no partial second wave
  // leaves SMs idle at the tail; CGB_VEC_CTAS caps the CTAs per SM for A/B measurements
  int occ_u = 8, occ_p = 8;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_u, cg_update_kernel, kBlock, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_p, cg_pupdate_kernel, kBlock, 0);
  int cap = 8;
  if (const char *e = getenv("CGB_VEC_CTAS")) cap = atoi(e) > 0 ? atoi(e) : cap;
  A->grid_vec = sms * (occ_u < cap ? (occ_u > 0 ? occ_u : 1) : cap);
  A->grid_pup = sms * (occ_p < cap ? (occ_p > 0 ? occ_p : 1) : cap);
  // never launch more CTAs than there is work for
  const int64_t rows_per_cta = (int64_t)(kBlock / lanes) * 2;
  const int64_t need = (A->n + rows_per_cta - 1) / rows_per_cta;
  if (need < A->grid_spmv) A->grid_spmv = (int)(need > 0 ? need : 1);
  const int64_t need_v = (A->n / 2 + kBlock - 1) / kBlock;
  if (need_v < A->grid_vec) A->grid_vec = (int)(need_v > 0 ? need_v : 1);
  if (need_v < A->grid_pup) A->grid_pup = (int)(need_v > 0 ? need_v : 1);
  // The same two passes with TMA-staged tiles: 10 % faster than the loops above on long vectors (C4: K3 0.875 -> 0.78 ms),
  // slower on short ones (the two-stage pipeline wants >= 8 tiles per CTA).  CGB_VEC_TMA=0/1 forces it (A/B runs).
  {
    int occ_pt = 0, occ_ut = 0;
    cudaFuncSetAttribute(cg_pupdate_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vec_tma_smem_bytes(3));
    cudaFuncSetAttribute(cg_update_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vec_tma_smem_bytes(2));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_pt, cg_pupdate_tma_kernel, kTmaThreads, vec_tma_smem_bytes(3));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_ut, cg_update_tma_kernel, kTmaThreads, vec_tma_smem_bytes(2));
    cudaGetLastError();
    A->grid_pup_tma = sms * std::max(1, std::min(occ_pt, 4));
    A->grid_vec_tma = std::min(sms * std::max(1, std::min(occ_ut, 4)), kMaxPartials);
    A->vec_tma = occ_pt > 0 && occ_ut > 0 && A->n >= (int64_t)A->grid_pup_tma * 8 * kVecTile;
    if (const char *e = getenv("CGB_VEC_TMA")) A->vec_tma = atoi(e) != 0 && occ_pt > 0 && occ_ut > 0;
  }
  // TMA engine: a tile of 256 / L rows should carry about 7 * 256 non-zeros
  int lt = 1;
  while (lt < 32 && (double)lt * 7.0 < mean) lt <<= 1;
  if (const char *e = getenv("CGB_TMA_LANES")) lt = atoi(e) > 0 ? atoi(e) : lt; // A/B measurements
  A->lanes_tma = lt;
  const int tile_rows = kTmaConsumers / lt;
  const size_t smem = tma_smem_bytes(tile_rows) + 1024;
  int per_sm = (int)((size_t)227 * 1024 / smem);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  A->grid_tma = sms * per_sm;
  const int64_t ntiles = (A->n + tile_rows - 1) / tile_rows;
  if (ntiles < A->grid_tma) A->grid_tma = (int)(ntiles > 0 ? ntiles : 1);
  // bulk copies need 16-byte aligned array bases; tiny systems gain nothing from staging
  auto al16 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  const char *env = getenv("CGB_SPMV_ENGINE"); // "vec" forces direct loads (A/B measurements)
  A->use_tma = al16(A->vals) && al16(A->cols) && al16(A->rowptr) && ntiles >= 2 && !(env && !strcmp(env, "vec"));
  return CGB_OK;
}

// max row length and number of TMA tiles that overflow a stage: one pass over the row offsets
__global__ void csr_stats_kernel(int64_t n, const int *__restrict__ rowptr, int tile_rows, int cap, int *max_row,
                                 unsigned long long *overflow) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int len = 0;
  if (t < n) len = rowptr[t + 1] - rowptr[t];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) len = max(len, __shfl_xor_sync(0xffffffffu, len, o));
  if ((threadIdx.x & 31) == 0 && len > 0) atomicMax(max_row, len);
  if (t < n && t % tile_rows == 0) {
    const int64_t e = t + tile_rows < n ? t + tile_rows : n;
    if (rowptr[e] - rowptr[t] + 8 > cap) atomicAdd(overflow, 1ull);
  }
}

constexpr int kSellDefaultSigma = 32768;
constexpr int64_t kPanelMinBytes = (int64_t)48 << 20; // gathered vectors up to this size are left in one piece
constexpr int64_t kPanelBytes = (int64_t)40 << 20;    // target slice of the gathered vector per panel (C5: 2 x 40 MB wins)

int csr_prepare(cgb_csr *A) {
  if (A->n == 0) return CGB_OK;
  cgb_ctx *ctx = A->ctx;
  if (!A->stats_done) {
    int *d_max = nullptr;
    CGB_CUDA(cudaMalloc(&d_max, 16));
    CGB_CUDA(cudaMemsetAsync(d_max, 0, 16, ctx->stream));
    const int tile_rows = kTmaConsumers / A->lanes_tma;
    csr_stats_kernel<<<(unsigned)((A->n + 255) / 256), 256, 0, ctx->stream>>>(A->n, A->rowptr, tile_rows, kTmaCap, d_max,
                                                                             (unsigned long long *)(d_max + 2));
    int h[4] = {0, 0, 0, 0};
    cudaError_t e = cudaMemcpyAsync(h, d_max, 16, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_max);
    if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "csr_prepare: %s", cudaGetErrorString(e));
    A->max_row = h[0];
    memcpy(&A->overflow_tiles, h + 2, 8);
    A->stats_done = true;
    // CGB_FMT_AUTO: irregular row lengths (tiles that overflow the TMA stage, or a long tail) -> SELL-32-sigma
    const double mean = (double)A->nnz / (double)A->n;
    const int64_t ntiles = (A->n + tile_rows - 1) / tile_rows;
    A->auto_sell = A->n >= 4096 && (A->overflow_tiles * 50 > ntiles || (double)A->max_row > 8.0 * mean + 64.0);
    if (const char *env = getenv("CGB_AUTO_SELL")) A->auto_sell = atoi(env) != 0;
  }
  if (A->format == CGB_FMT_DICT && !A->dict_ids) CGB_TRY(csr_build_dict(A));
  if (csr_wants_sell(A)) {
    // Column panels: the gathered vector is read at random columns; when it is larger than what L2 keeps
    // (the matrix streams through with evict-first, but a B200 L2 half holds ~60 MB), most gathers go to
    // DRAM at a 32-byte sector per 8-byte value.  Splitting A by column ranges of <= kPanelBytes of the
    // vector makes each pass's gathers L2 hits, for one extra read+write of the output per extra panel.
    int P = A->panels_req;
    if (const char *env = getenv("CGB_SELL_PANELS")) P = atoi(env);
    if (P <= 0) {
      const int64_t bytes = 8 * A->n_cols;
      P = bytes > kPanelMinBytes ? (int)((bytes + kPanelBytes - 1) / kPanelBytes) : 1;
    }
    if (!A->panels_ok) P = 1;
    if (P > 64) P = 64;
    if (P > 1) {
      if ((int)A->panels.size() != P) CGB_TRY(csr_build_panels(A, P, A->sell_sigma > 0 ? A->sell_sigma : kSellDefaultSigma));
    } else {
      csr_release_panels(A);
      if (!A->sell_vals) CGB_TRY(cgb_csr_build_sell(A, kSellDefaultSigma));
    }
  }
  return CGB_OK;
}

void workspace_release(cgb_csr *A) {
  Workspace &w = A->ws;
  if (w.graph) cudaGraphExecDestroy(w.graph);
  cudaFree(w.r);
  cudaFree(w.p0);
  cudaFree(w.p1);
  cudaFree(w.Ap);
  cudaFree(w.dinv);
  cudaFree(w.st);
  cudaFree(w.partials);
  cudaFree(w.hist);
  if (w.h_st) cached_host_free(w.h_st);
  for (auto &e : w.ev)
    if (e) cudaEventDestroy(e);
  w = Workspace();
}

int workspace_ensure(cgb_csr *A, int64_t n_ext, bool need_p1, int64_t hist_cap) {
  Workspace &w = A->ws;
  cgb_ctx *ctx = A->ctx;
  if (w.n != A->n || w.n_ext != n_ext) {
    workspace_release(A);
    const size_t vb = sizeof(double) * (size_t)(n_ext > 0 ? n_ext : 1);
    CGB_CUDA(cudaMalloc(&w.r, vb));
    CGB_CUDA(cudaMalloc(&w.p0, vb));
    CGB_CUDA(cudaMalloc(&w.Ap, vb));
    CGB_CUDA(cudaMalloc(&w.st, sizeof(CgState)));
    CGB_CUDA(cudaMalloc(&w.partials, sizeof(double) * 2 * kMaxPartials));
    CGB_CUDA(cached_host_alloc((void **)&w.h_st, 2 * sizeof(CgState)));
    CGB_CUDA(cudaMemsetAsync(w.st, 0, sizeof(CgState), ctx->stream));
    w.n = A->n;
    w.n_ext = n_ext;
  }
  if (need_p1 && !w.p1) {
    CGB_CUDA(cudaMalloc(&w.p1, sizeof(double) * (size_t)(n_ext > 0 ? n_ext : 1)));
    if (w.graph) {
      cudaGraphExecDestroy(w.graph);
      w.graph = nullptr;
    }
  }
  if (hist_cap > w.hist_cap) {
    cudaFree(w.hist);
    w.hist = nullptr;
    CGB_CUDA(cudaMalloc(&w.hist, sizeof(double) * (size_t)hist_cap));
    w.hist_cap = hist_cap;
    if (w.graph) {
      cudaGraphExecDestroy(w.graph);
      w.graph = nullptr;
    }
  }
  return CGB_OK;
}

// ---- stencil generation -----------------------------------------------------------------------
// kind 0: 2D 5-point (diag 4), 1: 3D 7-point (diag 6), 2: 3D 27-point (diag 26); neighbours -1;
// natural ordering row = i + m*j + m*m*k; Dirichlet truncation at the boundary.
__device__ __forceinline__ int stencil_row_len(int kind, int64_t m, int64_t mz, int64_t row) {
  const int64_t i = row % m, j = (row / m) % m, k = row / (m * m);
  const int nx = 1 + (i > 0) + (i < m - 1), ny = 1 + (j > 0) + (j < m - 1);
  if (kind == 0) return 1 + (nx - 1) + (ny - 1);
  const int nz = 1 + (k > 0) + (k < mz - 1);
  if (kind == 1) return 1 + (nx - 1) + (ny - 1) + (nz - 1);
  return nx * ny * nz;
}

__global__ void stencil_len_kernel(int kind, int64_t m, int64_t mz, int64_t row_begin, int64_t nloc, int *len) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nloc) len[t] = stencil_row_len(kind, m, mz, row_begin + t);
}

__global__ void stencil_fill_kernel(int kind, int64_t m, int64_t mz, int64_t row_begin, int64_t nloc,
                                    const int *__restrict__ rowptr, int *__restrict__ cols,
                                    double *__restrict__ vals) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nloc) return;
  const int64_t row = row_begin + t;
  const int64_t i = row % m, j = (row / m) % m, k = row / (m * m);
  int64_t o = rowptr[t];
  if (kind == 2) {
    for (int dz = -1; dz <= 1; dz++) {
      if (k + dz < 0 || k + dz >= mz) continue;
      for (int dy = -1; dy <= 1; dy++) {
        if (j + dy < 0 || j + dy >= m) continue;
        for (int dx = -1; dx <= 1; dx++) {
          if (i + dx < 0 || i + dx >= m) continue;
          const int64_t c = row + dx + dy * m + dz * m * m;
          cols[o] = (int)c;
          vals[o] = (c == row) ? 26.0 : -1.0;
          o++;
        }
      }
    }
    return;
  }
  const double diag = kind == 0 ? 4.0 : 6.0;
  if (kind == 1 && k > 0) { cols[o] = (int)(row - m * m); vals[o++] = -1.0; }
  if (j > 0) { cols[o] = (int)(row - m); vals[o++] = -1.0; }
  if (i > 0) { cols[o] = (int)(row - 1); vals[o++] = -1.0; }
  cols[o] = (int)row; vals[o++] = diag;
  if (i < m - 1) { cols[o] = (int)(row + 1); vals[o++] = -1.0; }
  if (j < m - 1) { cols[o] = (int)(row + m); vals[o++] = -1.0; }
  if (kind == 1 && k < mz - 1) { cols[o] = (int)(row + m * m); vals[o++] = -1.0; }
}

} // namespace cgb

extern "C" {

int cgb_csr_from_device(cgb_ctx *ctx, int64_t n, int64_t nnz, double *d_vals, int32_t *d_cols, int32_t *d_rowptr,
                        int take_ownership, cgb_csr **out) {
  CGB_REQUIRE(ctx && out, "cgb_csr_from_device: NULL argument");
  CGB_REQUIRE(n >= 0 && nnz >= 0 && nnz < ((int64_t)1 << 31) && n < ((int64_t)1 << 31),
              "cgb_csr: n and nnz must fit the reference's 32-bit indices");
  CGB_REQUIRE(d_rowptr && (nnz == 0 || (d_vals && d_cols)), "cgb_csr_from_device: NULL array");
  cgb_csr *A = new cgb_csr();
  A->ctx = ctx;
  A->n = n;
  A->nnz = nnz;
  A->n_cols = n;
  A->vals = d_vals;
  A->cols = d_cols;
  A->rowptr = d_rowptr;
  A->owns = take_ownership != 0;
  csr_plan(A);
  *out = A;
  return CGB_OK;
}

int cgb_csr_create(cgb_ctx *ctx, int64_t n, int64_t nnz, const double *h_vals, const int32_t *h_cols,
                   const int32_t *h_rowptr, cgb_csr **out) {
  CGB_REQUIRE(ctx && out, "cgb_csr_create: NULL argument");
  CGB_REQUIRE(h_rowptr && (nnz == 0 || (h_vals && h_cols)), "cgb_csr_create: NULL array");
  CGB_REQUIRE(n >= 0 && nnz >= 0 && nnz < ((int64_t)1 << 31) && n < ((int64_t)1 << 31),
              "cgb_csr: n and nnz must fit the reference's 32-bit indices");
  ScopedDevice sd(ctx->device);
  double *dv = nullptr;
  int32_t *dc = nullptr, *dr = nullptr;
  // +16 bytes of slack so 128-bit loads that start inside the array never leave the allocation
  cudaError_t e = cudaMalloc(&dv, sizeof(double) * (size_t)nnz + 64);
  if (e == cudaSuccess) e = cudaMalloc(&dc, sizeof(int32_t) * (size_t)nnz + 64);
  if (e == cudaSuccess) e = cudaMalloc(&dr, sizeof(int32_t) * (size_t)(n + 1) + 64);
  if (e == cudaSuccess && nnz) e = copy_h2d(ctx, dv, h_vals, sizeof(double) * (size_t)nnz);
  if (e == cudaSuccess && nnz) e = copy_h2d(ctx, dc, h_cols, sizeof(int32_t) * (size_t)nnz);
  if (e == cudaSuccess) e = copy_h2d(ctx, dr, h_rowptr, sizeof(int32_t) * (size_t)(n + 1));
  // the host arrays are only borrowed until this call returns (reference waits in setTarget, src/CG.hpp:169)
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    cudaFree(dv);
    cudaFree(dc);
    cudaFree(dr);
    return fail(CGB_ERR_CUDA, "cgb_csr_create: %s", cudaGetErrorString(e));
  }
  return cgb_csr_from_device(ctx, n, nnz, dv, dc, dr, 1, out);
}

int cgb_csr_destroy(cgb_csr *A) {
  if (!A) return CGB_OK;
  ScopedDevice sd(A->ctx->device);
  cudaStreamSynchronize(A->ctx->stream);
  workspace_release(A);
  csr_release_panels(A);
  csr_release_dict(A);
  cudaFree(A->sell_vals);
  cudaFree(A->sell_cols);
  cudaFree(A->sell_off);
  cudaFree(A->sell_width);
  cudaFree(A->sell_perm);
  cudaFree(A->sell_long);
  cudaFree(A->sell_chunk);
  if (A->owns) {
    cudaFree(A->vals);
    cudaFree(A->cols);
    cudaFree(A->rowptr);
  }
  delete A;
  return CGB_OK;
}

int64_t cgb_csr_n(const cgb_csr *A) { return A ? A->n : 0; }
int64_t cgb_csr_nnz(const cgb_csr *A) { return A ? A->nnz : 0; }
double *cgb_csr_vals(cgb_csr *A) { return A ? A->vals : nullptr; }
int32_t *cgb_csr_cols(cgb_csr *A) { return A ? A->cols : nullptr; }
int32_t *cgb_csr_rowptr(cgb_csr *A) { return A ? A->rowptr : nullptr; }

int cgb_csr_set_format(cgb_csr *A, int format) {
  CGB_REQUIRE(A, "cgb_csr_set_format: NULL matrix");
  CGB_REQUIRE(format == CGB_FMT_AUTO || format == CGB_FMT_CSR || format == CGB_FMT_SELL || format == CGB_FMT_DICT,
              "unknown format");
  if (format != A->format && A->ws.graph) { // cached graphs hold the other engine's kernels
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
  const int before = A->format;
  A->format = format;
  if (format == CGB_FMT_SELL || format == CGB_FMT_DICT) { // builds the SELL copy / column panels / packed copy now
    const int rc = csr_prepare(A);
    if (rc != CGB_OK) {
      A->format = before; // e.g. CGB_ERR_UNSUPPORTED: the matrix keeps streaming the way it did
      return rc;
    }
  }
  return CGB_OK;
}
int cgb_csr_get_format(const cgb_csr *A) { return A ? A->format : 0; }

int cgb_csr_set_panels(cgb_csr *A, int panels) {
  CGB_REQUIRE(A, "cgb_csr_set_panels: NULL matrix");
  CGB_REQUIRE(panels >= 0 && panels <= 64, "cgb_csr_set_panels: 0 (automatic) .. 64 panels");
  A->panels_req = panels;
  return CGB_OK;
}
int cgb_csr_get_panels(const cgb_csr *A) { return A ? (int)csr_panels(A) : 0; }

int cgb_spmv(cgb_ctx *ctx, cgb_csr *A, const double *d_x, double *d_y) {
  CGB_REQUIRE(ctx && A && d_x && d_y, "cgb_spmv: NULL argument");
  if (A->n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_TRY(csr_prepare(A));
  const SpmvLaunch lc = spmv_launch_cfg(A);
  launch_panel_partials(A, lc, ctx->stream, d_x, d_y, nullptr);
  dispatch_lanes(lc.lanes, lc.eng, [&](auto Lc, auto Tc) {
    launch_kernel(spmv_kernel<decltype(Lc)::value, decltype(Tc)::value>, lc, ctx->stream, csr_view_final(A, d_y), d_x, d_y,
                  (const CgState *)nullptr);
  });
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int cgb_residual(cgb_ctx *ctx, cgb_csr *A, const double *d_x, const double *d_b, double *rr, double *xx) {
  CGB_REQUIRE(ctx && A && d_x && d_b && rr && xx, "cgb_residual: NULL argument");
  ScopedDevice sd(ctx->device);
  CGB_TRY(ctx_scratch(ctx));
  if (A->n == 0) {
    *rr = 0.0;
    *xx = 0.0;
    return CGB_OK;
  }
  double *out2 = ctx->partials + 2 * kMaxPartials - 2; // tail of the scratch is never a partial slot
  CGB_TRY(csr_prepare(A));
  const SpmvLaunch lc = spmv_launch_cfg(A);
  double *acc = nullptr;
  if (csr_panels(A) > 1) { // the pass has no output vector of its own: the workspace's Ap is the accumulator
    if (!A->ws.Ap) CGB_TRY(workspace_ensure(A, A->n, false, 0));
    acc = A->ws.Ap;
    launch_panel_partials(A, lc, ctx->stream, d_x, acc, nullptr);
  }
  dispatch_lanes(lc.lanes, lc.eng, [&](auto Lc, auto Tc) {
    launch_kernel(residual_kernel<decltype(Lc)::value, decltype(Tc)::value>, lc, ctx->stream, csr_view_final(A, acc), d_x, d_b, out2,
                  ctx->partials, ctx->ticket);
  });
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaMemcpyAsync(ctx->pinned, out2, 2 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  *rr = ctx->pinned[0];
  *xx = ctx->pinned[1];
  return CGB_OK;
}

int cgb_gen_stencil(cgb_ctx *ctx, int kind, int64_t m, int64_t mz, int64_t row_begin, int64_t row_end, cgb_csr **out) {
  CGB_REQUIRE(ctx && out, "cgb_gen_stencil: NULL argument");
  CGB_REQUIRE(kind >= 0 && kind <= 2 && m >= 1, "cgb_gen_stencil: bad kind or size");
  if (kind == 0) mz = 1;
  if (mz <= 0) mz = m;
  const int64_t n = m * m * mz;
  if (row_end < 0) row_end = n;
  CGB_REQUIRE(row_begin >= 0 && row_begin <= row_end && row_end <= n, "cgb_gen_stencil: bad row range");
  CGB_REQUIRE(n < ((int64_t)1 << 31), "cgb_gen_stencil: n must fit 32-bit column indices");
  ScopedDevice sd(ctx->device);
  const int64_t nloc = row_end - row_begin;
  // row offsets are the reference's 32-bit ints (and the row lengths are scanned as such): bound the block's entries
  CGB_REQUIRE(nloc * (kind == 2 ? 27 : kind == 1 ? 7 : 5) < ((int64_t)1 << 31),
              "cgb_gen_stencil: this row block would hold 2^31 or more entries; generate it in smaller row ranges (one per GPU)");
  int32_t *rowptr = nullptr, *cols = nullptr;
  double *vals = nullptr;
  CGB_CUDA(cudaMalloc(&rowptr, sizeof(int32_t) * (size_t)(nloc + 1) + 64));
  CGB_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int32_t), ctx->stream));
  const int tb = 256;
  const unsigned gb = (unsigned)((nloc + tb - 1) / tb);
  int64_t nnz = 0;
  if (nloc > 0) {
    stencil_len_kernel<<<gb, tb, 0, ctx->stream>>>(kind, m, mz, row_begin, nloc, rowptr + 1);
    size_t tmp_bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, rowptr + 1, rowptr + 1, (int)nloc, ctx->stream);
    void *tmp = nullptr;
    CGB_CUDA(cudaMalloc(&tmp, tmp_bytes));
    cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, rowptr + 1, rowptr + 1, (int)nloc, ctx->stream);
    int last = 0;
    CGB_CUDA(cudaMemcpyAsync(&last, rowptr + nloc, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CGB_CUDA(cudaStreamSynchronize(ctx->stream));
    cudaFree(tmp);
    nnz = last;
  }
  CGB_CUDA(cudaMalloc(&cols, sizeof(int32_t) * (size_t)nnz + 64));
  CGB_CUDA(cudaMalloc(&vals, sizeof(double) * (size_t)nnz + 64));
  if (nloc > 0) {
    stencil_fill_kernel<<<gb, tb, 0, ctx->stream>>>(kind, m, mz, row_begin, nloc, rowptr, cols, vals);
    CGB_CUDA(cudaGetLastError());
  }
  int rc = cgb_csr_from_device(ctx, nloc, nnz, vals, cols, rowptr, 1, out);
  if (rc == CGB_OK) (*out)->n_cols = n;
  return rc;
}

} // extern "C"
