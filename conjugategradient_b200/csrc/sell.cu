// sell.cu — device-side CSR -> SELL-32-sigma conversion (the shadow copy the kEngSell row engine streams).
//
// The reference has one storage format, CSR (src/LinearAlgebraTypes.hpp:57-132), and a thread-per-row
// SpMV (src/VectorOperations.hpp:455-462) whose loads are uncoalesced and whose warps wait for their
// longest row.  SELL-C-sigma (C = 32 = one warp, sigma = sorting window) is the layout that fixes both
// for irregular row lengths (BASELINE config C5): see spmv.cuh for the kernel side.
//
//   1. slot order: rows sorted by DESCENDING length inside windows of sigma rows (stable radix sort of
//      (window, ~length) keys); sigma = 1 keeps the natural order and needs no permutation array;
//   2. slice width = longest row of its 32 slots; slice offsets = exclusive scan of 32 * width (64-bit);
//   3. fill: entry j of slot l of slice s at off[s] + 32 j + l; padding = (0.0, column -1).
//   4. rows longer than kSellLongRow (head of a power-law distribution) are kept OUT of the slices (they
//      would make their whole slice as wide as they are and serialise one lane for thousands of steps):
//      their slot is marked empty (perm[slot] = ~row) and they go to a short list that the engine
//      processes a WARP per row straight from the CSR arrays (coalesced, shuffle-reduced).
// The permutation stays inside the library: x, b, r, p keep the caller's row order; only the order
// in which rows are processed (and the row written by the epilogue) is permuted.
#include "matrix.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <stdlib.h>

using namespace cgb;

namespace {

__device__ __forceinline__ int sell_len(const int *__restrict__ rowptr, int64_t r, int long_row) {
  const int len = rowptr[r + 1] - rowptr[r];
  return len > long_row ? 0 : len; // long rows do not live in the slices
}

__global__ void sell_keys_kernel(int64_t n, const int *__restrict__ rowptr, int sigma, int long_row,
                                 uint64_t *__restrict__ keys, int *__restrict__ rows) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t len = (uint32_t)sell_len(rowptr, i, long_row);
  keys[i] = ((uint64_t)(i / sigma) << 32) | (uint64_t)(0xffffffffu - len);
  rows[i] = (int)i;
}

__global__ void sell_identity_kernel(int64_t n, int *__restrict__ perm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) perm[i] = (int)i;
}

// perm[slot] = ~row for the rows that live in the long-row list
__global__ void sell_mark_long_kernel(int64_t n, const int *__restrict__ rowptr, int long_row, int *__restrict__ perm) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int r = perm[t];
  if (rowptr[r + 1] - rowptr[r] > long_row) perm[t] = ~r;
}

struct IsLongRow {
  const int *rowptr;
  int long_row;
  __host__ __device__ bool operator()(const int &r) const { return rowptr[r + 1] - rowptr[r] > long_row; }
};

// a warp per slice: width = max row length over its slots
__global__ void sell_width_kernel(int64_t n, int64_t slices, const int *__restrict__ rowptr, const int *__restrict__ perm,
                                  int long_row, int64_t *__restrict__ entries, int *__restrict__ widths) {
  const int lane = threadIdx.x & 31;
  const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (s >= slices) return;
  const int64_t slot = s * 32 + lane;
  int len = 0;
  if (slot < n) {
    const int64_t r = perm ? perm[slot] : slot;
    len = r >= 0 ? sell_len(rowptr, r, long_row) : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) len = max(len, __shfl_xor_sync(0xffffffffu, len, o));
  if (lane == 0) {
    entries[s] = (int64_t)len * 32;
    widths[s] = len;
  }
}

__global__ void sell_fill_kernel(int64_t n, int64_t slices, const int *__restrict__ rowptr, const int *__restrict__ cols,
                                 const double *__restrict__ vals, const int *__restrict__ perm,
                                 const int64_t *__restrict__ off, int *__restrict__ scols, double *__restrict__ svals) {
  const int lane = threadIdx.x & 31;
  const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (s >= slices) return;
  const int64_t slot = s * 32 + lane;
  int b = 0, len = 0;
  if (slot < n) {
    const int64_t r = perm ? perm[slot] : slot;
    if (r >= 0) { // negative: a long row, served from the CSR arrays
      b = rowptr[r];
      len = rowptr[r + 1] - b;
    }
  }
  const int64_t o0 = off[s];
  const int width = (int)((off[s + 1] - o0) >> 5);
  for (int j = 0; j < width; j++) {
    const int64_t d = o0 + (int64_t)j * 32 + lane;
    if (j < len) {
      scols[d] = cols[b + j];
      svals[d] = vals[b + j];
    } else {
      scols[d] = -1;
      svals[d] = 0.0;
    }
  }
}

// chunk[b] = first slice of warp b: contiguous slice ranges of (nearly) equal COST, so that the static assignment
// is balanced whatever the row-length distribution and sorting window are.  cost(slice) = its column-major steps
// + kSliceCost for the per-slice work (slot -> row lookup, epilogue), which dominates when rows are short.
constexpr int64_t kSliceCost = 2;
__global__ void sell_chunk_kernel(int64_t slices, const int64_t *__restrict__ off, int nb, int64_t *__restrict__ chunk) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nb) return;
  const int64_t total = (off[slices] >> 5) + kSliceCost * slices;
  const int64_t target = (int64_t)(((__int128)total * b) / nb);
  int64_t lo = 0, hi = slices; // first slice whose cost prefix is >= target
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if ((off[mid] >> 5) + kSliceCost * mid < target) lo = mid + 1; else hi = mid;
  }
  chunk[b] = b == nb ? slices : lo;
}

struct Buf {
  void *p = nullptr;
  ~Buf() { cudaFree(p); }
  template <class T> T *as() { return (T *)p; }
};

} // namespace

extern "C" int cgb_csr_build_sell(cgb_csr *A, int sigma) {
  CGB_REQUIRE(A, "cgb_csr_build_sell: NULL matrix");
  CGB_REQUIRE(sigma >= 1, "cgb_csr_build_sell: sigma must be >= 1");
  if (sigma > 1) sigma = ((sigma + 31) / 32) * 32; // windows are whole slices
  if (A->sell_vals && A->sell_sigma == sigma) return CGB_OK;
  csr_release_panels(A); // column panels (if any) are rebuilt with the new window by the next csr_prepare
  int long_row = kSellLongRow;
  if (const char *e = getenv("CGB_SELL_LONG")) long_row = atoi(e) > 0 ? atoi(e) : long_row;
  cgb_ctx *ctx = A->ctx;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  CGB_CUDA(cudaStreamSynchronize(s));
  cudaFree(A->sell_vals);
  cudaFree(A->sell_cols);
  cudaFree(A->sell_off);
  cudaFree(A->sell_width);
  A->sell_width = nullptr;
  cudaFree(A->sell_perm);
  cudaFree(A->sell_long);
  cudaFree(A->sell_chunk);
  A->sell_chunk = nullptr;
  A->sell_long = nullptr;
  A->sell_nlong = 0;
  A->sell_vals = nullptr;
  A->sell_cols = nullptr;
  A->sell_off = nullptr;
  A->sell_perm = nullptr;
  A->sell_sigma = 0;
  if (A->ws.graph) { // cached graphs hold the old engine
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
  const int64_t n = A->n;
  if (n == 0) return CGB_OK;
  const int64_t slices = (n + 31) / 32;
  const int tb = 256;
  auto blocks = [&](int64_t m) { return (unsigned)((m + tb - 1) / tb); };

  // the long-row list (usually empty or tiny)
  {
    Buf ids, cnt, tmp;
    CGB_CUDA(cudaMalloc(&ids.p, sizeof(int) * (size_t)n));
    CGB_CUDA(cudaMalloc(&cnt.p, sizeof(int)));
    CGB_CUDA(cudaMalloc(&A->sell_long, sizeof(int) * (size_t)n));
    sell_identity_kernel<<<blocks(n), tb, 0, s>>>(n, ids.as<int>());
    size_t bytes = 0;
    cub::DeviceSelect::If(nullptr, bytes, ids.as<int>(), A->sell_long, cnt.as<int>(), (int)n, IsLongRow{A->rowptr, long_row}, s);
    CGB_CUDA(cudaMalloc(&tmp.p, bytes));
    cub::DeviceSelect::If(tmp.p, bytes, ids.as<int>(), A->sell_long, cnt.as<int>(), (int)n, IsLongRow{A->rowptr, long_row}, s);
    int nl = 0;
    CGB_CUDA(cudaMemcpyAsync(&nl, cnt.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    A->sell_nlong = nl;
    if (nl == 0) {
      cudaFree(A->sell_long);
      A->sell_long = nullptr;
    }
  }
  if (sigma > 1) {
    Buf k0, k1, r0, tmp;
    CGB_CUDA(cudaMalloc(&k0.p, sizeof(uint64_t) * (size_t)n));
    CGB_CUDA(cudaMalloc(&k1.p, sizeof(uint64_t) * (size_t)n));
    CGB_CUDA(cudaMalloc(&r0.p, sizeof(int) * (size_t)n));
    CGB_CUDA(cudaMalloc(&A->sell_perm, sizeof(int) * (size_t)n));
    sell_keys_kernel<<<blocks(n), tb, 0, s>>>(n, A->rowptr, sigma, long_row, k0.as<uint64_t>(), r0.as<int>());
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k0.as<uint64_t>(), k1.as<uint64_t>(), r0.as<int>(), A->sell_perm, (int)n, 0,
                                    64, s);
    CGB_CUDA(cudaMalloc(&tmp.p, bytes));
    cub::DeviceRadixSort::SortPairs(tmp.p, bytes, k0.as<uint64_t>(), k1.as<uint64_t>(), r0.as<int>(), A->sell_perm, (int)n, 0, 64,
                                    s);
    CGB_CUDA(cudaGetLastError());
    CGB_CUDA(cudaStreamSynchronize(s));
  }
  if (A->sell_nlong > 0) {
    if (!A->sell_perm) {
      CGB_CUDA(cudaMalloc(&A->sell_perm, sizeof(int) * (size_t)n));
      sell_identity_kernel<<<blocks(n), tb, 0, s>>>(n, A->sell_perm);
    }
    sell_mark_long_kernel<<<blocks(n), tb, 0, s>>>(n, A->rowptr, long_row, A->sell_perm);
  }
  CGB_CUDA(cudaMalloc(&A->sell_off, sizeof(int64_t) * (size_t)(slices + 1)));
  CGB_CUDA(cudaMalloc(&A->sell_width, sizeof(int) * (size_t)slices + 256));
  sell_width_kernel<<<blocks(slices * 32), tb, 0, s>>>(n, slices, A->rowptr, A->sell_perm, long_row, A->sell_off, A->sell_width);
  CGB_CUDA(cudaMemsetAsync(A->sell_off + slices, 0, sizeof(int64_t), s));
  {
    Buf tmp;
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, A->sell_off, A->sell_off, (int)(slices + 1), s);
    CGB_CUDA(cudaMalloc(&tmp.p, bytes));
    cub::DeviceScan::ExclusiveSum(tmp.p, bytes, A->sell_off, A->sell_off, (int)(slices + 1), s);
    CGB_CUDA(cudaGetLastError());
    CGB_CUDA(cudaStreamSynchronize(s));
  }
  int64_t padded = 0;
  CGB_CUDA(cudaMemcpy(&padded, A->sell_off + slices, sizeof(int64_t), cudaMemcpyDeviceToHost));
  CGB_CUDA(cudaMalloc(&A->sell_cols, sizeof(int) * (size_t)padded + 64));
  CGB_CUDA(cudaMalloc(&A->sell_vals, sizeof(double) * (size_t)padded + 64));
  sell_fill_kernel<<<blocks(slices * 32), tb, 0, s>>>(n, slices, A->rowptr, A->cols, A->vals, A->sell_perm, A->sell_off,
                                                      A->sell_cols, A->sell_vals);
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaStreamSynchronize(s));
  // one contiguous, cost-balanced slice range per WARP (kBlock / 32 warps per CTA)
  const int units = A->grid_sell * (kBlock / 32);
  CGB_CUDA(cudaMalloc(&A->sell_chunk, sizeof(int64_t) * (size_t)(units + 1)));
  sell_chunk_kernel<<<blocks(units + 1), tb, 0, s>>>(slices, A->sell_off, units, A->sell_chunk);
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaStreamSynchronize(s));
  A->sell_slices = slices;
  A->sell_padded = padded;
  A->sell_sigma = sigma;
  return CGB_OK;
}

extern "C" int64_t cgb_csr_sell_padded(const cgb_csr *A) {
  if (!A) return 0;
  if (A->panels.empty()) return A->sell_padded;
  int64_t t = 0;
  for (const cgb_csr *p : A->panels) t += p->sell_padded;
  return t;
}

// ---- column panels ------------------------------------------------------------------------------
namespace {

// entries of row t with column in [c_lo, c_hi): counted into cnt[t + 1] (cnt[0] = 0 is set by the caller)
__global__ void panel_count_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols, int c_lo, int c_hi,
                                   int *__restrict__ cnt) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  int c = 0;
  for (int j = rowptr[t], e = rowptr[t + 1]; j < e; j++) {
    const int col = cols[j];
    c += (col >= c_lo && col < c_hi) ? 1 : 0;
  }
  cnt[t + 1] = c;
}

// copies them in their original order (the reference's left-to-right row sum is kept inside a panel)
__global__ void panel_fill_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols,
                                  const double *__restrict__ vals, int c_lo, int c_hi, const int *__restrict__ prow,
                                  int *__restrict__ pcols, double *__restrict__ pvals) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  int o = prow[t];
  for (int j = rowptr[t], e = rowptr[t + 1]; j < e; j++) {
    const int col = cols[j];
    if (col >= c_lo && col < c_hi) {
      pcols[o] = col;
      pvals[o] = vals[j];
      o++;
    }
  }
}

} // namespace

namespace cgb {

void csr_release_panels(cgb_csr *A) {
  if (A->panels.empty()) return;
  for (cgb_csr *p : A->panels) cgb_csr_destroy(p);
  A->panels.clear();
  if (A->ws.graph) {
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
}

// A = [A_0 | ... | A_{P-1}], panel k = columns [n_cols k / P, n_cols (k+1) / P): each a CSR sub-matrix over all
// rows (owned by its own handle) with its SELL-32-sigma copy; the handles share A's grid so one launch
// configuration serves every panel.
int csr_build_panels(cgb_csr *A, int P, int sigma) {
  csr_release_panels(A);
  if (P < 2 || A->n == 0) return CGB_OK;
  cgb_ctx *ctx = A->ctx;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  const int64_t n = A->n;
  const int tb = 256;
  const unsigned gb = (unsigned)((n + tb - 1) / tb);
  for (int k = 0; k < P; k++) {
    const int c_lo = (int)((A->n_cols * k) / P), c_hi = (int)((A->n_cols * (int64_t)(k + 1)) / P);
    int *prow = nullptr, *pcols = nullptr;
    double *pvals = nullptr;
    Buf g_row, g_cols, g_vals, tmp;
    CGB_CUDA(cudaMalloc(&prow, sizeof(int) * (size_t)(n + 1) + 64));
    g_row.p = prow;
    CGB_CUDA(cudaMemsetAsync(prow, 0, sizeof(int), s));
    panel_count_kernel<<<gb, tb, 0, s>>>(n, A->rowptr, A->cols, c_lo, c_hi, prow);
    size_t bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, bytes, prow + 1, prow + 1, (int)n, s);
    CGB_CUDA(cudaMalloc(&tmp.p, bytes));
    cub::DeviceScan::InclusiveSum(tmp.p, bytes, prow + 1, prow + 1, (int)n, s);
    int nnz_k = 0;
    CGB_CUDA(cudaMemcpyAsync(&nnz_k, prow + n, sizeof(int), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    CGB_CUDA(cudaMalloc(&pcols, sizeof(int) * (size_t)nnz_k + 64));
    g_cols.p = pcols;
    CGB_CUDA(cudaMalloc(&pvals, sizeof(double) * (size_t)nnz_k + 64));
    g_vals.p = pvals;
    panel_fill_kernel<<<gb, tb, 0, s>>>(n, A->rowptr, A->cols, A->vals, c_lo, c_hi, prow, pcols, pvals);
    CGB_CUDA(cudaGetLastError());
    CGB_CUDA(cudaStreamSynchronize(s));
    cgb_csr *S = nullptr;
    CGB_TRY(cgb_csr_from_device(ctx, n, nnz_k, pvals, pcols, prow, 1, &S));
    g_row.p = g_cols.p = g_vals.p = nullptr; // owned by the panel's handle now
    S->n_cols = A->n_cols;
    S->panels_ok = false;
    S->format = CGB_FMT_SELL;
    S->stats_done = true; // never planned on its own: always streamed through the SELL engine
    S->grid_sell = A->grid_sell;
    A->panels.push_back(S);
    CGB_TRY(cgb_csr_build_sell(S, sigma));
  }
  if (A->ws.graph) {
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
  return CGB_OK;
}

} // namespace cgb
