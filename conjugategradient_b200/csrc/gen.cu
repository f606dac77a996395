// gen.cu — on-device generator of the synthetic random SPD matrix with power-law row lengths
// (BASELINE.json config C5).  The reference ships no inputs at all; this builds the
// load-imbalance stress case directly in HBM (n = 10^7 would take minutes and tens of GB on the host).
//
// Construction (deterministic, counter-based splitmix64 keyed by (seed, row, slot)):
//   * strictly-lower degree of row i: deg_i = min(d_max, floor(d_min * u^(-1/(gamma-1)))), capped at i,
//     drawn by inverse transform through a table of P(deg >= d) built on the host (no device pow);
//   * its columns: one uniform pick inside each of deg_i equal strata of [0, i) -> distinct, ascending;
//   * values uniform in (-1, 0); every entry is mirrored; a_ii = 1 + sum_j |a_ij| (strictly diagonally
//     dominant => SPD, well conditioned).
// Assembly: (row << 32 | col) keys of lower + mirrored + diagonal entries, one 64-bit radix sort,
// row offsets from the sorted keys, diagonal filled by a sequential per-row sum.
#include "matrix.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <math.h>

#include <vector>

using namespace cgb;

namespace {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint64_t rnd(uint64_t seed, uint64_t row, uint64_t slot) {
  return splitmix64(splitmix64(seed ^ (row * 0xD1342543DE82EF95ull)) + slot);
}
// (0, 1]
__host__ __device__ __forceinline__ double unit_open0(uint64_t h) { return (double)((h >> 11) + 1) * (1.0 / 9007199254740992.0); }

// tail[d] = P(deg >= d), decreasing; deg = number of d in [1, dmax] with u <= tail[d]
__global__ void degree_kernel(int64_t n, uint64_t seed, const double *__restrict__ tail, int dmax, int *__restrict__ deg) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double u = unit_open0(rnd(seed, (uint64_t)i, 0));
  int lo = 0, hi = dmax; // invariant: u <= tail[lo] (tail[0] = 1), u > tail[hi + 1]
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (u <= tail[mid]) lo = mid; else hi = mid - 1;
  }
  deg[i] = (int)(lo < i ? lo : i);
}

__global__ void lower_fill_kernel(int64_t n, uint64_t seed, const int *__restrict__ deg, const int64_t *__restrict__ loff,
                                  int64_t nnz_lower, uint64_t *__restrict__ keys, double *__restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int d = deg[i];
  const int64_t base = 2 * loff[i];
  for (int s = 0; s < d; s++) {
    const int64_t a = (i * (int64_t)s) / d, b = (i * (int64_t)(s + 1)) / d; // stratum [a, b), b > a because d <= i
    const uint64_t h = rnd(seed, (uint64_t)i, (uint64_t)(2 * s + 1));
    const int64_t c = a + (int64_t)(h % (uint64_t)(b - a));
    const double v = -unit_open0(rnd(seed, (uint64_t)i, (uint64_t)(2 * s + 2)));
    keys[base + 2 * s] = ((uint64_t)i << 32) | (uint64_t)c;
    vals[base + 2 * s] = v;
    keys[base + 2 * s + 1] = ((uint64_t)c << 32) | (uint64_t)i;
    vals[base + 2 * s + 1] = v;
  }
  keys[2 * nnz_lower + i] = ((uint64_t)i << 32) | (uint64_t)i; // diagonal, value set later
  vals[2 * nnz_lower + i] = 0.0;
}

// every row has its diagonal, so each row id occurs: rowptr[row] = first entry of that row
__global__ void rowptr_kernel(int64_t total, const uint64_t *__restrict__ keys, int *__restrict__ rowptr, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int64_t r = (int64_t)(keys[e] >> 32);
  if (e == 0 || (int64_t)(keys[e - 1] >> 32) != r) rowptr[r] = (int)e;
  if (e == total - 1) rowptr[n] = (int)total;
}

__global__ void split_keys_kernel(int64_t total, const uint64_t *__restrict__ keys, int *__restrict__ cols) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < total) cols[e] = (int)(keys[e] & 0xffffffffull);
}

__global__ void diag_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols, double *__restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  int dpos = -1;
  for (int j = rowptr[i]; j < rowptr[i + 1]; j++) {
    if (cols[j] == (int)i) dpos = j; else s += fabs(vals[j]);
  }
  if (dpos >= 0) vals[dpos] = 1.0 + s;
}

__global__ void rebase_rowptr_kernel(int64_t nloc, const int *__restrict__ src, int *__restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= nloc) dst[i] = src[i] - src[0];
}

struct Buf {
  void *p = nullptr;
  ~Buf() { cudaFree(p); }
  template <class T> T *as() { return (T *)p; }
};

} // namespace

extern "C" int cgb_gen_random_spd(cgb_ctx *ctx, int64_t n, uint64_t seed, int d_min, int d_max, double gamma,
                                  int64_t row_begin, int64_t row_end, cgb_csr **out) {
  CGB_REQUIRE(ctx && out, "cgb_gen_random_spd: NULL argument");
  CGB_REQUIRE(n >= 1 && n < ((int64_t)1 << 31) && d_min >= 1 && d_max >= d_min && gamma > 1.0, "cgb_gen_random_spd: bad parameters");
  if (row_end < 0) row_end = n;
  CGB_REQUIRE(row_begin >= 0 && row_begin <= row_end && row_end <= n, "cgb_gen_random_spd: bad row range");
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  const int tb = 256;
  auto blocks = [&](int64_t m) { return (unsigned)((m + tb - 1) / tb); };

  // P(deg >= d) = P(d_min * u^(-1/(gamma-1)) >= d) = min(1, (d_min/d)^(gamma-1))
  std::vector<double> tail((size_t)d_max + 2, 0.0);
  tail[0] = 1.0;
  for (int d = 1; d <= d_max; d++) tail[d] = d <= d_min ? 1.0 : pow((double)d_min / (double)d, gamma - 1.0);
  Buf b_tail, b_deg, b_off, b_tmp, b_k0, b_k1, b_v0, b_v1;
  CGB_CUDA(cudaMalloc(&b_tail.p, tail.size() * sizeof(double)));
  CGB_CUDA(cudaMemcpyAsync(b_tail.p, tail.data(), tail.size() * sizeof(double), cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMalloc(&b_deg.p, sizeof(int) * (size_t)n));
  CGB_CUDA(cudaMalloc(&b_off.p, sizeof(int64_t) * (size_t)(n + 1)));
  degree_kernel<<<blocks(n), tb, 0, s>>>(n, seed, b_tail.as<double>(), d_max, b_deg.as<int>());
  CGB_CUDA(cudaGetLastError());
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, b_deg.as<int>(), b_off.as<int64_t>(), (int)n, s);
  CGB_CUDA(cudaMalloc(&b_tmp.p, tmp_bytes));
  cub::DeviceScan::ExclusiveSum(b_tmp.p, tmp_bytes, b_deg.as<int>(), b_off.as<int64_t>(), (int)n, s);
  int64_t last_off = 0;
  int last_deg = 0;
  CGB_CUDA(cudaMemcpyAsync(&last_off, b_off.as<int64_t>() + (n - 1), sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(&last_deg, b_deg.as<int>() + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  const int64_t nnz_lower = last_off + last_deg;
  const int64_t total = 2 * nnz_lower + n;
  CGB_REQUIRE(total < ((int64_t)1 << 31), "cgb_gen_random_spd: more than 2^31 non-zeros (32-bit row offsets)");

  CGB_CUDA(cudaMalloc(&b_k0.p, sizeof(uint64_t) * (size_t)total));
  CGB_CUDA(cudaMalloc(&b_k1.p, sizeof(uint64_t) * (size_t)total));
  CGB_CUDA(cudaMalloc(&b_v0.p, sizeof(double) * (size_t)total));
  double *vals_sorted = nullptr;
  CGB_CUDA(cudaMalloc(&vals_sorted, sizeof(double) * (size_t)total + 64));
  Buf g_vals;
  g_vals.p = vals_sorted; // released on any early return below
  lower_fill_kernel<<<blocks(n), tb, 0, s>>>(n, seed, b_deg.as<int>(), b_off.as<int64_t>(), nnz_lower, b_k0.as<uint64_t>(),
                                             b_v0.as<double>());
  CGB_CUDA(cudaGetLastError());
  int bits = 33;
  while (bits < 64 && ((int64_t)1 << (bits - 32)) < n) bits++;
  size_t sort_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, b_k0.as<uint64_t>(), b_k1.as<uint64_t>(), b_v0.as<double>(),
                                  vals_sorted, (int)total, 0, bits, s);
  Buf b_sort;
  CGB_CUDA(cudaMalloc(&b_sort.p, sort_bytes));
  cub::DeviceRadixSort::SortPairs(b_sort.p, sort_bytes, b_k0.as<uint64_t>(), b_k1.as<uint64_t>(), b_v0.as<double>(), vals_sorted,
                                  (int)total, 0, bits, s);
  int *rowptr = nullptr, *cols = nullptr;
  CGB_CUDA(cudaMalloc(&rowptr, sizeof(int) * (size_t)(n + 1) + 64));
  Buf g_rowptr;
  g_rowptr.p = rowptr;
  CGB_CUDA(cudaMalloc(&cols, sizeof(int) * (size_t)total + 64));
  Buf g_cols;
  g_cols.p = cols;
  rowptr_kernel<<<blocks(total), tb, 0, s>>>(total, b_k1.as<uint64_t>(), rowptr, n);
  split_keys_kernel<<<blocks(total), tb, 0, s>>>(total, b_k1.as<uint64_t>(), cols);
  diag_kernel<<<blocks(n), tb, 0, s>>>(n, rowptr, cols, vals_sorted);
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaStreamSynchronize(s));

  if (row_begin == 0 && row_end == n) {
    g_vals.p = g_rowptr.p = g_cols.p = nullptr; // ownership moves to the handle
    return cgb_csr_from_device(ctx, n, total, vals_sorted, cols, rowptr, 1, out);
  }
  // a row block with GLOBAL column indices (multi-GPU): slice the arrays
  const int64_t nloc = row_end - row_begin;
  int lo = 0, hi = 0;
  CGB_CUDA(cudaMemcpyAsync(&lo, rowptr + row_begin, sizeof(int), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(&hi, rowptr + row_end, sizeof(int), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  const int64_t nnz_loc = hi - lo;
  int *rp2 = nullptr, *c2 = nullptr;
  double *v2 = nullptr;
  CGB_CUDA(cudaMalloc(&rp2, sizeof(int) * (size_t)(nloc + 1) + 64));
  CGB_CUDA(cudaMalloc(&c2, sizeof(int) * (size_t)nnz_loc + 64));
  CGB_CUDA(cudaMalloc(&v2, sizeof(double) * (size_t)nnz_loc + 64));
  rebase_rowptr_kernel<<<blocks(nloc + 1), tb, 0, s>>>(nloc, rowptr + row_begin, rp2);
  CGB_CUDA(cudaMemcpyAsync(c2, cols + lo, sizeof(int) * (size_t)nnz_loc, cudaMemcpyDeviceToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(v2, vals_sorted + lo, sizeof(double) * (size_t)nnz_loc, cudaMemcpyDeviceToDevice, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  int rc = cgb_csr_from_device(ctx, nloc, nnz_loc, v2, c2, rp2, 1, out);
  if (rc == CGB_OK) (*out)->n_cols = n;
  return rc;
}
