// coo.cu — triplets (COO) -> device CSR, entirely on the GPU: mirror, sort, merge duplicates, row offsets.
//
// What it replaces.  The reference builds its CSR on the host: read_real_coordinate_matrix pushes 16-byte tuples
// into a std::vector, mirrors every off-diagonal entry, std::sorts them and walks the result
// (test/mm_reader.cpp:45-107); its other caller pulls a matrix out of HiFlow3 one element at a time
// (src/CG.hpp:104-151).  Both are the step right before CG::setMatrix.  Here the
// triplets are uploaded once and everything else is radix sort + scans in HBM; the result is an ordinary cgb_csr
// (rows ascending, columns ascending inside a row), so the CG path behind it is unchanged.
//
// Flags reproduce the reference reader's dialect when asked to (CGB_COO_REFERENCE_DIALECT): every off-diagonal
// is mirrored whatever the banner said (:68-74), indices are 1-based (:62-66), and a row offset is emitted per
// DISTINCT row id, so empty rows vanish (:92-104).
#include "matrix.cuh"

#include <chrono>
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

using namespace cgb;

namespace {

struct Buf {
  void *p = nullptr;
  ~Buf() { cudaFree(p); }
  template <class T> T *as() { return (T *)p; }
};

// range check + off-diagonal flags.  stat[0] = min index, stat[1] = max index (after the base shift)
__global__ void coo_scan_kernel(int64_t h, const int *__restrict__ r, const int *__restrict__ c, int base, int mirror,
                                int *__restrict__ stat, int *__restrict__ offdiag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int lo = 0x7fffffff, hi = -0x7fffffff - 1;
  if (i < h) {
    const int a = r[i] - base, b = c[i] - base;
    lo = min(a, b);
    hi = max(a, b);
    if (offdiag) offdiag[i] = (mirror && a != b) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(stat + 0, lo);
    atomicMax(stat + 1, hi);
  }
}

// key = (row << 32) | column; the mirror of entry i goes to slot h + moff[i]
__global__ void coo_keys_kernel(int64_t h, const int *__restrict__ r, const int *__restrict__ c,
                                const double *__restrict__ v, int base, const int *__restrict__ offdiag,
                                const int *__restrict__ moff, uint64_t *__restrict__ keys, double *__restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= h) return;
  const uint32_t a = (uint32_t)(r[i] - base), b = (uint32_t)(c[i] - base);
  const double x = v[i];
  keys[i] = ((uint64_t)a << 32) | b;
  vals[i] = x;
  if (offdiag && offdiag[i]) {
    const int64_t m = h + moff[i];
    keys[m] = ((uint64_t)b << 32) | a;
    vals[m] = x;
  }
}

__global__ void coo_head_kernel(int64_t t, const uint64_t *__restrict__ keys, int shift, int *__restrict__ head) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= t) return;
  head[i] = (i == 0 || (keys[i] >> shift) != (keys[i - 1] >> shift)) ? 1 : 0;
}

// duplicates: the first entry of a run of equal (row, column) keys sums the run left to right (the stable sort
// keeps the input order inside a run, so the sum is reproducible) and writes the compacted entry
__global__ void coo_merge_kernel(int64_t t, const uint64_t *__restrict__ keys, const double *__restrict__ vals,
                                 const int *__restrict__ head, const int *__restrict__ uidx, uint64_t *__restrict__ okeys,
                                 double *__restrict__ ovals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= t || !head[i]) return;
  double s = vals[i];
  for (int64_t j = i + 1; j < t && !head[j]; j++) s += vals[j];
  okeys[uidx[i]] = keys[i];
  ovals[uidx[i]] = s;
}

__global__ void coo_cols_kernel(int64_t t, const uint64_t *__restrict__ keys, int *__restrict__ cols) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < t) cols[i] = (int)(keys[i] & 0xffffffffu);
}

// full row range [0, n): rowptr[q] = first entry whose row is >= q
__global__ void coo_rowptr_kernel(int64_t t, int64_t n, const uint64_t *__restrict__ keys, int *__restrict__ rowptr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > t) return;
  const int64_t prev = i == 0 ? -1 : (int64_t)(keys[i - 1] >> 32);
  const int64_t cur = i == t ? n : (int64_t)(keys[i] >> 32);
  for (int64_t q = prev + 1; q <= cur; q++) rowptr[q] = (int)i;
}

int coo_to_csr_device(cgb_ctx *ctx, int64_t n_rows, int64_t h, const int *d_r, const int *d_c, const double *d_v, int flags,
                      cgb_csr **out) {
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  const int tb = 256;
  auto blocks = [&](int64_t m) { return (unsigned)((m + tb - 1) / tb); };
  const int base = (flags & CGB_COO_ONE_BASED) ? 1 : 0;
  const bool mirror = (flags & CGB_COO_MIRROR) != 0;
  const bool compact = (flags & CGB_COO_COMPACT_ROWS) != 0;
  CGB_REQUIRE(h > 0, "triplets: no entries");
  CGB_REQUIRE(h < ((int64_t)1 << 31), "triplets: too many entries for 32-bit offsets");

  Buf stat, offd, moff, tmp;
  CGB_CUDA(cudaMalloc(&stat.p, 2 * sizeof(int)));
  const int init[2] = {0x7fffffff, -0x7fffffff - 1};
  CGB_CUDA(cudaMemcpyAsync(stat.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
  if (mirror) {
    CGB_CUDA(cudaMalloc(&offd.p, sizeof(int) * (size_t)h));
    CGB_CUDA(cudaMalloc(&moff.p, sizeof(int) * (size_t)h));
  }
  coo_scan_kernel<<<blocks(h), tb, 0, s>>>(h, d_r, d_c, base, mirror ? 1 : 0, stat.as<int>(), offd.as<int>());
  CGB_CUDA(cudaGetLastError());
  int64_t mirrored = 0;
  if (mirror) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, offd.as<int>(), moff.as<int>(), (int)h, s);
    CGB_CUDA(cudaMalloc(&tmp.p, bytes));
    cub::DeviceScan::ExclusiveSum(tmp.p, bytes, offd.as<int>(), moff.as<int>(), (int)h, s);
    int last_off = 0, last_flag = 0;
    CGB_CUDA(cudaMemcpyAsync(&last_off, moff.as<int>() + (h - 1), sizeof(int), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaMemcpyAsync(&last_flag, offd.as<int>() + (h - 1), sizeof(int), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    mirrored = (int64_t)last_off + last_flag;
  }
  int hstat[2];
  CGB_CUDA(cudaMemcpyAsync(hstat, stat.p, sizeof(hstat), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  if (hstat[0] < 0) return fail(CGB_ERR_INVALID, "triplets: index below %d", base);
  if (n_rows >= 0 && hstat[1] >= n_rows) return fail(CGB_ERR_INVALID, "triplets: index %d outside the %lld x %lld matrix", hstat[1] + base, (long long)n_rows, (long long)n_rows);
  const int64_t span = (int64_t)hstat[1] + 1; // rows (and columns) the entries reach
  int64_t total = h + mirrored;
  CGB_REQUIRE(total < ((int64_t)1 << 31), "triplets: more than 2^31 entries after mirroring (32-bit row offsets)");

  Buf k0, k1, v0, v1, sorttmp;
  CGB_CUDA(cudaMalloc(&k0.p, sizeof(uint64_t) * (size_t)total));
  CGB_CUDA(cudaMalloc(&k1.p, sizeof(uint64_t) * (size_t)total));
  CGB_CUDA(cudaMalloc(&v0.p, sizeof(double) * (size_t)total));
  CGB_CUDA(cudaMalloc(&v1.p, sizeof(double) * (size_t)total + 64));
  coo_keys_kernel<<<blocks(h), tb, 0, s>>>(h, d_r, d_c, d_v, base, offd.as<int>(), moff.as<int>(), k0.as<uint64_t>(), v0.as<double>());
  CGB_CUDA(cudaGetLastError());
  int bits = 33;
  while (bits < 64 && ((int64_t)1 << (bits - 32)) < span) bits++;
  {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k0.as<uint64_t>(), k1.as<uint64_t>(), v0.as<double>(), v1.as<double>(), (int)total, 0, bits, s);
    CGB_CUDA(cudaMalloc(&sorttmp.p, bytes));
    cub::DeviceRadixSort::SortPairs(sorttmp.p, bytes, k0.as<uint64_t>(), k1.as<uint64_t>(), v0.as<double>(), v1.as<double>(), (int)total, 0, bits, s);
    CGB_CUDA(cudaGetLastError());
  }
  uint64_t *keys = k1.as<uint64_t>();
  double *vals = v1.as<double>();
  Buf head, uidx, scantmp;
  CGB_CUDA(cudaMalloc(&head.p, sizeof(int) * (size_t)(total + 1)));
  CGB_CUDA(cudaMalloc(&uidx.p, sizeof(int) * (size_t)(total + 1)));
  auto scan_heads = [&](int64_t count) -> int { // exclusive scan of head[0..count] (head[count] = 0) -> uidx; returns the sum
    cudaMemsetAsync(head.as<int>() + count, 0, sizeof(int), s);
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, head.as<int>(), uidx.as<int>(), (int)(count + 1), s);
    if (!scantmp.p) cudaMalloc(&scantmp.p, bytes + 1024);
    cub::DeviceScan::ExclusiveSum(scantmp.p, bytes, head.as<int>(), uidx.as<int>(), (int)(count + 1), s);
    int sum = 0;
    cudaMemcpyAsync(&sum, uidx.as<int>() + count, sizeof(int), cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
    return sum;
  };
  if (flags & CGB_COO_SUM_DUPLICATES) {
    coo_head_kernel<<<blocks(total), tb, 0, s>>>(total, keys, 0, head.as<int>());
    const int uniq = scan_heads(total);
    if (uniq < total) {
      coo_merge_kernel<<<blocks(total), tb, 0, s>>>(total, keys, vals, head.as<int>(), uidx.as<int>(), k0.as<uint64_t>(), v0.as<double>());
      CGB_CUDA(cudaGetLastError());
      keys = k0.as<uint64_t>();
      vals = v0.as<double>();
      total = uniq;
    }
  }
  // CSR arrays
  int *cols = nullptr, *rowptr = nullptr;
  double *ovals = nullptr;
  Buf g_cols, g_row, g_vals;
  CGB_CUDA(cudaMalloc(&cols, sizeof(int) * (size_t)total + 64));
  g_cols.p = cols;
  CGB_CUDA(cudaMalloc(&ovals, sizeof(double) * (size_t)total + 64));
  g_vals.p = ovals;
  coo_cols_kernel<<<blocks(total), tb, 0, s>>>(total, keys, cols);
  CGB_CUDA(cudaMemcpyAsync(ovals, vals, sizeof(double) * (size_t)total, cudaMemcpyDeviceToDevice, s));
  int64_t n = 0;
  if (compact) { // an offset per distinct row id (reference: test/mm_reader.cpp:92-104)
    coo_head_kernel<<<blocks(total), tb, 0, s>>>(total, keys, 32, head.as<int>());
    n = scan_heads(total);
    CGB_CUDA(cudaMalloc(&rowptr, sizeof(int) * (size_t)(n + 1) + 64));
    g_row.p = rowptr;
    size_t bytes = 0;
    Buf seltmp, cnt;
    CGB_CUDA(cudaMalloc(&cnt.p, sizeof(int)));
    thrust::counting_iterator<int> idx(0);
    cub::DeviceSelect::Flagged(nullptr, bytes, idx, head.as<int>(), rowptr, cnt.as<int>(), (int)total, s);
    CGB_CUDA(cudaMalloc(&seltmp.p, bytes));
    cub::DeviceSelect::Flagged(seltmp.p, bytes, idx, head.as<int>(), rowptr, cnt.as<int>(), (int)total, s);
    const int tot32 = (int)total;
    CGB_CUDA(cudaMemcpyAsync(rowptr + n, &tot32, sizeof(int), cudaMemcpyHostToDevice, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    // Column ids are left alone (test/mm_reader.cpp:92-104): with an empty row they reach past the compacted dimension
    // and the result is a rectangular n x span matrix (n_cols below) -- fine for cgb_spmv with a vector of that length,
    // refused by the solver and by cgb_csr_from_mm_file.
  } else {
    n = n_rows >= 0 ? n_rows : span;
    CGB_CUDA(cudaMalloc(&rowptr, sizeof(int) * (size_t)(n + 1) + 64));
    g_row.p = rowptr;
    coo_rowptr_kernel<<<blocks(total + 1), tb, 0, s>>>(total, n, keys, rowptr);
  }
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaStreamSynchronize(s));
  CGB_TRY(cgb_csr_from_device(ctx, n, total, ovals, cols, rowptr, 1, out));
  g_cols.p = g_row.p = g_vals.p = nullptr; // owned by the handle now
  if (span > n && (compact || n_rows < 0)) (*out)->n_cols = span;
  return CGB_OK;
}

} // namespace

namespace cgb {
// mm_reader.cpp: the reference dialect's triplets (1-based) of a Matrix Market file, parsed on all host cores
int mm_parse_chunks(const char *path, std::vector<std::vector<int32_t>> &pr, std::vector<std::vector<int32_t>> &pc,
                    std::vector<std::vector<double>> &pv);
} // namespace cgb

extern "C" {

int cgb_csr_from_triplets_device(cgb_ctx *ctx, int64_t n_rows, int64_t count, const int32_t *d_rows, const int32_t *d_cols,
                                 const double *d_vals, int flags, cgb_csr **out) {
  CGB_REQUIRE(ctx && out && d_rows && d_cols && d_vals, "cgb_csr_from_triplets_device: NULL argument");
  CGB_REQUIRE(n_rows < ((int64_t)1 << 31), "cgb_csr_from_triplets: n must fit the reference's 32-bit indices");
  CGB_REQUIRE(!((flags & CGB_COO_COMPACT_ROWS) && n_rows >= 0), "cgb_csr_from_triplets: CGB_COO_COMPACT_ROWS derives the row count itself (pass n_rows = -1)");
  *out = nullptr;
  return coo_to_csr_device(ctx, n_rows, count, d_rows, d_cols, d_vals, flags, out);
}

int cgb_csr_from_triplets(cgb_ctx *ctx, int64_t n_rows, int64_t count, const int32_t *h_rows, const int32_t *h_cols,
                          const double *h_vals, int flags, cgb_csr **out) {
  CGB_REQUIRE(ctx && out && h_rows && h_cols && h_vals, "cgb_csr_from_triplets: NULL argument");
  CGB_REQUIRE(count > 0 && count < ((int64_t)1 << 31), "cgb_csr_from_triplets: entry count must be in [1, 2^31)");
  *out = nullptr;
  ScopedDevice sd(ctx->device);
  Buf r, c, v;
  CGB_CUDA(cudaMalloc(&r.p, sizeof(int) * (size_t)count));
  CGB_CUDA(cudaMalloc(&c.p, sizeof(int) * (size_t)count));
  CGB_CUDA(cudaMalloc(&v.p, sizeof(double) * (size_t)count));
  CGB_CUDA(cudaMemcpyAsync(r.p, h_rows, sizeof(int) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  CGB_CUDA(cudaMemcpyAsync(c.p, h_cols, sizeof(int) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  CGB_CUDA(cudaMemcpyAsync(v.p, h_vals, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  return cgb_csr_from_triplets_device(ctx, n_rows, count, r.as<int>(), c.as<int>(), v.as<double>(), flags, out);
}

int cgb_csr_from_mm_file(cgb_ctx *ctx, const char *path, cgb_csr **out) {
  CGB_REQUIRE(ctx && path && out, "cgb_csr_from_mm_file: NULL argument");
  *out = nullptr;
  std::vector<std::vector<int32_t>> pr, pc;
  std::vector<std::vector<double>> pv;
  const bool timing = getenv("CGB_INGEST_TIMING") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  CGB_TRY(mm_parse_chunks(path, pr, pc, pv));
  const auto t1 = std::chrono::steady_clock::now();
  size_t total = 0;
  for (auto &x : pr) total += x.size();
  if (total == 0) return fail(CGB_ERR_IO, "cgb_csr_from_mm_file: no entries in '%s'", path);
  CGB_REQUIRE(total < ((size_t)1 << 31), "cgb_csr_from_mm_file: too many entries for 32-bit offsets");
  // the parser's per-thread chunks go up as they are (no concatenation on the host)
  ScopedDevice sd(ctx->device);
  Buf r, c, v;
  CGB_CUDA(cudaMalloc(&r.p, sizeof(int) * total));
  CGB_CUDA(cudaMalloc(&c.p, sizeof(int) * total));
  CGB_CUDA(cudaMalloc(&v.p, sizeof(double) * total));
  size_t o = 0;
  for (size_t k = 0; k < pr.size(); k++) {
    const size_t m = pr[k].size();
    if (m == 0) continue;
    CGB_CUDA(cudaMemcpyAsync(r.as<int>() + o, pr[k].data(), sizeof(int) * m, cudaMemcpyHostToDevice, ctx->stream));
    CGB_CUDA(cudaMemcpyAsync(c.as<int>() + o, pc[k].data(), sizeof(int) * m, cudaMemcpyHostToDevice, ctx->stream));
    CGB_CUDA(cudaMemcpyAsync(v.as<double>() + o, pv[k].data(), sizeof(double) * m, cudaMemcpyHostToDevice, ctx->stream));
    o += m;
  }
  const int rc = cgb_csr_from_triplets_device(ctx, -1, (int64_t)total, r.as<int>(), c.as<int>(), v.as<double>(),
                                              CGB_COO_REFERENCE_DIALECT, out);
  if (timing) {
    const auto t2 = std::chrono::steady_clock::now();
    fprintf(stderr, "cgb_csr_from_mm_file: parse %.3f s (%zu chunks), upload + GPU assembly %.3f s (%zu triplets)\n",
            std::chrono::duration<double>(t1 - t0).count(), pr.size(), std::chrono::duration<double>(t2 - t1).count(), total);
  }
  if (rc != CGB_OK) return rc;
  // The reference dialect drops empty rows but leaves the column ids alone: the reference's own spmv would then read
  // outside its vector (undefined behaviour there, a device fault here).  Such a file cannot be an SPD system: refuse it.
  if ((*out)->n_cols > (*out)->n) {
    const long long n = (long long)(*out)->n, span = (long long)(*out)->n_cols;
    cgb_csr_destroy(*out);
    *out = nullptr;
    return fail(CGB_ERR_INVALID, "matrix has %lld empty row(s): column ids reach %lld but only %lld rows hold entries "
                "(the reference reader compacts rows without renumbering columns)", span - n, span, n);
  }
  return CGB_OK;
}

} // extern "C"
