// dict.cu — device-side CSR -> row-pattern dictionary (format and kernel: spmv_dict.cuh).
//
//   1. one pass hashes every row's pattern — the sequence of (column - row, value bits) pairs — and inserts the
//      64-bit hash into a small open-addressing table (atomicCAS) together with the smallest row that carries it;
//      more than kDictMaxPatterns distinct hashes, or a row longer than kDictMaxRow, and the matrix is reported
//      as not compressible (CGB_ERR_UNSUPPORTED: the caller keeps CSR);
//   2. the table comes to the host (a few KB); patterns are numbered in order of first occurrence, their entries
//      are fetched from the representative rows, and the pattern table goes back to the device;
//   3. a second pass gives every row its pattern id (binary search of its hash) and VERIFIES the row against the
//      pattern entry by entry, so a hash collision can never change a result: it only makes the matrix
//      "not compressible".
#include "matrix.cuh"
#include "spmv_dict.cuh"

#include <algorithm>
#include <vector>

using namespace cgb;

namespace {

constexpr int kTab = 4096; // hash-table slots (power of two, >= 4 * kDictMaxPatterns)
constexpr unsigned long long kEmpty = 0xffffffffffffffffull;

struct Buf {
  void *p = nullptr;
  ~Buf() { cudaFree(p); }
  template <class T> T *as() { return (T *)p; }
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long h, unsigned long long x) {
  h ^= x + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
  h *= 0xff51afd7ed558ccdull;
  h ^= h >> 33;
  return h;
}

__device__ __forceinline__ unsigned long long row_hash(int64_t row, const int *__restrict__ rowptr,
                                                       const int *__restrict__ cols, const double *__restrict__ vals) {
  const int b = rowptr[row], e = rowptr[row + 1];
  unsigned long long h = mix64(0x243f6a8885a308d3ull, (unsigned long long)(e - b));
  for (int j = b; j < e; j++) {
    h = mix64(h, (unsigned long long)(unsigned)(int)((int64_t)cols[j] - row));
    h = mix64(h, (unsigned long long)__double_as_longlong(vals[j]));
  }
  return h == kEmpty ? kEmpty - 1 : h;
}

// counters: [0] distinct patterns, [1] overflow flag
__global__ void dict_collect_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols,
                                    const double *__restrict__ vals, unsigned long long *tab, int *rep, int *counters) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += stride) {
    if (*(volatile int *)(counters + 1)) return;
    const unsigned long long h = row_hash(row, rowptr, cols, vals);
    unsigned s = (unsigned)(h >> 20) & (kTab - 1);
    for (int probe = 0; probe < kTab; probe++, s = (s + 1) & (kTab - 1)) {
      unsigned long long cur = *(volatile unsigned long long *)(tab + s);
      if (cur == kEmpty) {
        cur = atomicCAS(tab + s, kEmpty, h);
        if (cur == kEmpty) {
          if (atomicAdd(counters + 0, 1) >= kDictMaxPatterns) atomicExch(counters + 1, 1);
          cur = h;
        }
      }
      if (cur == h) {
        if (*(volatile int *)(rep + s) > (int)row) atomicMin(rep + s, (int)row);
        break;
      }
    }
  }
}

// representative rows -> staging: len[p], off[p * kDictMaxRow + k], val[...]
__global__ void dict_fetch_kernel(int npat, const int *__restrict__ reps, const int *__restrict__ rowptr,
                                  const int *__restrict__ cols, const double *__restrict__ vals, int *__restrict__ len,
                                  int *__restrict__ off, double *__restrict__ val) {
  const int p = blockIdx.x;
  if (p >= npat) return;
  const int64_t row = reps[p];
  const int b = rowptr[row], l = rowptr[row + 1] - b;
  if (threadIdx.x == 0) len[p] = l;
  for (int k = threadIdx.x; k < l && k < kDictMaxRow; k += blockDim.x) {
    off[p * kDictMaxRow + k] = (int)((int64_t)cols[b + k] - row);
    val[p * kDictMaxRow + k] = vals[b + k];
  }
}

// ids[row] = pattern of the row; *bad = 1 when a row is not exactly its pattern (hash collision) or has none
__global__ void dict_assign_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols,
                                   const double *__restrict__ vals, int npat, const unsigned long long *__restrict__ hs,
                                   const int *__restrict__ pid_of, const int *__restrict__ ptr, const int *__restrict__ off,
                                   const double *__restrict__ val, uint16_t *__restrict__ ids, int *bad) {
  const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n) return;
  const unsigned long long h = row_hash(row, rowptr, cols, vals);
  int lo = 0, hi = npat - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (hs[mid] < h) lo = mid + 1; else hi = mid;
  }
  bool ok = hs[lo] == h;
  const int p = pid_of[lo];
  if (ok) {
    const int b = rowptr[row], l = rowptr[row + 1] - b, pb = ptr[p] & 0xffff;
    ok = l == ptr[p] >> 16;
    for (int k = 0; ok && k < l; k++)
      ok = (int)((int64_t)cols[b + k] - row) == off[pb + k] &&
           __double_as_longlong(vals[b + k]) == __double_as_longlong(val[pb + k]);
  }
  if (!ok) atomicExch(bad, 1);
  ids[row] = (uint16_t)p;
}

} // namespace

namespace cgb {

void csr_release_dict(cgb_csr *A) {
  if (!A->dict_ids) return;
  cudaFree(A->dict_ids);
  cudaFree(A->dict_ptr);
  cudaFree(A->dict_off);
  cudaFree(A->dict_val);
  A->dict_ids = nullptr;
  A->dict_ptr = A->dict_off = nullptr;
  A->dict_val = nullptr;
  A->dict_npat = A->dict_nent = 0;
  if (A->ws.graph) {
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
}

// CGB_ERR_UNSUPPORTED (with the reason in cgb_last_error) when the matrix does not fit the format.
int csr_build_dict(cgb_csr *A) {
  if (A->dict_ids || A->n == 0) return CGB_OK;
  cgb_ctx *ctx = A->ctx;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  if (A->max_row > kDictMaxRow)
    return fail(CGB_ERR_UNSUPPORTED, "row-pattern dictionary: longest row has %d entries (limit %d)", A->max_row, kDictMaxRow);
  const int64_t n = A->n;
  Buf tab, rep, cnt;
  CGB_CUDA(cudaMalloc(&tab.p, sizeof(unsigned long long) * kTab));
  CGB_CUDA(cudaMalloc(&rep.p, sizeof(int) * kTab));
  CGB_CUDA(cudaMalloc(&cnt.p, sizeof(int) * 4));
  std::vector<unsigned long long> h_tab(kTab, kEmpty);
  std::vector<int> h_rep(kTab, 0x7fffffff);
  CGB_CUDA(cudaMemcpyAsync(tab.p, h_tab.data(), sizeof(unsigned long long) * kTab, cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(rep.p, h_rep.data(), sizeof(int) * kTab, cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(int) * 4, s));
  const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
  dict_collect_kernel<<<grid, 256, 0, s>>>(n, A->rowptr, A->cols, A->vals, tab.as<unsigned long long>(), rep.as<int>(),
                                           cnt.as<int>());
  CGB_CUDA(cudaGetLastError());
  int h_cnt[4] = {0, 0, 0, 0};
  CGB_CUDA(cudaMemcpyAsync(h_cnt, cnt.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(h_tab.data(), tab.p, sizeof(unsigned long long) * kTab, cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(h_rep.data(), rep.p, sizeof(int) * kTab, cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  if (h_cnt[1] || h_cnt[0] > kDictMaxPatterns)
    return fail(CGB_ERR_UNSUPPORTED, "row-pattern dictionary: more than %d distinct row patterns", kDictMaxPatterns);
  std::vector<std::pair<int, unsigned long long>> pats; // (representative row, hash): numbered by first occurrence
  for (int i = 0; i < kTab; i++)
    if (h_tab[i] != kEmpty) pats.push_back({h_rep[i], h_tab[i]});
  std::sort(pats.begin(), pats.end());
  const int npat = (int)pats.size();
  if (npat == 0) return fail(CGB_ERR_UNSUPPORTED, "row-pattern dictionary: matrix has no rows");
  // entries of the representative rows
  std::vector<int> reps(npat), len(npat), off_st((size_t)npat * kDictMaxRow);
  std::vector<double> val_st((size_t)npat * kDictMaxRow);
  for (int p = 0; p < npat; p++) reps[p] = pats[p].first;
  Buf d_reps, d_len, d_off_st, d_val_st;
  CGB_CUDA(cudaMalloc(&d_reps.p, sizeof(int) * npat));
  CGB_CUDA(cudaMalloc(&d_len.p, sizeof(int) * npat));
  CGB_CUDA(cudaMalloc(&d_off_st.p, sizeof(int) * off_st.size()));
  CGB_CUDA(cudaMalloc(&d_val_st.p, sizeof(double) * val_st.size()));
  CGB_CUDA(cudaMemcpyAsync(d_reps.p, reps.data(), sizeof(int) * npat, cudaMemcpyHostToDevice, s));
  dict_fetch_kernel<<<npat, 64, 0, s>>>(npat, d_reps.as<int>(), A->rowptr, A->cols, A->vals, d_len.as<int>(), d_off_st.as<int>(),
                                        d_val_st.as<double>());
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaMemcpyAsync(len.data(), d_len.p, sizeof(int) * npat, cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(off_st.data(), d_off_st.p, sizeof(int) * off_st.size(), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaMemcpyAsync(val_st.data(), d_val_st.p, sizeof(double) * val_st.size(), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  // device table: header = first entry | length << 16, patterns padded to kDictAlign entries (spmv_dict.cuh)
  std::vector<int> ptr(npat + 1, 0), off;
  std::vector<double> val;
  for (int p = 0; p < npat; p++) {
    if (off.size() > 0xffff) break; // caught by the size check below
    ptr[p] = (int)off.size() | len[p] << 16;
    for (int k = 0; k < len[p]; k++) {
      off.push_back(off_st[(size_t)p * kDictMaxRow + k]);
      val.push_back(val_st[(size_t)p * kDictMaxRow + k]);
    }
    while (off.size() % kDictAlign) {
      off.push_back(0);
      val.push_back(0.0);
    }
  }
  const int nent = (int)off.size();
  if (nent > kDictMaxEntries)
    return fail(CGB_ERR_UNSUPPORTED, "row-pattern dictionary: patterns hold %d entries (limit %d)", nent, kDictMaxEntries);
  // lookup arrays sorted by hash
  std::vector<std::pair<unsigned long long, int>> by_hash(npat);
  for (int p = 0; p < npat; p++) by_hash[p] = {pats[p].second, p};
  std::sort(by_hash.begin(), by_hash.end());
  std::vector<unsigned long long> hs(npat);
  std::vector<int> pid_of(npat);
  for (int p = 0; p < npat; p++) {
    hs[p] = by_hash[p].first;
    pid_of[p] = by_hash[p].second;
  }
  Buf d_hs, d_pid, d_bad;
  uint16_t *ids = nullptr;
  int *d_ptr = nullptr, *d_off = nullptr;
  double *d_val = nullptr;
  Buf g_ids, g_ptr, g_off, g_val; // released on an early return
  CGB_CUDA(cudaMalloc(&d_hs.p, sizeof(unsigned long long) * npat));
  CGB_CUDA(cudaMalloc(&d_pid.p, sizeof(int) * npat));
  CGB_CUDA(cudaMalloc(&d_bad.p, sizeof(int)));
  CGB_CUDA(cudaMalloc(&ids, sizeof(uint16_t) * (size_t)n + 64));
  g_ids.p = ids;
  CGB_CUDA(cudaMalloc(&d_ptr, sizeof(int) * (npat + 1)));
  g_ptr.p = d_ptr;
  CGB_CUDA(cudaMalloc(&d_off, sizeof(int) * std::max(nent, 1)));
  g_off.p = d_off;
  CGB_CUDA(cudaMalloc(&d_val, sizeof(double) * std::max(nent, 1)));
  g_val.p = d_val;
  CGB_CUDA(cudaMemcpyAsync(d_hs.p, hs.data(), sizeof(unsigned long long) * npat, cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(d_pid.p, pid_of.data(), sizeof(int) * npat, cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(d_ptr, ptr.data(), sizeof(int) * (npat + 1), cudaMemcpyHostToDevice, s));
  if (nent) CGB_CUDA(cudaMemcpyAsync(d_off, off.data(), sizeof(int) * nent, cudaMemcpyHostToDevice, s));
  if (nent) CGB_CUDA(cudaMemcpyAsync(d_val, val.data(), sizeof(double) * nent, cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemsetAsync(d_bad.p, 0, sizeof(int), s));
  dict_assign_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, A->rowptr, A->cols, A->vals, npat,
                                                                 d_hs.as<unsigned long long>(), d_pid.as<int>(), d_ptr, d_off,
                                                                 d_val, ids, d_bad.as<int>());
  CGB_CUDA(cudaGetLastError());
  int bad = 0;
  CGB_CUDA(cudaMemcpyAsync(&bad, d_bad.p, sizeof(int), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  if (bad) return fail(CGB_ERR_UNSUPPORTED, "row-pattern dictionary: pattern hash collision, matrix left in CSR");
  g_ids.p = g_ptr.p = g_off.p = g_val.p = nullptr;
  A->dict_ids = ids;
  A->dict_ptr = d_ptr;
  A->dict_off = d_off;
  A->dict_val = d_val;
  A->dict_npat = npat;
  A->dict_nent = nent;
  if (A->ws.graph) {
    cudaGraphExecDestroy(A->ws.graph);
    A->ws.graph = nullptr;
  }
  return CGB_OK;
}

} // namespace cgb

extern "C" int64_t cgb_csr_dict_bytes(const cgb_csr *A) {
  return A && A->dict_ids ? 2 * A->n + (int64_t)dict_smem_bytes(A->dict_npat, A->dict_nent) : 0;
}
extern "C" int cgb_csr_dict_patterns(const cgb_csr *A) { return A ? A->dict_npat : 0; }
