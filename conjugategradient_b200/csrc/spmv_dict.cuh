// spmv_dict.cuh — row-pattern dictionary engine (kEngDict): 2 bytes per ROW instead of 12 per non-zero.
//
// SURVEY.md §8f row 4 ("byte-reduction formats: stencil-aware value dedup, short column indices").  The CG
// iteration is HBM-bound and at C4 (3D 7-point Poisson 512^3) 88 of the 168 bytes per row and iteration are
// the CSR arrays (12 B per non-zero, src/LinearAlgebraTypes.hpp:57-132).  A matrix that comes from a
// constant-coefficient stencil repeats a handful of ROW PATTERNS — the sequence of (column - row, value) pairs
// of a row: 27 of them for the 7-point Laplacian on a box (interior + faces, edges, corners), 27 for the
// 27-point one.  When a matrix has at most kDictMaxPatterns distinct patterns it is stored losslessly as
//
//   ids[n]        uint16   pattern of each row
//   ptr[P + 1]    int32    pattern p owns entries [ptr[p], ptr[p+1])          \  a few KB, copied to shared
//   off[E], val[E]         its column offsets (int32) and values (fp64)       /  memory by every CTA
//
// Entries keep their CSR order inside the row and the row sum is the same left-to-right FMA chain as the
// one-lane-per-row CSR engines, so results are bit-identical to them.  A warp whose 32 rows share one pattern
// (all of them, away from the boundary) reads offsets and values as shared-memory broadcasts: per entry one
// gather, one FMA, one address add.  Explicit option (CGB_FMT_DICT); the default engines stream the caller's
// CSR arrays as they are.
#pragma once
#include "spmv.cuh"

namespace cgb {

constexpr int kDictMaxPatterns = 1024; // distinct row patterns
constexpr int kDictMaxEntries = 3072;  // entries over all patterns (36 KB of shared memory)
constexpr int kDictMaxRow = 256;       // longest row the format accepts

struct DictView {
  const uint16_t *ids; // n
  const int *ptr;      // npat + 1
  const int *off;      // nent
  const double *val;   // nent
  int npat, nent;
};

__host__ __device__ constexpr size_t dict_smem_bytes(int npat, int nent) {
  return (size_t)nent * 8 + (size_t)nent * 4 + (size_t)(npat + 1) * 4 + 16;
}

// One thread per row.
template <class G, class E>
__device__ __forceinline__ void spmv_rows_dict(int64_t n, const DictView &D, const G &g, const E &e, int bid, int nb,
                                               double &a0, double &a1, unsigned char *smem) {
  double *s_val = reinterpret_cast<double *>(smem);
  int *s_off = reinterpret_cast<int *>(s_val + D.nent);
  int *s_ptr = s_off + D.nent;
  for (int i = threadIdx.x; i < D.nent; i += blockDim.x) {
    s_val[i] = D.val[i];
    s_off[i] = D.off[i];
  }
  for (int i = threadIdx.x; i <= D.npat; i += blockDim.x) s_ptr[i] = D.ptr[i];
  __syncthreads();
  // Rows are dealt 32 per warp round-robin over the grid: neighbouring row blocks are in flight at the same time, so
  // the far stencil neighbours of a row (the +-m^2 planes) are still in L2 when they are needed.  (Contiguous runs per
  // CTA, which keep the +-m neighbours in L1, measured slower: C4 1.21 ms against 1.05 ms.)
  const int64_t stride = (int64_t)nb * blockDim.x;
  const int64_t base0 = (int64_t)bid * blockDim.x + (threadIdx.x & ~31);
  const int lane = threadIdx.x & 31;
  for (int64_t base = base0; base < n; base += stride) { // warp-uniform
    const int64_t row = base + lane;
    const bool live = row < n;
    const int pid = live ? (int)__ldg(D.ids + row) : -1;
    const int pid0 = __shfl_sync(0xffffffffu, pid, 0);
    double sum = 0.0;
    if (__all_sync(0xffffffffu, pid == pid0 || !live)) { // one pattern for the whole warp: broadcast reads
      int k = s_ptr[pid0];
      const int ke = s_ptr[pid0 + 1];
      if (live) {
        for (; k + 4 <= ke; k += 4) {
          const double x0 = g((int)(row + s_off[k])), x1 = g((int)(row + s_off[k + 1]));
          const double x2 = g((int)(row + s_off[k + 2])), x3 = g((int)(row + s_off[k + 3]));
          sum = fma(s_val[k], x0, sum);
          sum = fma(s_val[k + 1], x1, sum);
          sum = fma(s_val[k + 2], x2, sum);
          sum = fma(s_val[k + 3], x3, sum);
        }
        for (; k < ke; k++) sum = fma(s_val[k], g((int)(row + s_off[k])), sum);
      }
    } else if (live) { // boundary rows: every lane walks its own pattern
      for (int k = s_ptr[pid], ke = s_ptr[pid + 1]; k < ke; k++) sum = fma(s_val[k], g((int)(row + s_off[k])), sum);
    }
    if (live) e.row(row, sum, a0, a1);
  }
}

} // namespace cgb
