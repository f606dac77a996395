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
//   hdr[P]        int32    pattern p owns entries [hdr & 0xffff, + (hdr >> 16))   \  a few KB, copied to shared
//   off[E], val[E]         its column offsets (int32) and values (fp64)           /  memory by every CTA
//
// Entries keep their CSR order inside the row and the row sum is the same left-to-right FMA chain as the
// one-lane-per-row CSR engines, so results are bit-identical to them.  A warp whose 32 rows share one pattern
// (all of them, away from the boundary) reads offsets and values as shared-memory broadcasts: per entry one
// gather, one FMA, one address multiply-add.  Explicit option (CGB_FMT_DICT); the default engines stream the caller's
// CSR arrays as they are.
#pragma once
#include "spmv.cuh"

namespace cgb {

constexpr int kDictMaxPatterns = 1024; // distinct row patterns
constexpr int kDictMaxEntries = 3072;  // entries over all patterns (36 KB of shared memory)
constexpr int kDictMaxRow = 256;       // longest row the format accepts

struct DictView {
  const uint16_t *ids; // n
  const int *ptr;      // npat headers: first entry | length << 16
  const int *off;      // nent (patterns padded to kDictAlign entries)
  const double *val;   // nent
  int npat, nent;
};

__host__ __device__ constexpr size_t dict_smem_bytes(int npat, int nent) {
  return (size_t)nent * 8 + (size_t)nent * 4 + (size_t)(npat + 1) * 4 + 16;
}

// Device table: hdr[p] = first entry | length << 16; every pattern starts at a multiple of 4 entries (the gap is
// never read as data), so offsets come as one 16-byte and values as 16-byte shared-memory loads.

#ifndef CGB_DICT_PREFETCH
#define CGB_DICT_PREFETCH 1 // the next round's pattern id is loaded a round ahead
#endif

constexpr int kDictAlign = 4;
#ifndef CGB_DICT_MINB
#define CGB_DICT_MINB 5
#endif
constexpr int kDictMinCtas = CGB_DICT_MINB; // resident CTAs per SM the register budget is sized for

// sum += C entries (C <= 8) of the pattern block at (sv, so): every gather is issued before the first FMA; the FMA
// chain runs in entry order.
template <int C, class At>
__device__ __forceinline__ double dict_dot(const double *__restrict__ sv, const int *__restrict__ so, const At &x, double sum) {
  int o[(C + 3) / 4 * 4];
#pragma unroll
  for (int q = 0; q < (C + 3) / 4; q++) {
    const int4 t = reinterpret_cast<const int4 *>(so)[q];
    o[4 * q] = t.x, o[4 * q + 1] = t.y, o[4 * q + 2] = t.z, o[4 * q + 3] = t.w;
  }
  double xs[C];
#pragma unroll
  for (int j = 0; j < C; j++) xs[j] = x(o[j]);
  double v[(C + 1) / 2 * 2];
#pragma unroll
  for (int q = 0; q < (C + 1) / 2; q++) {
    const double2 t = reinterpret_cast<const double2 *>(sv)[q];
    v[2 * q] = t.x, v[2 * q + 1] = t.y;
  }
#pragma unroll
  for (int j = 0; j < C; j++) sum = fma(v[j], xs[j], sum);
  return sum;
}

// The same for a warp whose lanes carry different patterns (it straddles a boundary): every lane walks its own
// pattern under one warp-uniform trip count C, entries past its own length c predicated off — no divergence, one
// round of gathers.  (Taking the branch of each length in turn made such warps twice as slow, and with a row
// stride that is a multiple of the grid line they were always the same warps: a third of the kernel was spent
// waiting for them at the final barrier.)
template <int C, class At>
__device__ __forceinline__ double dict_dot_ragged(const double *__restrict__ sv, const int *__restrict__ so, const At &x, int c,
                                                  double sum) {
  int o[(C + 3) / 4 * 4];
#pragma unroll
  for (int q = 0; q < (C + 3) / 4; q++) {
    int4 t = make_int4(0, 0, 0, 0);
    if (4 * q < c) t = reinterpret_cast<const int4 *>(so)[q];
    o[4 * q] = t.x, o[4 * q + 1] = t.y, o[4 * q + 2] = t.z, o[4 * q + 3] = t.w;
  }
  double xs[C];
#pragma unroll
  for (int j = 0; j < C; j++) xs[j] = j < c ? x(o[j]) : 0.0;
  double v[(C + 1) / 2 * 2];
#pragma unroll
  for (int q = 0; q < (C + 1) / 2; q++) {
    double2 t = make_double2(0.0, 0.0);
    if (2 * q < c) t = reinterpret_cast<const double2 *>(sv)[q];
    v[2 * q] = t.x, v[2 * q + 1] = t.y;
  }
#pragma unroll
  for (int j = 0; j < C; j++)
    if (j < c) sum = fma(v[j], xs[j], sum);
  return sum;
}

#define CGB_DICT_SWITCH(FN, N, ...)                                                                                              \
  switch (N) {                                                                                                                   \
  case 1: sum = FN<1>(__VA_ARGS__); break;                                                                                       \
  case 2: sum = FN<2>(__VA_ARGS__); break;                                                                                       \
  case 3: sum = FN<3>(__VA_ARGS__); break;                                                                                       \
  case 4: sum = FN<4>(__VA_ARGS__); break;                                                                                       \
  case 5: sum = FN<5>(__VA_ARGS__); break;                                                                                       \
  case 6: sum = FN<6>(__VA_ARGS__); break;                                                                                       \
  case 7: sum = FN<7>(__VA_ARGS__); break;                                                                                       \
  case 8: sum = FN<8>(__VA_ARGS__); break;                                                                                       \
  default: break;                                                                                                                \
  }

// One thread per row.  Lanes of a warp that share a pattern (all of them, away from the boundary) read the table as
// shared-memory broadcasts.
template <class G, class E>
__device__ __forceinline__ void spmv_rows_dict(int64_t n, const DictView &D, const G &g, const E &e, int bid, int nb,
                                               double &a0, double &a1, unsigned char *smem) {
  double *s_val = reinterpret_cast<double *>(smem);
  int *s_off = reinterpret_cast<int *>(s_val + D.nent);
  int *s_hdr = s_off + D.nent;
  for (int i = threadIdx.x; i < D.nent; i += blockDim.x) {
    s_val[i] = D.val[i];
    s_off[i] = D.off[i];
  }
  for (int i = threadIdx.x; i < D.npat; i += blockDim.x) s_hdr[i] = D.ptr[i];
  __syncthreads();
  // Row blocks of blockDim.x rows are dealt round-robin over the grid: neighbouring blocks are in flight at the same
  // time, so the far stencil neighbours of a row (the +-m^2 planes) are still in L2 when they are needed.  (Contiguous
  // runs per CTA, which keep the +-m neighbours in L1, measured slower: C4 1.21 ms against 1.05 ms.)  Inside a block
  // the 32-row groups rotate over the warps from round to round, so a group position that is always the expensive one
  // (the grid-line ends when the stride is a multiple of the line) is not always the same warp's.
  // Row numbers fit 32 bits (columns are int32), so the bookkeeping is 32-bit; base + stride stays below 2^32.
  const uint32_t un = (uint32_t)n, stride = (uint32_t)nb * blockDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int slot = warp;
#if CGB_DICT_PREFETCH
  int pid_next;
  {
    const uint32_t row = (uint32_t)bid * blockDim.x + slot * 32 + lane;
    pid_next = row < un ? (int)__ldg(D.ids + row) : -1;
  }
#endif
  for (uint32_t base = (uint32_t)bid * blockDim.x; base < un; base += stride) { // warp-uniform
    const uint32_t row = base + slot * 32 + lane;
    slot = slot + 1 == nwarps ? 0 : slot + 1;
#if CGB_DICT_PREFETCH
    const int pid = pid_next;
    {
      const uint32_t next = base + stride + slot * 32 + lane;
      pid_next = next < un ? (int)__ldg(D.ids + next) : -1;
    }
#else
    const int pid = row < un ? (int)__ldg(D.ids + row) : -1;
#endif
    const int hdr = pid >= 0 ? s_hdr[pid] : 0; // rows past the end: no entries
    int k = hdr & 0xffff, c = hdr >> 16;
    const auto x = g.at((int64_t)row);
    double sum = 0.0;
    if (__all_sync(0xffffffffu, hdr == __shfl_sync(0xffffffffu, hdr, 0))) {
      for (; c > 8; c -= 8, k += 8) sum = dict_dot<8>(s_val + k, s_off + k, x, sum);
      CGB_DICT_SWITCH(dict_dot, c, s_val + k, s_off + k, x, sum)
    } else {
      int cmax = __reduce_max_sync(0xffffffffu, c);
      for (; cmax > 8; cmax -= 8, c -= 8, k += 8) sum = dict_dot_ragged<8>(s_val + k, s_off + k, x, c, sum);
      CGB_DICT_SWITCH(dict_dot_ragged, cmax, s_val + k, s_off + k, x, c, sum)
    }
    if (row < un) e.row((int64_t)row, sum, a0, a1);
  }
}

} // namespace cgb
