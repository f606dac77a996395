// spmv.cuh — CSR SpMV device code with pluggable gather and per-row epilogue.
//
// One body serves every SpMV-shaped pass of the path:
//   y = A x                         VectorOperations::spmv        (reference src/VectorOperations.hpp:455-462)
//   Ap = A p, partial p.Ap          CG loop steps 3+4             (src/CG.hpp:374-379)
//   r = b - A x0, partial r.r       CG set-up                     (src/CG.hpp:324-341)
//   sum (b - A x)^2, sum x^2        CG::accuracy                  (src/CG.hpp:477-499)
// Row sums use L lanes per row (L = 1 .. 32, chosen from the mean row length), values and column
// indices are streamed with L1::no_allocate loads, the gathered vector goes through L1/L2.
#pragma once
#include "common.cuh"

namespace cgb {

// ---- gathers --------------------------------------------------------------------------------
struct GatherPlain {
  const double *x;
  __device__ __forceinline__ double operator()(int c) const { return __ldg(x + c); }
  // gathers relative to one row (dictionary engine): the row's address is formed once
  struct At {
    const double *xr;
    __device__ __forceinline__ double operator()(int off) const { return __ldg(xr + off); }
  };
  __device__ __forceinline__ At at(int64_t row) const {
    const double *xr = x + row;
    asm("" : "+l"(xr)); // keep (row address) + offset: one multiply-add per gather instead of a 64-bit add and a scale
    return At{xr};
  }
};
// p = r + beta * p_old formed on the fly (fused p-update); same rounding as the epilogue's p_new.
struct GatherFusedP {
  const double *r;
  const double *p_old;
  double beta;
  __device__ __forceinline__ double operator()(int c) const { return fma(beta, __ldg(p_old + c), __ldg(r + c)); }
  struct At {
    const double *rr, *pr;
    double beta;
    __device__ __forceinline__ double operator()(int off) const { return fma(beta, __ldg(pr + off), __ldg(rr + off)); }
  };
  __device__ __forceinline__ At at(int64_t row) const {
    const double *rr = r + row, *pr = p_old + row;
    asm("" : "+l"(rr), "+l"(pr));
    return At{rr, pr, beta};
  }
};

// ---- per-row epilogues: row(i, sum) consumes the finished row sum; acc0/acc1 are the thread's
// contributions to up to two fused reductions ----------------------------------------------------
struct EpiStore {
  static constexpr int NRED = 0;
  double *y;
  __device__ __forceinline__ void row(int64_t i, double s, double &, double &) const { y[i] = s; }
};
struct EpiDotP { // Ap[i] = s ; acc0 += p[i] * s
  static constexpr int NRED = 1;
  double *y;
  const double *p;
  __device__ __forceinline__ void row(int64_t i, double s, double &a0, double &) const {
    y[i] = s;
    a0 += p[i] * s;
  }
};
// x[i] += alpha_prev p_old[i] (deferred) ; p_new[i] = r[i] + beta p_old[i] ; Ap[i] = s ; acc0 += p_new[i] * s
struct EpiDotFusedP {
  static constexpr int NRED = 1;
  double *y;
  const double *r;
  const double *p_old;
  double *p_new;
  double *x;
  double beta;
  double alpha_prev;
  __device__ __forceinline__ void row(int64_t i, double s, double &a0, double &) const {
    const double po = p_old[i];
    x[i] = fma(alpha_prev, po, x[i]);
    const double pn = fma(beta, po, r[i]);
    p_new[i] = pn;
    y[i] = s;
    a0 += pn * s;
  }
};
struct EpiInitResidual { // r[i] = b[i] - s ; acc0 += r[i]^2      (src/CG.hpp:324-332, :341)
  static constexpr int NRED = 1;
  const double *b;
  double *r;
  __device__ __forceinline__ void row(int64_t i, double s, double &a0, double &) const {
    const double ri = b[i] - s;
    r[i] = ri;
    a0 += ri * ri;
  }
};
struct EpiAccuracy { // acc0 += (b[i]-s)^2 ; acc1 += x[i]^2        (src/CG.hpp:487-498)
  static constexpr int NRED = 2;
  const double *b;
  const double *x;
  __device__ __forceinline__ void row(int64_t i, double s, double &a0, double &a1) const {
    const double a = b[i] - s;
    a0 += a * a;
    a1 += x[i] * x[i];
  }
};

// ---- L lanes per row, direct loads ------------------------------------------------------------
// Rows [row_lo, row_hi) are dealt to lane groups round-robin over (bid, nb) CTAs; two rows per
// group are in flight at a time.  All 32 lanes of a warp take every shuffle.
template <int L, class G, class E>
__device__ __forceinline__ void spmv_rows_vec(int64_t row_lo, int64_t row_hi, const int *__restrict__ rowptr,
                                              const int *__restrict__ cols, const double *__restrict__ vals,
                                              const G &g, const E &e, int bid, int nb, double &a0, double &a1) {
  constexpr int GPW = 32 / L; // groups per warp
  const int lane = threadIdx.x & 31;
  const int sub = lane % L;
  const int grp = lane / L;
  const int64_t warp_id = (int64_t)bid * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)nb * (blockDim.x >> 5);
  const int64_t stride = nwarps * GPW;
  for (int64_t base = row_lo + warp_id * GPW; base < row_hi; base += 2 * stride) {
    const int64_t r0 = base + grp, r1 = base + stride + grp;
    int s0 = 0, e0 = 0, s1 = 0, e1 = 0;
    if (r0 < row_hi) {
      s0 = __ldg(rowptr + r0);
      e0 = __ldg(rowptr + r0 + 1);
    }
    if (r1 < row_hi) {
      s1 = __ldg(rowptr + r1);
      e1 = __ldg(rowptr + r1 + 1);
    }
    double sum0 = 0.0, sum1 = 0.0;
    int j0 = s0 + sub, j1 = s1 + sub;
    // both rows advance together while both have entries, so their loads overlap
    while (j0 < e0 && j1 < e1) {
      const int c0 = ld_stream_s32(cols + j0), c1 = ld_stream_s32(cols + j1);
      const double v0 = ld_stream_f64(vals + j0), v1 = ld_stream_f64(vals + j1);
      sum0 = fma(v0, g(c0), sum0);
      sum1 = fma(v1, g(c1), sum1);
      j0 += L;
      j1 += L;
    }
    for (; j0 < e0; j0 += L) sum0 = fma(ld_stream_f64(vals + j0), g(ld_stream_s32(cols + j0)), sum0);
    for (; j1 < e1; j1 += L) sum1 = fma(ld_stream_f64(vals + j1), g(ld_stream_s32(cols + j1)), sum1);
#pragma unroll
    for (int o = L >> 1; o > 0; o >>= 1) {
      sum0 += __shfl_down_sync(0xffffffffu, sum0, o, L);
      sum1 += __shfl_down_sync(0xffffffffu, sum1, o, L);
    }
    if (sub == 0) {
      if (r0 < row_hi) e.row(r0, sum0, a0, a1);
      if (r1 < row_hi) e.row(r1, sum1, a0, a1);
    }
  }
}


// ---- SELL-32-sigma engine -------------------------------------------------------------------------
// Shadow copy of the CSR matrix built by cgb_csr_build_sell (sell.cu): rows are sorted by length inside
// windows of sigma rows, grouped in slices of 32 slots, every slice padded to its longest row and stored
// column-major (entry j of slot l at off[s] + 32 j + l).  A warp owns a slice, lane l owns slot l: every
// load of values / column indices is a fully coalesced 256 / 128-byte warp access whatever the row
// lengths are, and a long row costs its slice only (sigma keeps rows of similar length together).  The
// row sum is the reference's own left-to-right order (src/VectorOperations.hpp:455-462).  Padding entries
// carry column -1 and are skipped (never multiplied: 0 * Inf would poison the sum).
__device__ __forceinline__ double ld_stream_f64_ef(const double *p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ int ld_stream_s32_ef(const int *p, uint64_t pol) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}

#ifndef CGB_SELL_U
#define CGB_SELL_U 4
#endif
#ifndef CGB_SELL_MINB
#define CGB_SELL_MINB 4
#endif
constexpr int kSellUnroll = CGB_SELL_U;  // column-major steps per load group
constexpr int kSellMinCtas = CGB_SELL_MINB; // resident CTAs per SM the register budget is sized for

// The warp streams its contiguous range of slices as ONE sequence of column-major steps: the loads of a group of
// kSellUnroll steps are issued while the previous group's gathers come back, whatever the slice boundaries are,
// so narrow slices (short rows, or a column panel of a longer row) cost no extra memory round trip.  Slice
// bookkeeping rides along: widths arrive 32 slices at a time (lane i holds slice sb + i), the slot -> row map is
// read two slices ahead and the panel accumulator one slice ahead of its use.
template <class G, class E>
__device__ __forceinline__ void spmv_rows_sell(int64_t n, int64_t slices, const int64_t *__restrict__ off,
                                               const int *__restrict__ width, const int *__restrict__ cols,
                                               const double *__restrict__ vals, const int *__restrict__ perm, int64_t nlong,
                                               const int *__restrict__ long_rows, const int64_t *__restrict__ chunk,
                                               const int *__restrict__ rowptr, const int *__restrict__ ccols,
                                               const double *__restrict__ cvals, const double *acc, const G &g, const E &e,
                                               int bid, int nb, double &a0, double &a1) {
  constexpr int U = kSellUnroll;
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  const int lane = threadIdx.x & 31;
  const int64_t warp_id = (int64_t)bid * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)nb * (blockDim.x >> 5);
  // every warp streams its own contiguous range of slices of (nearly) equal cost: static, deterministic
  int64_t s = chunk[warp_id];
  const int64_t s_end = chunk[warp_id + 1];
  if (s < s_end) {
    auto slot_row = [&](int64_t sl) -> int { // < 0: no row behind this slot (tail of the last slice, or a long row)
      const int64_t slot = sl * 32 + lane;
      return slot < n ? (perm ? perm[slot] : (int)slot) : -1;
    };
    int row = slot_row(s);
    int row_next = s + 1 < s_end ? slot_row(s + 1) : -1;
    double accv = (acc && row >= 0) ? acc[row] : 0.0; // column panels: what the earlier panels summed for this row
    int64_t sb = s;
    int myw = sb + lane < s_end ? width[sb + lane] : 0;
    int q = 0;
    int left = __shfl_sync(0xffffffffu, myw, 0);
    double sum = 0.0;
    auto finish = [&]() { // slice s is complete (warp-uniform)
      if (row >= 0) e.row(row, sum + accv, a0, a1);
      sum = 0.0;
      s++;
      if (++q == 32) {
        sb += 32;
        myw = sb + lane < s_end ? width[sb + lane] : 0;
        q = 0;
      }
      left = __shfl_sync(0xffffffffu, myw, q);
      row = row_next;
      accv = (acc && row >= 0) ? acc[row] : 0.0;
      row_next = s + 1 < s_end ? slot_row(s + 1) : -1;
    };
    while (s < s_end && left == 0) finish(); // leading empty slices
    const int64_t j0 = off[chunk[warp_id]];
    const int64_t T = (off[s_end] - j0) >> 5; // column-major steps of this range
    const int *cp = cols + j0 + lane;
    const double *vp = vals + j0 + lane;
    int c[U];
    double v[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (u < T) {
        c[u] = ld_stream_s32_ef(cp + 32 * u, pol);
        v[u] = ld_stream_f64_ef(vp + 32 * u, pol);
      } else {
        c[u] = -1;
        v[u] = 0.0;
      }
    }
    for (int64_t t = 0; t < T; t += U) {
      double x[U];
#pragma unroll
      for (int u = 0; u < U; u++) x[u] = c[u] >= 0 ? g(c[u]) : 0.0; // padding carries column -1: never multiplied
      int cn[U];
      double vn[U];
      cp += 32 * U;
      vp += 32 * U;
#pragma unroll
      for (int u = 0; u < U; u++) {
        if (t + U + u < T) {
          cn[u] = ld_stream_s32_ef(cp + 32 * u, pol);
          vn[u] = ld_stream_f64_ef(vp + 32 * u, pol);
        } else {
          cn[u] = -1;
          vn[u] = 0.0;
        }
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        if (t + u < T) { // warp-uniform
          if (c[u] >= 0) sum = fma(v[u], x[u], sum);
          if (--left == 0) {
            finish();
            while (s < s_end && left == 0) finish();
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        c[u] = cn[u];
        v[u] = vn[u];
      }
    }
  }
  // long rows: a warp per row over the CSR arrays (coalesced), two steps in flight, shuffle reduction
  for (int64_t k = warp_id; k < nlong; k += nwarps) {
    const int64_t row = long_rows[k];
    const int b = __ldg(rowptr + row), en = __ldg(rowptr + row + 1);
    double s0 = 0.0, s1 = 0.0;
    int j = b + lane;
    for (; j + 32 < en; j += 64) {
      const int c0 = ld_stream_s32_ef(ccols + j, pol), c1 = ld_stream_s32_ef(ccols + j + 32, pol);
      const double v0 = ld_stream_f64_ef(cvals + j, pol), v1 = ld_stream_f64_ef(cvals + j + 32, pol);
      s0 = fma(v0, g(c0), s0);
      s1 = fma(v1, g(c1), s1);
    }
    if (j < en) s0 = fma(ld_stream_f64_ef(cvals + j, pol), g(ld_stream_s32_ef(ccols + j, pol)), s0);
    double sum = warp_sum(s0 + s1);
    if (lane == 0) {
      if (acc) sum += acc[row];
      e.row(row, sum, a0, a1);
    }
  }
}

template <class G, class E>
__device__ __forceinline__ void spmv_rows_dispatch(int lanes, int64_t row_lo, int64_t row_hi, const int *rowptr,
                                                   const int *cols, const double *vals, const G &g, const E &e,
                                                   int bid, int nb, double &a0, double &a1) {
  switch (lanes) {
  case 1: spmv_rows_vec<1>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  case 2: spmv_rows_vec<2>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  case 4: spmv_rows_vec<4>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  case 8: spmv_rows_vec<8>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  case 16: spmv_rows_vec<16>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  default: spmv_rows_vec<32>(row_lo, row_hi, rowptr, cols, vals, g, e, bid, nb, a0, a1); break;
  }
}

} // namespace cgb
