// persistent.cuh — the whole CG loop as ONE cooperative kernel (CGB_LOOP_PERSISTENT), for systems
// small enough that launch and inter-kernel latency, not HBM, set the iteration time (BASELINE config
// C1: n = 16,384 — 2.5 MB per iteration, everything lives in L2).
//
// The reference runs 12 submissions and a host wait per iteration (src/CG.hpp:359-436); the graph path
// runs 3 kernels per iteration; here an iteration is two phases separated by grid-wide barriers:
//
//   phase B   p = r + beta p_old formed in the gather and written once (ping-pong buffers), the deferred
//             x += alpha_prev p_old, Ap = A p, per-CTA partial of p.Ap            (:374-379, :390, :418)
//   -- grid.sync --   every CTA sums the partials in the same fixed order: alpha = rr / p.Ap   (:385-386)
//   phase C   r -= alpha Ap, per-CTA partial of r.r                                (:392, :406)
//   -- grid.sync --   every CTA forms r.r, applies the reference's stop rule to the OLD rr (:396-404),
//             beta = rr'/rr (:414), advances the trip counter (:436)
//
// All CTAs carry identical copies of the scalars in registers (same partials, same order => same
// bits), so no broadcast through memory is needed and the result is reproducible run to run.
// Rows are served by the direct-load CSR engine (spmv_rows_vec): at these sizes the matrix is
// L2-resident and TMA staging would only add latency.
#pragma once
#include <cooperative_groups.h>

#include "cg_kernels.cuh"

namespace cgb {

namespace cg = cooperative_groups;

// block-wide sum of `v`, then the fixed-order sum of all CTAs' partials after the grid barrier;
// returns the total in EVERY thread
__device__ __forceinline__ double grid_sum(cg::grid_group &grid, double v, double *partials, double *red, double *bcast) {
  v = block_sum(v, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = v;
  grid.sync();
  const double s = reduce_partials(partials, gridDim.x, red);
  if (threadIdx.x == 0) *bcast = s;
  __syncthreads();
  const double out = *bcast;
  __syncthreads();
  return out;
}

template <int L>
__global__ void __launch_bounds__(kBlock) cg_persistent_kernel(CsrView A, double *__restrict__ x, double *__restrict__ r,
                                                                double *pbuf0, double *pbuf1, double *__restrict__ Ap,
                                                                CgState *st, double *partials, double *hist) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double red[32];
  __shared__ double bcast;
  // K0 has armed the state; every thread keeps its own copy of the scalars
  double rr = st->rr, alpha = 0.0, beta = 0.0;
  const double tol = st->tol;
  const long long max_iter = st->max_iter;
  const int record = st->record_history;
  const long long hist_cap = st->hist_cap;
  long long it = 0;
  int done = st->done, nan_stop = 0;
  double *p_old = pbuf0, *p_new = pbuf1;
  double *part_b = partials, *part_c = partials + kMaxPartials;
  while (!done && it < max_iter) {
    double a0 = 0.0, a1 = 0.0;
    {
      GatherFusedP g{r, p_old, beta};
      EpiDotFusedP e{Ap, r, p_old, p_new, x, beta, alpha};
      spmv_rows_vec<L>(0, A.n, A.rowptr, A.cols, A.vals, g, e, blockIdx.x, gridDim.x, a0, a1);
    }
    const double pAp = grid_sum(grid, a0, part_b, red, &bcast);
    alpha = rr / pAp;
    const double acc = cg_update_body(A.n, r, Ap, alpha);
    const double rr_new = grid_sum(grid, acc, part_c, red, &bcast);
    if (isnan(rr) || sqrt(rr) <= tol) { // the rr from BEFORE this trip's update (one-trip lag)
      done = 1;
      nan_stop = isnan(rr) ? 1 : 0;
    }
    beta = rr_new / rr;
    rr = rr_new;
    it++;
    if (record && it < hist_cap && blockIdx.x == 0 && threadIdx.x == 0) hist[it] = rr_new;
    double *t = p_old;
    p_old = p_new;
    p_new = t;
  }
  // the x update of the last trip; its p is the buffer written last (now p_old)
  if (it > 0) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = tid; i < A.n; i += nthr) x[i] = fma(alpha, p_old[i], x[i]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->rr = rr;
    st->alpha = alpha;
    st->beta = beta;
    st->iter = it;
    st->done = done;
    st->nan_stop = nan_stop;
  }
}

} // namespace cgb
