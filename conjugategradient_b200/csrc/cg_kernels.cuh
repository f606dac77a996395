// cg_kernels.cuh — the kernels of one CG iteration (reference loop: src/CG.hpp:359-436).
//
//   K0  cg_init_kernel      r = b - A x0, rr = r.r                       (:314-341)
//   K3  cg_pupdate_kernel   x += alpha_prev p_old ; p = r + beta p_old   (only when not fused)   (:390, :418)
//   K1  cg_spmv_kernel      Ap = A p, pAp = p.Ap, alpha = rr / pAp        (:374-386)
//       fused flavour: p = r + beta p_old is formed in the gather and written once, and the
//       deferred x += alpha_prev p_old rides on the same row-aligned read of p_old
//   K2  cg_update_kernel    r -= alpha Ap, rr' = r.r,                     (:392-416)
//                           stop test on the OLD rr, beta = rr'/rr, rr = rr', ++iter
//   KX  cg_xflush_kernel    x += alpha p for the last trip (once, after the loop)   (:390)
//   K4  residual_kernel     sum (b - A x)^2, sum x^2                      (:477-499)
//
// Deferred x update: the reference applies x += alpha_k p_k inside trip k (:390).  x is never read
// inside the loop, so the same fma is applied one kernel later, where p_k is read anyway (the next
// p-update), and once more after the loop for the last trip.  Same operands, same rounding, one
// vector read fewer per iteration (the K2 pass no longer touches x or p).
//
// Every kernel is gated on the device-side state (no host round trip per iteration): once the
// stop test has fired or the trip cap is reached, later launches return immediately, so a CUDA
// graph holding a batch of iterations can overshoot harmlessly.
#pragma once
#include "spmv_dict.cuh"
#include "spmv_tma.cuh"
#include "vec_tma.cuh"

namespace cgb {

struct CsrView {
  int64_t n;
  int64_t nnz;
  const int *rowptr;
  const int *cols;
  const double *vals;
  // SELL-32-sigma shadow copy (engine kEngSell only)
  int64_t sell_slices;
  const int64_t *sell_off;
  const int *sell_width;
  const int *sell_cols;
  const double *sell_vals;
  const int *sell_perm;
  int64_t sell_nlong;
  const int *sell_long;
  const int64_t *sell_chunk;
  // column panels (kEngSell): when set, the row sum handed to the epilogue is panel_acc[row] + (this panel's sum)
  const double *panel_acc;
  // row-pattern dictionary copy (engine kEngDict only)
  DictView dict;
};

// One entry point for the three row engines: CSR with direct loads and L lanes per row (256 threads),
// CSR with TMA-staged tiles (288 threads, dynamic shared memory), SELL-32-sigma (256 threads, a warp
// per slice).
template <int L, int ENG, class G, class E>
__device__ __forceinline__ void spmv_rows(const CsrView &A, const G &g, const E &e, double &a0, double &a1) {
  if constexpr (ENG == kEngTma) {
    extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
    spmv_rows_tma<L>(A.n, A.nnz, A.rowptr, A.cols, A.vals, g, e, blockIdx.x, gridDim.x, a0, a1, cgb_dyn_smem);
  } else if constexpr (ENG == kEngSell) {
    spmv_rows_sell(A.n, A.sell_slices, A.sell_off, A.sell_width, A.sell_cols, A.sell_vals, A.sell_perm, A.sell_nlong, A.sell_long,
                   A.sell_chunk, A.rowptr, A.cols, A.vals, A.panel_acc, g, e, blockIdx.x, gridDim.x, a0, a1);
  } else if constexpr (ENG == kEngDict) {
    extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
    spmv_rows_dict(A.n, A.dict, g, e, blockIdx.x, gridDim.x, a0, a1, cgb_dyn_smem);
  } else {
    spmv_rows_vec<L>(0, A.n, A.rowptr, A.cols, A.vals, g, e, blockIdx.x, gridDim.x, a0, a1);
  }
}
// min-CTAs 0 = unspecified: an explicit 1 lets ptxas grow the TMA kernel from 54 to 90 registers (2 CTAs per SM
// instead of 3: C3 894 -> 720 it/s)
#define CGB_BOUNDS(ENG) \
  __launch_bounds__((ENG) == kEngTma ? kTmaThreads : kBlock, (ENG) == kEngSell ? kSellMinCtas : (ENG) == kEngDict ? kDictMinCtas : 0)

__device__ __forceinline__ bool cg_active(const CgState *st) {
  // reference: do { ... } while (counter++ < N && !*is_done)   (src/CG.hpp:436)
  return !st->done && st->iter < st->max_iter;
}

// ---- K0 -------------------------------------------------------------------------------------
template <int L, int ENG>
__global__ void CGB_BOUNDS(ENG) cg_init_kernel(CsrView A, const double *__restrict__ x,
                                                          const double *__restrict__ b, double *__restrict__ r,
                                                          CgState *st, double *partials, double *hist) {
  __shared__ double red[32];
  double a0 = 0.0, a1 = 0.0;
  GatherPlain g{x};
  EpiInitResidual e{b, r};
  spmv_rows<L, ENG>(A, g, e, a0, a1);
  a0 = block_sum(a0, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = a0;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rr = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) {
      st->rr = rr;
      st->rr0 = rr;
      st->pAp = 0.0;
      st->alpha = 0.0;
      st->beta = 0.0; // first p-update yields p = r
      st->iter = 0;
      st->done = 0;
      st->nan_stop = 0;
      if (st->record_history && st->hist_cap > 0) hist[0] = rr;
    }
  }
}

// ---- residual replacement (optional; not in the reference) ------------------------------------------
// r = b - A x recomputed from the current iterate, rr = r.r: run every `residual_replace_every` trips, after the
// pending x update has been flushed.  The recurrence residual the reference iterates on (src/CG.hpp:392-416)
// drifts away from b - A x once rounding dominates (972 trips at Tester's 1e-24 with the true
// residual stuck at 2.8e-12); replacing it keeps the stop rule honest.  alpha is cleared because the flush has
// already applied it (the next p-update would apply it again).
template <int L, int ENG>
__global__ void CGB_BOUNDS(ENG) cg_replace_kernel(CsrView A, const double *__restrict__ x, const double *__restrict__ b,
                                                  double *r, CgState *st, double *partials) {
  if (!cg_active(st)) return;
  __shared__ double red[32];
  double a0 = 0.0, a1 = 0.0;
  GatherPlain g{x};
  EpiInitResidual e{b, r};
  spmv_rows<L, ENG>(A, g, e, a0, a1);
  a0 = block_sum(a0, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = a0;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rr = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) {
      st->rr = rr;
      st->alpha = 0.0;
    }
  }
}

// ---- K3 (unfused only) -------------------------------------------------------------------------
static __global__ void __launch_bounds__(kBlock) cg_pupdate_kernel(int64_t n, const double *__restrict__ r,
                                                             double *__restrict__ p, double *__restrict__ x,
                                                             const CgState *st) {
  if (!cg_active(st)) return;
  const double beta = st->beta;
  const double alpha = st->alpha; // of the previous trip (0 before the first: p is zero then)
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = n >> 1;
  const double2 *r2 = reinterpret_cast<const double2 *>(r);
  double2 *p2 = reinterpret_cast<double2 *>(p);
  double2 *x2 = reinterpret_cast<double2 *>(x);
  for (int64_t i = tid; i < n2; i += nthr) {
    const double2 rv = r2[i];
    double2 pv = p2[i];
    double2 xv = x2[i];
    xv.x = fma(alpha, pv.x, xv.x);
    xv.y = fma(alpha, pv.y, xv.y);
    pv.x = fma(beta, pv.x, rv.x);
    pv.y = fma(beta, pv.y, rv.y);
    x2[i] = xv;
    p2[i] = pv;
  }
  if ((n & 1) && tid == 0) {
    const double po = p[n - 1];
    x[n - 1] = fma(alpha, po, x[n - 1]);
    p[n - 1] = fma(beta, po, r[n - 1]);
  }
}

// K3 with TMA-staged tiles (vec_tma.cuh): kTmaThreads threads, vec_tma_smem_bytes(3) of dynamic shared memory.  Same
// arithmetic, element for element; chosen for vectors long enough to fill the two-stage pipeline (csr_plan).
static __global__ void __launch_bounds__(kTmaThreads) cg_pupdate_tma_kernel(int64_t n, const double *r, double *p, double *x,
                                                                           const CgState *st) {
  if (!cg_active(st)) return;
  extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
  vec_tma_pupdate(n, r, p, x, st->alpha, st->beta, cgb_dyn_smem);
}

// ---- K1 -------------------------------------------------------------------------------------
// FUSED: p buffers ping-pong on the iteration parity; p_old = pbuf[iter & 1], p_new = pbuf[~iter & 1].
template <int L, int ENG, bool FUSED>
__global__ void CGB_BOUNDS(ENG) cg_spmv_kernel(CsrView A, const double *__restrict__ r, double *pbuf0,
                                                          double *pbuf1, double *__restrict__ Ap, double *x,
                                                          CgState *st, double *partials, double *local_out,
                                                          const double *gather) {
  // gather != NULL (multi-GPU all-gather mode, unfused only): the matrix has GLOBAL columns and gathers from the
  // all-gathered copy of p, while pbuf0 stays this rank's own slice (the p[i] of the p.Ap epilogue)
  if (!cg_active(st)) return;
  __shared__ double red[32];
  double a0 = 0.0, a1 = 0.0;
  if (FUSED) {
    const bool odd = (st->iter & 1) != 0;
    const double *p_old = odd ? pbuf1 : pbuf0;
    double *p_new = odd ? pbuf0 : pbuf1;
    const double beta = st->beta;
    GatherFusedP g{r, p_old, beta};
    EpiDotFusedP e{Ap, r, p_old, p_new, x, beta, st->alpha};
    spmv_rows<L, ENG>(A, g, e, a0, a1);
  } else {
    GatherPlain g{gather ? gather : pbuf0};
    EpiDotP e{Ap, pbuf0};
    spmv_rows<L, ENG>(A, g, e, a0, a1);
  }
  a0 = block_sum(a0, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = a0;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double pAp = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) {
      if (local_out) {
        *local_out = pAp; // multi-GPU: all-reduced before K2 forms alpha
      } else {
        st->pAp = pAp;
        st->alpha = (st->precond ? st->rho : st->rr) / pAp; // src/CG.hpp:385-386 (rho = r.z with Jacobi)
      }
    }
  }
}

// ---- K2 -------------------------------------------------------------------------------------
// Single GPU (pAp_in == NULL, local_out == NULL): alpha comes from the state and the last CTA advances
// the state.  Multi-GPU: alpha = rr / *pAp_in (the all-reduced p.Ap) is also stored to the state for
// the deferred x update, the local r.r partial goes to *local_out and cg_advance_kernel advances the
// state after the second all-reduce.
// r -= alpha Ap over a grid-stride range; returns this thread's share of r.r
__device__ __forceinline__ double cg_update_body(int64_t n, double *__restrict__ r, const double *__restrict__ Ap,
                                                 double alpha) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = n >> 1;
  double2 *r2 = reinterpret_cast<double2 *>(r);
  const double2 *q2 = reinterpret_cast<const double2 *>(Ap);
  double acc = 0.0;
  for (int64_t i = tid; i < n2; i += nthr) {
    double2 rv = r2[i];
    const double2 qv = q2[i];
    rv.x = fma(-alpha, qv.x, rv.x);
    rv.y = fma(-alpha, qv.y, rv.y);
    r2[i] = rv;
    acc = fma(rv.x, rv.x, acc);
    acc = fma(rv.y, rv.y, acc);
  }
  if ((n & 1) && tid == 0) {
    const double rv = fma(-alpha, Ap[n - 1], r[n - 1]);
    r[n - 1] = rv;
    acc = fma(rv, rv, acc);
  }
  return acc;
}

// End of a trip (one thread): stop test on the rr from BEFORE this trip's update (src/CG.hpp:396-404),
// beta = rr'/rr (:414), rr = rr' (:415), ++trip counter.
__device__ __forceinline__ void cg_advance_state(CgState *st, double rr_new, double *hist) {
  const double rr = st->rr;
  if (isnan(rr) || sqrt(rr) <= st->tol) {
    st->done = 1;
    if (isnan(rr)) st->nan_stop = 1;
  }
  st->beta = rr_new / rr;
  st->rr = rr_new;
  const long long it = st->iter + 1;
  st->iter = it;
  if (st->record_history && it < st->hist_cap) hist[it] = rr_new;
}

static __global__ void __launch_bounds__(kBlock) cg_update_kernel(int64_t n, double *__restrict__ r,
                                                            const double *__restrict__ Ap, CgState *st,
                                                            double *partials, double *hist, const double *pAp_in,
                                                            double *local_out) {
  if (!cg_active(st)) return;
  __shared__ double red[32];
  const double alpha = pAp_in ? st->rr / *pAp_in : st->alpha;
  double acc = cg_update_body(n, r, Ap, alpha);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rr_new = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) {
      if (!local_out) {
        cg_advance_state(st, rr_new, hist);
      } else {
        st->alpha = alpha; // every CTA used this value; kept for the deferred x update
        *local_out = rr_new;
      }
    }
  }
}

// K2 with TMA-staged tiles (single GPU: alpha from the state, the last CTA advances the state).
static __global__ void __launch_bounds__(kTmaThreads) cg_update_tma_kernel(int64_t n, double *r, const double *Ap, CgState *st,
                                                                          double *partials, double *hist) {
  if (!cg_active(st)) return;
  extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
  __shared__ double red[32];
  double acc = vec_tma_update(n, r, Ap, st->alpha, cgb_dyn_smem);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rr_new = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) cg_advance_state(st, rr_new, hist);
  }
}

// ---- Jacobi-preconditioned variants (optional; not in the reference: SURVEY.md §8f row 4) --------------
// z = D^-1 r is never stored: it is formed on the fly from dinv = 1 / diag(A) where it is needed.
//   p-update   p = z + beta p           (and the deferred x update, as in cg_pupdate_kernel)
//   r-update   r -= alpha Ap ; rr' = r.r (stop rule, unchanged) ; rho' = r.z ; beta = rho'/rho
static __global__ void __launch_bounds__(kBlock) cg_pupdate_jacobi_kernel(int64_t n, const double *__restrict__ r,
                                                                    const double *__restrict__ dinv, double *__restrict__ p,
                                                                    double *__restrict__ x, const CgState *st) {
  if (!cg_active(st)) return;
  const double beta = st->beta, alpha = st->alpha;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    const double po = p[i];
    x[i] = fma(alpha, po, x[i]);
    p[i] = fma(beta, po, dinv[i] * r[i]);
  }
}

static __global__ void __launch_bounds__(kBlock) cg_update_jacobi_kernel(int64_t n, double *__restrict__ r,
                                                                   const double *__restrict__ Ap,
                                                                   const double *__restrict__ dinv, CgState *st,
                                                                   double *partials, double *hist) {
  if (!cg_active(st)) return;
  __shared__ double red[32];
  const double alpha = st->alpha;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double a_rr = 0.0, a_rz = 0.0;
  for (int64_t i = tid; i < n; i += nthr) {
    const double rv = fma(-alpha, Ap[i], r[i]);
    r[i] = rv;
    a_rr = fma(rv, rv, a_rr);
    a_rz = fma(dinv[i] * rv, rv, a_rz);
  }
  a_rr = block_sum(a_rr, red);
  a_rz = block_sum(a_rz, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = a_rr;
    partials[kMaxPartials + blockIdx.x] = a_rz;
  }
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rr_new = reduce_partials(partials, gridDim.x, red);
    const double rho_new = reduce_partials(partials + kMaxPartials, gridDim.x, red);
    if (threadIdx.x == 0) {
      const double rho = st->rho;
      cg_advance_state(st, rr_new, hist); // stop rule on the un-preconditioned residual, as the reference's
      st->beta = rho_new / rho;
      st->rho = rho_new;
    }
  }
}

// rho = r.z from the current r (after K0 and after a residual replacement)
static __global__ void __launch_bounds__(kBlock) cg_rho_kernel(int64_t n, const double *__restrict__ r,
                                                         const double *__restrict__ dinv, CgState *st, double *partials,
                                                         int gated) {
  if (gated && !cg_active(st)) return;
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double a = 0.0;
  for (int64_t i = tid; i < n; i += nthr) a = fma(dinv[i] * r[i], r[i], a);
  a = block_sum(a, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = a;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double rho = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) st->rho = rho;
  }
}

// dinv[i] = 1 / A_ii; *bad = 1 when a row has no (or a zero) diagonal entry
static __global__ void diag_inv_kernel(int64_t n, const int *__restrict__ rowptr, const int *__restrict__ cols,
                                       const double *__restrict__ vals, double *__restrict__ dinv, int *bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double d = 0.0;
  for (int j = rowptr[i], e = rowptr[i + 1]; j < e; j++)
    if (cols[j] == i) d += vals[j];
  if (d == 0.0) atomicExch(bad, 1);
  dinv[i] = 1.0 / d;
}

// ---- KX -------------------------------------------------------------------------------------
// The x update of the LAST trip (the ones before it were applied by the following p-update).
// Launched once after the loop; p of the last trip is pbuf0 (unfused) or the buffer the parity
// of the trip counter selects (fused).
static __global__ void __launch_bounds__(kBlock) cg_xflush_kernel(int64_t n, double *__restrict__ x, const double *pbuf0,
                                                            const double *pbuf1, int fused, const CgState *st, int gated) {
  if (st->iter <= 0) return; // no trip ran: nothing pending
  if (gated && !cg_active(st)) return; // inside the loop (residual replacement): the flush after the loop does the last trip
  const double alpha = st->alpha;
  const double *p = (fused && (st->iter & 1)) ? pbuf1 : pbuf0;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = n >> 1;
  double2 *x2 = reinterpret_cast<double2 *>(x);
  const double2 *p2 = reinterpret_cast<const double2 *>(p);
  for (int64_t i = tid; i < n2; i += nthr) {
    double2 xv = x2[i];
    const double2 pv = p2[i];
    xv.x = fma(alpha, pv.x, xv.x);
    xv.y = fma(alpha, pv.y, xv.y);
    x2[i] = xv;
  }
  if ((n & 1) && tid == 0) x[n - 1] = fma(alpha, p[n - 1], x[n - 1]);
}

// Multi-GPU (NCCL transport): advance the state from the all-reduced scalars (one thread).
static __global__ void cg_advance_kernel(CgState *st, const double *rr_global, double *hist) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (!cg_active(st)) return;
  cg_advance_state(st, *rr_global, hist);
}

// ---- K4 -------------------------------------------------------------------------------------
template <int L, int ENG>
__global__ void CGB_BOUNDS(ENG) residual_kernel(CsrView A, const double *__restrict__ x,
                                                           const double *__restrict__ b, double *out2,
                                                           double *partials, unsigned int *ticket) {
  __shared__ double red[32];
  double a0 = 0.0, a1 = 0.0;
  GatherPlain g{x};
  EpiAccuracy e{b, x};
  spmv_rows<L, ENG>(A, g, e, a0, a1);
  a0 = block_sum(a0, red);
  a1 = block_sum(a1, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = a0;
    partials[kMaxPartials + blockIdx.x] = a1;
  }
  if (publish_and_elect(ticket, gridDim.x)) {
    const double s0 = reduce_partials(partials, gridDim.x, red);
    const double s1 = reduce_partials(partials + kMaxPartials, gridDim.x, red);
    if (threadIdx.x == 0) {
      out2[0] = s0;
      out2[1] = s1;
    }
  }
}

// y = A x (plain).  Also the partial pass of a column panel (A.panel_acc == y: y += A_k x; y is therefore
// not __restrict__), gated on the CG state when it runs inside the loop.
template <int L, int ENG>
__global__ void CGB_BOUNDS(ENG) spmv_kernel(CsrView A, const double *__restrict__ x, double *y, const CgState *gate) {
  if (gate && !cg_active(gate)) return;
  double a0 = 0.0, a1 = 0.0;
  GatherPlain g{x};
  EpiStore e{y};
  spmv_rows<L, ENG>(A, g, e, a0, a1);
}

} // namespace cgb
