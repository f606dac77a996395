// cluster.cuh — the whole CG loop inside ONE thread-block cluster, matrix and vectors resident in the cluster's
// shared memory (distributed shared memory, sm_90+; 16 CTAs on B200 with the non-portable size allowed).
//
// For systems of BASELINE config C1's size (n = 16,384, 1.0 MB of CSR) the iteration time is pure latency: the
// graph loop pays three kernel boundaries per trip, the cooperative kernel (persistent.cuh) two grid-wide barriers
// of ~2 us each.  A cluster barrier is a hardware barrier (barrier.cluster) and a load from a neighbour CTA's
// shared memory costs ~215 cycles, so when 12 nnz + 24 n bytes fit the cluster's shared memory nothing has to leave
// the SMs during the loop:
//
//   resident, per CTA (R = rows per CTA):   the CTA's rows as a column-major ELL slice (W x R values + packed
//       columns: owner CTA << 16 | index in the owner's slice), its slices of r and of the two p buffers;
//       x, r, p, Ap of a row live in the registers of the thread that owns the row
//   phase B   p_i = r_i + beta p_i; p_new_s[i] = p_i; Ap_i = sum_j a_ij (r_s[c] + beta p_old_s[c]) with the gathers going
//             to the owner CTA's shared memory (p formed in the gather like persistent.cuh, so no barrier is
//             needed between the p-update and the product); partial of p.Ap            (src/CG.hpp:374-379, :418)
//   -- cluster barrier: every CTA has written its partial into every CTA's slot array; warp 0 of each sums them
//   phase C   alpha = rr / p.Ap; x_i += alpha p_i; r_i -= alpha Ap_i; r_s[i] = r_i; partial of r.r  (:385-392, :406)
//   -- cluster barrier, the same exchange: rr', the reference's stop rule on the OLD rr (:396-404), beta = rr'/rr (:414)
//
// Two barriers per trip, no global-memory traffic inside the loop.  Measured at C1: 3.4 us per trip against 6.8 us for
// the cooperative kernel; what is left is the two reductions (two block-level trees, the exchange, the barrier, a
// butterfly and a division: ~2,500 cycles each, tools/cluster_sync_bench.cu) whatever the number of rows.
// Reductions are fixed-order (warp tree, warps in order, a butterfly over the CTAs): bit-reproducible, every CTA
// derives identical scalars.  Row sums are the same left-to-right FMA chain as the one-lane CSR engines.
#pragma once
#include <cooperative_groups.h>

#include "cg_kernels.cuh"

namespace cgb {

namespace cg = cooperative_groups;

constexpr int kClusterMaxCtas = 16;
constexpr size_t kClusterSmemMax = 220 * 1024; // per CTA, dynamic

__host__ __device__ constexpr size_t cluster_smem_bytes(int R, int W) {
  return (size_t)3 * R * 8 + (3 * 32 + 8) * 8 + (size_t)W * R * 12;
}

// Publishes this CTA's sum of v in slot [rank] of every CTA's `part` array and waits for everybody's (cluster
// barrier).  After it, part[0 .. ncta) holds the same numbers in every CTA; the caller sums them in rank order in
// ONE thread and hands the derived scalars to the CTA through shared memory.  (One SM serves 1,024 rows here: 32 warps
// share four schedulers, and two fp64 divisions plus the square root of the stop rule repeated by every thread
// issued more instructions than the SpMV.)
__device__ __forceinline__ void cluster_publish(cg::cluster_group &cl, double v, double *red, double *part, int rank, int ncta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) red[warp] = v;
  __syncthreads();
  if (warp == 0) {
    double s = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.0;
    s = warp_sum(s);
    s = __shfl_sync(0xffffffffu, s, 0);
    if (lane < ncta) cl.map_shared_rank(part, lane)[rank] = s; // my partial into CTA `lane`'s slot array
  }
  cl.sync(); // release / acquire: the remote stores are visible; also fences `red` for the next call
}
// Rank-symmetric tree over the slots (a butterfly: every lane of warp 0 ends with the same bits, 4 dependent adds
// instead of 15); valid in warp 0.
__device__ __forceinline__ double cluster_total(const double *part, int ncta) {
  const int lane = threadIdx.x & 31;
  double t = lane < ncta ? part[lane] : 0.0;
#pragma unroll
  for (int o = kClusterMaxCtas / 2; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;
}

template <int RPT, int TPB>
__global__ void __launch_bounds__(TPB, 1)
    cg_cluster_kernel(CsrView A, double *__restrict__ x, const double *__restrict__ r_g, CgState *st, double *hist, int R, int W) {
  cg::cluster_group cl = cg::this_cluster();
  extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
  const int rank = (int)cl.block_rank(), ncta = (int)cl.num_blocks();
  double *vec_s = reinterpret_cast<double *>(cgb_dyn_smem); // r_s[R], pbuf0[R], pbuf1[R]
  double *part_b = vec_s + (size_t)3 * R, *part_c = part_b + 32, *red = part_c + 32;
  double *sc = red + 32;                                     // [8]: alpha | beta, rr, stop, nan
  double *vals_s = red + 32 + 8;                             // [W][R]
  int *cols_s = reinterpret_cast<int *>(vals_s + (size_t)W * R);
  const int t = threadIdx.x;
  const int64_t base = (int64_t)rank * R;

  double xi[RPT], ri[RPT], pi[RPT], api[RPT];
#pragma unroll
  for (int k = 0; k < RPT; k++) {
    const int lt = t + k * TPB;
    const int64_t row = base + lt;
    const bool live = lt < R && row < A.n;
    xi[k] = live ? x[row] : 0.0;
    ri[k] = live ? r_g[row] : 0.0;
    pi[k] = 0.0;
    api[k] = 0.0;
    if (lt < R) {
      vec_s[lt] = ri[k];
      vec_s[R + lt] = 0.0;
      vec_s[2 * R + lt] = 0.0;
      const int b = live ? A.rowptr[row] : 0, len = live ? A.rowptr[row + 1] - b : 0;
      for (int j = 0; j < W; j++) {
        const size_t idx = (size_t)j * R + lt;
        if (j < len) {
          const int c = A.cols[b + j];
          cols_s[idx] = (c / R) << 16 | (c % R); // owner CTA | index in the owner's slice
          vals_s[idx] = A.vals[b + j];
        } else { // no entry: a harmless address (this row's own slot) with the sign bit set
          cols_s[idx] = (int)(0x80000000u | (unsigned)rank << 16 | (unsigned)lt);
          vals_s[idx] = 0.0;
        }
      }
    }
  }
  cl.sync();

  // K0 has armed the state; every thread keeps its own copy of the scalars
  double rr = st->rr, alpha = 0.0, beta = 0.0;
  const double tol = st->tol;
  const long long max_iter = st->max_iter;
  const int record = st->record_history;
  const long long hist_cap = st->hist_cap;
  long long it = 0;
  int done = st->done, nan_stop = 0, cur = 0;
  double stop_now = isnan(rr) ? 2.0 : (sqrt(rr) <= tol ? 1.0 : 0.0); // used by thread 0 only
  while (!done && it < max_iter) {
    const int off_old = R * (1 + cur);
    double *p_new_s = vec_s + (size_t)R * (1 + (cur ^ 1));
    double a = 0.0;
#pragma unroll
    for (int k = 0; k < RPT; k++) {
      const int lt = t + k * TPB;
      if (lt < R) {
        pi[k] = fma(beta, pi[k], ri[k]);
        p_new_s[lt] = pi[k];
      }
      api[k] = 0.0;
    }
    for (int j = 0; j < W; j++) { // the rows of a thread advance together: their gathers overlap
#pragma unroll
      for (int k = 0; k < RPT; k++) {
        const int lt = t + k * TPB;
        if (lt < R) {
          const int c = cols_s[(size_t)j * R + lt];
          const double *owner = cl.map_shared_rank(vec_s, (c >> 16) & 0x7fff);
          const int l = c & 0xffff;
          const double pj = fma(beta, owner[off_old + l], owner[l]); // p = r + beta p_old, as the owner rounds it
          const double s1 = fma(vals_s[(size_t)j * R + lt], pj, api[k]);
          api[k] = c < 0 ? api[k] : s1;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < RPT; k++) a = fma(pi[k], api[k], a);
    cluster_publish(cl, a, red, part_b, rank, ncta);
    if (t < 32) {
      const double pAp = cluster_total(part_b, ncta);
      if (t == 0) sc[0] = rr / pAp;
    }
    __syncthreads();
    alpha = sc[0];
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < RPT; k++) {
      const int lt = t + k * TPB;
      if (lt < R) {
        xi[k] = fma(alpha, pi[k], xi[k]);
        ri[k] = fma(-alpha, api[k], ri[k]);
        vec_s[lt] = ri[k];
        acc = fma(ri[k], ri[k], acc);
      }
    }
    cluster_publish(cl, acc, red, part_c, rank, ncta);
    if (t < 32) {
      const double rr_new = cluster_total(part_c, ncta);
      if (t == 0) {
        sc[1] = rr_new / rr;
        sc[2] = rr_new;
        sc[3] = stop_now; // the stop rule on the rr from BEFORE this trip's update (one-trip lag), formed a trip ago
      }
    }
    __syncthreads();
    beta = sc[1];
    rr = sc[2];
    if (sc[3] != 0.0) {
      done = 1;
      nan_stop = sc[3] > 1.5 ? 1 : 0;
    }
    // off the broadcast path: what the next trip will say about the rr just formed (the square root overlaps phase B)
    if (t == 0) stop_now = isnan(rr) ? 2.0 : (sqrt(rr) <= tol ? 1.0 : 0.0);
    it++;
    if (record && it < hist_cap && rank == 0 && t == 0) hist[it] = rr;
    cur ^= 1;
  }
#pragma unroll
  for (int k = 0; k < RPT; k++) {
    const int lt = t + k * TPB;
    const int64_t row = base + lt;
    if (lt < R && row < A.n) x[row] = xi[k];
  }
  if (rank == 0 && t == 0) {
    st->rr = rr;
    st->alpha = alpha;
    st->beta = beta;
    st->iter = it;
    st->done = done;
    st->nan_stop = nan_stop;
  }
  cl.sync(); // no CTA leaves while a neighbour may still read its shared memory
}

} // namespace cgb
