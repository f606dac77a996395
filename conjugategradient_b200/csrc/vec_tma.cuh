// vec_tma.cuh — the two vector passes of a CG trip with TMA-staged streaming, for kernels that run them on the SpMV's
// launch shape (dist_persistent.cuh: 3 CTAs of 288 threads per SM).
//
//   p-update   x += alpha p ; p = r + beta p          (reference src/CG.hpp:390, :418; x deferred by one trip)
//   r-update   r -= alpha Ap ; partial r.r            (:392, :406)
//
// With so few threads per SM a grid-stride loop of 16-byte loads does not keep enough bytes in flight (measured on
// C4 at 2 GPUs: both passes 10 % slower than the stand-alone kernels, which run 2,048 threads per SM; unrolling the
// loops made it worse).  Here the arrays move as 9 KB tiles: one thread issues 1-D bulk copies global -> shared for
// the next tile (completion on an mbarrier) while all threads update the current one in shared memory, and the
// results go back with bulk stores shared -> global.  Two stages; a stage is reloaded only after the bulk store
// that read it has finished reading (cp.async.bulk.wait_group.read).  Tiles are dealt round-robin over the CTAs
// (static: the order in which a thread adds its squares is fixed, results are bitwise reproducible).
//
// Proxy rules: the tiles are written with ordinary stores and read by the copy engine -> fence.proxy.async before
// the stores are issued; the arrays were written by other CTAs with ordinary stores in the previous phase and are
// read here by the copy engine -> the caller has passed a grid barrier (acquire) and vec_tma_begin() adds the proxy
// fence; the bulk stores are complete (wait_group 0) and fenced before the next grid barrier releases them.
#pragma once
#include "spmv_tma.cuh"

namespace cgb {

constexpr int kVecTile = 1152; // doubles per tile and array: 4 per thread of a 288-thread CTA, 9 KB

// dynamic shared memory of a pass over NARR arrays: two stages of NARR tiles + the two stage barriers
__host__ __device__ constexpr size_t vec_tma_smem_bytes(int narr) { return (size_t)2 * narr * kVecTile * sizeof(double) + 64; }

__device__ __forceinline__ void tma_load_1d_plain(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void proxy_fence_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// NARR arrays per stage, 2 stages, then the two stage barriers.  All threads of the CTA.
template <int NARR>
__device__ __forceinline__ uint64_t *vec_tma_begin(unsigned char *smem) {
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)2 * NARR * kVecTile * sizeof(double));
  if (threadIdx.x == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    proxy_fence_async(); // what other CTAs stored before the barrier we came through is visible to the copy engine
  }
  __syncthreads();
  return full;
}
__device__ __forceinline__ void vec_tma_end(uint64_t *full) {
  if (threadIdx.x == 0) {
    tma_store_wait_all(); // every bulk store of this CTA has been performed
    proxy_fence_async();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[0])) : "memory");
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[1])) : "memory");
  }
}

// x += alpha p ; p = r + beta p over [0, n).  r, p, x 16-byte aligned.
__device__ __forceinline__ void vec_tma_pupdate(int64_t n, const double *r, double *p, double *x, double alpha, double beta,
                                                unsigned char *smem) {
  constexpr int T = kVecTile;
  constexpr uint32_t TB = T * sizeof(double);
  const int bid = blockIdx.x, nb = gridDim.x, tid = threadIdx.x;
  const int64_t ntiles = n / T;
  uint64_t *full = vec_tma_begin<3>(smem);
  auto stage = [&](int s) { return reinterpret_cast<double *>(smem + (size_t)s * 3 * TB); };
  auto issue = [&](int64_t t, int s) {
    double *sb = stage(s);
    mbar_expect_tx(&full[s], 3 * TB);
    tma_load_1d_plain(sb, r + t * T, TB, &full[s]);
    tma_load_1d_plain(sb + T, p + t * T, TB, &full[s]);
    tma_load_1d_plain(sb + 2 * T, x + t * T, TB, &full[s]);
  };
  if (tid == 0 && bid < ntiles) issue(bid, 0);
  int s = 0;
  uint32_t phase[2] = {0u, 0u};
  for (int64_t t = bid; t < ntiles; t += nb, s ^= 1) {
    if (tid == 0 && t + nb < ntiles) {
      tma_store_wait_read(); // the store that read the other stage (previous tile) is done with it
      issue(t + nb, s ^ 1);
    }
    mbar_wait(&full[s], phase[s]);
    phase[s] ^= 1u;
    double2 *sr = reinterpret_cast<double2 *>(stage(s)), *sp = sr + T / 2, *sx = sp + T / 2;
#pragma unroll
    for (int k = tid; k < T / 2; k += kTmaThreads) {
      const double2 rv = sr[k];
      double2 pv = sp[k], xv = sx[k];
      xv.x = fma(alpha, pv.x, xv.x);
      xv.y = fma(alpha, pv.y, xv.y);
      pv.x = fma(beta, pv.x, rv.x);
      pv.y = fma(beta, pv.y, rv.y);
      sx[k] = xv;
      sp[k] = pv;
    }
    proxy_fence_async();
    __syncthreads();
    if (tid == 0) {
      tma_store_1d(p + t * T, sp, TB);
      tma_store_1d(x + t * T, sx, TB);
      tma_store_commit();
    }
  }
  vec_tma_end(full);
  // ragged tail (< one tile): plain code, spread over the grid
  const int64_t gt = (int64_t)bid * blockDim.x + tid, nthr = (int64_t)nb * blockDim.x;
  for (int64_t i = ntiles * T + gt; i < n; i += nthr) {
    const double po = p[i];
    x[i] = fma(alpha, po, x[i]);
    p[i] = fma(beta, po, r[i]);
  }
}

// r -= alpha Ap over [0, n); returns this thread's share of r.r.
__device__ __forceinline__ double vec_tma_update(int64_t n, double *r, const double *Ap, double alpha, unsigned char *smem) {
  constexpr int T = kVecTile;
  constexpr uint32_t TB = T * sizeof(double);
  const int bid = blockIdx.x, nb = gridDim.x, tid = threadIdx.x;
  const int64_t ntiles = n / T;
  uint64_t *full = vec_tma_begin<2>(smem);
  auto stage = [&](int s) { return reinterpret_cast<double *>(smem + (size_t)s * 2 * TB); };
  auto issue = [&](int64_t t, int s) {
    double *sb = stage(s);
    mbar_expect_tx(&full[s], 2 * TB);
    tma_load_1d_plain(sb, r + t * T, TB, &full[s]);
    tma_load_1d_plain(sb + T, Ap + t * T, TB, &full[s]);
  };
  if (tid == 0 && bid < ntiles) issue(bid, 0);
  int s = 0;
  uint32_t phase[2] = {0u, 0u};
  double acc = 0.0;
  for (int64_t t = bid; t < ntiles; t += nb, s ^= 1) {
    if (tid == 0 && t + nb < ntiles) {
      tma_store_wait_read();
      issue(t + nb, s ^ 1);
    }
    mbar_wait(&full[s], phase[s]);
    phase[s] ^= 1u;
    double2 *sr = reinterpret_cast<double2 *>(stage(s));
    const double2 *sq = sr + T / 2;
#pragma unroll
    for (int k = tid; k < T / 2; k += kTmaThreads) {
      double2 rv = sr[k];
      const double2 qv = sq[k];
      rv.x = fma(-alpha, qv.x, rv.x);
      rv.y = fma(-alpha, qv.y, rv.y);
      sr[k] = rv;
      acc = fma(rv.x, rv.x, acc);
      acc = fma(rv.y, rv.y, acc);
    }
    proxy_fence_async();
    __syncthreads();
    if (tid == 0) {
      tma_store_1d(r + t * T, sr, TB);
      tma_store_commit();
    }
  }
  vec_tma_end(full);
  const int64_t gt = (int64_t)bid * blockDim.x + tid, nthr = (int64_t)nb * blockDim.x;
  for (int64_t i = ntiles * T + gt; i < n; i += nthr) {
    const double rv = fma(-alpha, Ap[i], r[i]);
    r[i] = rv;
    acc = fma(rv, rv, acc);
  }
  return acc;
}

} // namespace cgb
