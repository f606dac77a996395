// spmv_tma.cuh — CSR SpMV with TMA-staged streaming (sm_100a).
//
// The matrix stream (values, column indices, row offsets) is the HBM-bound part of CG.  Here it is
// moved by the TMA engine, not by the SMs: a producer warp issues 1-D bulk copies
// (cp.async.bulk.shared::cluster.global, completion on an mbarrier) of one row tile per pipeline
// stage, STAGES tiles ahead; 256 consumer threads read the staged tile from shared memory, gather
// the vector through L1/L2 and form the row sums.  A tile is TILE_ROWS = 256 / L consecutive rows
// (L lanes per row), so its non-zeros are one contiguous range of the CSR arrays and no tile
// descriptors are needed.  Tiles that do not fit a stage (very long rows), the last tile and
// ranges that would read past the arrays fall back to direct loads inside the same kernel.
//
// Same gather / epilogue functors as spmv.cuh, so every SpMV-shaped pass of the path
// (reference src/VectorOperations.hpp:455-462, src/CG.hpp:324-341, :374-379, :477-499) uses it.
#pragma once
#include "spmv.cuh"

namespace cgb {

constexpr int kTmaConsumers = 256;
constexpr int kTmaThreads = kTmaConsumers + 32; // + one producer warp
constexpr int kTmaStages = 3;
constexpr int kTmaCap = 1856; // non-zeros per stage (incl. alignment slack)

struct TmaStageMeta {
  int j0a;      // first staged non-zero (aligned down to 4)
  int fallback; // 1: tile was not staged, consumers load directly
};

// dynamic shared memory layout per stage: vals[CAP] | cols[CAP] | rowptr[TILE_ROWS + 4]
__host__ __device__ constexpr size_t tma_stage_bytes(int tile_rows) {
  return (size_t)kTmaCap * 8 + (size_t)kTmaCap * 4 + (size_t)(tile_rows + 4) * 4;
}
__host__ __device__ constexpr size_t tma_smem_bytes(int tile_rows) {
  return kTmaStages * tma_stage_bytes(tile_rows) + 2 * kTmaStages * 8 + kTmaStages * sizeof(TmaStageMeta) + 64;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// The matrix stream is read exactly once per pass: evict-first in L2, so that it does not push the
// vectors (which are re-read by the next kernel, and gathered several times by this one) out.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// After spmv_rows_tma in a kernel that calls it more than once (persistent loops): all threads of the CTA; the stage
// barriers are invalidated so the next call may initialise them again (an mbarrier must be invalidated before its
// memory is re-initialised).
__device__ __forceinline__ void spmv_tma_teardown(unsigned char *smem_raw, int tile_rows) {
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + kTmaStages * tma_stage_bytes(tile_rows));
  __syncthreads(); // every staged tile has been consumed: no copy in flight, no arrival pending on `full`
  if (threadIdx.x == 0)
    for (int s = 0; s < 2 * kTmaStages; s++) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&bars[s])) : "memory");
}

// Row sums of `len` staged entries for one lane of an L-lane group; 4 gathers in flight.
template <int L, class G>
__device__ __forceinline__ double row_sum_staged(const double *sv, const int *sc, int s, int e, int sub, const G &g) {
  double sum = 0.0;
  for (int j = s + sub; j < e; j += 4 * L) {
    const int j1 = j + L, j2 = j + 2 * L, j3 = j + 3 * L;
    const int c0 = sc[j];
    const int c1 = j1 < e ? sc[j1] : c0, c2 = j2 < e ? sc[j2] : c0, c3 = j3 < e ? sc[j3] : c0;
    const double v0 = sv[j];
    const double v1 = j1 < e ? sv[j1] : 0.0, v2 = j2 < e ? sv[j2] : 0.0, v3 = j3 < e ? sv[j3] : 0.0;
    const double x0 = g(c0), x1 = g(c1), x2 = g(c2), x3 = g(c3);
    sum = fma(v0, x0, sum);
    sum = fma(v1, x1, sum);
    sum = fma(v2, x2, sum);
    sum = fma(v3, x3, sum);
  }
  return sum;
}

template <int L, class G>
__device__ __forceinline__ double row_sum_direct(const double *__restrict__ vals, const int *__restrict__ cols, int s,
                                                 int e, int sub, const G &g) {
  double sum = 0.0;
  for (int j = s + sub; j < e; j += L) sum = fma(ld_stream_f64(vals + j), g(ld_stream_s32(cols + j)), sum);
  return sum;
}

struct TmaNoSideJob {
  __device__ __forceinline__ void operator()() const {}
};

// Must be called by all kTmaThreads threads of the CTA.  `smem_raw` is the dynamic shared memory.
// `side` runs once per CTA on the producer lane, after it has put the first kTmaStages tiles in flight: the persistent
// multi-GPU loop hides the system-scope fence and the flag stores of its halo pack there.
template <int L, class G, class E, class SIDE = TmaNoSideJob>
__device__ __forceinline__ void spmv_rows_tma(int64_t n, int64_t nnz, const int *__restrict__ rowptr,
                                              const int *__restrict__ cols, const double *__restrict__ vals,
                                              const G &g, const E &e, int bid, int nb, double &a0, double &a1,
                                              unsigned char *smem_raw, const SIDE &side = SIDE()) {
  constexpr int TILE_ROWS = kTmaConsumers / L;
  constexpr size_t STAGE = tma_stage_bytes(TILE_ROWS);
  unsigned char *base = smem_raw;
  uint64_t *full = reinterpret_cast<uint64_t *>(base + kTmaStages * STAGE);
  uint64_t *empty = full + kTmaStages;
  TmaStageMeta *meta = reinterpret_cast<TmaStageMeta *>(empty + kTmaStages);
  const int tid = threadIdx.x;
  const int64_t ntiles = (n + TILE_ROWS - 1) / TILE_ROWS;

  if (tid == 0) {
    for (int s = 0; s < kTmaStages; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kTmaConsumers / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  if (tid >= kTmaConsumers) {
    // ---------------- producer warp: one lane drives the TMA ----------------
    // The side job runs on this same lane once the pipeline is primed (kTmaStages tiles in flight), so its latency
    // (a system-scope fence and a ticket in the multi-GPU pack) lies behind the consumers' work on those tiles.  On
    // another lane of this warp it would run first: the two branches of a warp do not overlap, and the fence would
    // sit in front of tile 0.
    if (tid == kTmaConsumers) {
      int primed = 0;
      const uint64_t pol = l2_policy_evict_first();
      int stage = 0;
      uint32_t phase = 0;
      // tile bounds are fetched one tile ahead so their latency never sits between two copies
      auto staged_ok = [&](int64_t t) { return t < ntiles - 1 && t * TILE_ROWS + TILE_ROWS + 4 <= n + 1; };
      int nj0 = 0, nj1 = 0;
      if (bid < ntiles && staged_ok(bid)) {
        nj0 = __ldg(rowptr + (int64_t)bid * TILE_ROWS);
        nj1 = __ldg(rowptr + (int64_t)bid * TILE_ROWS + TILE_ROWS);
      }
      for (int64_t t = bid; t < ntiles; t += nb) {
        const int j0 = nj0, j1 = nj1;
        const int64_t tn = t + nb;
        if (tn < ntiles && staged_ok(tn)) {
          nj0 = __ldg(rowptr + tn * TILE_ROWS);
          nj1 = __ldg(rowptr + tn * TILE_ROWS + TILE_ROWS);
        }
        mbar_wait(&empty[stage], phase ^ 1); // stage free (first pass falls through)
        unsigned char *sb = base + stage * STAGE;
        double *sv = reinterpret_cast<double *>(sb);
        int *sc = reinterpret_cast<int *>(sb + (size_t)kTmaCap * 8);
        int *sr = sc + kTmaCap;
        const int64_t r0 = t * TILE_ROWS;
        int fb = staged_ok(t) ? 0 : 1;
        int j0a = 0;
        if (!fb) {
          j0a = j0 & ~3;
          const int j1a = (j1 + 3) & ~3;
          if ((int64_t)j1a > nnz || j1a - j0a > kTmaCap) {
            fb = 1;
          } else {
            const uint32_t cnt = (uint32_t)(j1a - j0a);
            const uint32_t bytes = cnt * 12u + (uint32_t)(TILE_ROWS + 4) * 4u;
            meta[stage].j0a = j0a;
            meta[stage].fallback = 0;
            mbar_expect_tx(&full[stage], bytes);
            if (cnt) {
              tma_load_1d(sv, vals + j0a, cnt * 8u, &full[stage], pol);
              tma_load_1d(sc, cols + j0a, cnt * 4u, &full[stage], pol);
            }
            tma_load_1d(sr, rowptr + r0, (uint32_t)(TILE_ROWS + 4) * 4u, &full[stage], pol);
          }
        }
        if (fb) {
          meta[stage].j0a = 0;
          meta[stage].fallback = 1;
          mbar_arrive(&full[stage]);
        }
        if (++stage == kTmaStages) {
          stage = 0;
          phase ^= 1;
        }
        if (++primed == kTmaStages) side();
      }
      if (primed < kTmaStages) side(); // fewer tiles than stages (or none) for this CTA
    }
  } else {
    // ---------------- consumers ----------------
    const int sub = tid % L;
    const int lrow = tid / L;
    int stage = 0;
    uint32_t phase = 0;
    for (int64_t t = bid; t < ntiles; t += nb) {
      mbar_wait(&full[stage], phase);
      unsigned char *sb = base + stage * STAGE;
      const double *sv = reinterpret_cast<const double *>(sb);
      const int *sc = reinterpret_cast<const int *>(sb + (size_t)kTmaCap * 8);
      const int *sr = sc + kTmaCap;
      const int64_t row = t * TILE_ROWS + lrow;
      double sum = 0.0;
      if (!meta[stage].fallback) {
        const int j0a = meta[stage].j0a;
        const int s = sr[lrow] - j0a, en = sr[lrow + 1] - j0a;
        sum = row_sum_staged<L>(sv, sc, s, en, sub, g);
      } else if (row < n) {
        const int s = __ldg(rowptr + row), en = __ldg(rowptr + row + 1);
        sum = row_sum_direct<L>(vals, cols, s, en, sub, g);
      }
      // the stage can be refilled as soon as this warp's reads are done
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(&empty[stage]);
#pragma unroll
      for (int o = L >> 1; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o, L);
      if (sub == 0 && row < n) e.row(row, sum, a0, a1);
      if (++stage == kTmaStages) {
        stage = 0;
        phase ^= 1;
      }
    }
  }
}

} // namespace cgb
