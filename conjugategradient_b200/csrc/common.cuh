// common.cuh — shared host/device helpers of the cgb200 library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>

#include "../../include/cgb200.h"

namespace cgb {

// ---- error plumbing: nothing throws across the C boundary --------------------------------
void set_error(const char *fmt, ...);
int fail(int code, const char *fmt, ...);

#define CGB_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess)                                                                      \
      return ::cgb::fail(CGB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),  \
                         __FILE__, __LINE__);                                                   \
  } while (0)

#define CGB_TRY(expr)                 \
  do {                                \
    int _rc = (expr);                 \
    if (_rc != CGB_OK) return _rc;    \
  } while (0)

#define CGB_REQUIRE(cond, msg)                                       \
  do {                                                               \
    if (!(cond)) return ::cgb::fail(CGB_ERR_INVALID, "%s", msg);    \
  } while (0)

constexpr int kEngVec = 0, kEngTma = 1, kEngSell = 2, kEngDict = 3; // row engines of the SpMV-shaped kernels
constexpr int kSellLongRow = 256; // rows longer than this stay out of the SELL slices (sell.cu)
constexpr int kBlock = 256;        // threads per CTA for every streaming kernel
constexpr int kMaxPartials = 4096; // upper bound on CTAs that take part in a reduction

// Device-resident scalars of one CG solve.  One instance per workspace; kernels read it in
// their prologue, the last CTA of a reducing kernel (ticket pattern) updates it.
struct CgState {
  double rr;     // r.r at the start of the current iteration   (reference: rxr)
  double pAp;    // p.Ap of the current iteration               (reference: value2)
  double alpha;  // rr / pAp                                    (src/CG.hpp:385-386)
  double beta;   // rr_new / rr                                 (src/CG.hpp:413-416)
  double tol;    // absolute tolerance                          (reference: acc)
  double rr0;    // r0.r0
  double res_rr; // cgb_residual outputs
  double res_xx;
  long long iter;     // loop trips executed so far
  long long max_iter; // cap (reference: N+1)
  int done;           // stop test fired (reference: *is_done)
  int nan_stop;
  unsigned int ticket; // last-block counter, always returned to 0 by the last block
  int record_history;
  long long hist_cap;
  int comm_error; // multi-GPU: a peer wait timed out (the loop is stopped, the host reports it)
  int precond;    // 1: Jacobi-preconditioned loop (rho = r.z drives alpha and beta, rr only the stop rule)
  double rho;     // r.z of the current iteration (precond only)
};

struct Workspace;

} // namespace cgb

struct cgb_ctx {
  int device = 0;
  int sm_count = 0;
  int coop_ok = 0;
  size_t l2_bytes = 0;
  size_t l2_persist_max = 0;
  size_t l2_window_max = 0;   // largest access-policy window the device accepts
  size_t l2_persist_set = 0;  // persisting carve-out currently configured (cudaLimitPersistingL2CacheSize)
  cudaStream_t stream = nullptr;
  // scratch for the per-op reductions (lazy)
  double *partials = nullptr;
  unsigned int *ticket = nullptr;
  double *pinned = nullptr; // small pinned bounce buffer (scalars)
  cgb_progress_fn progress = nullptr; // called after every state poll of a solve (cgb_ctx_set_progress)
  void *progress_user = nullptr;
};

struct cgb_event {
  cudaEvent_t ev = nullptr;
};

namespace cgb {

struct ScopedDevice {
  int prev = -1;
  explicit ScopedDevice(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~ScopedDevice() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int ctx_scratch(cgb_ctx *ctx); // allocates ctx->partials / ticket on first use

// ---- device-memory cache (core.cu) -------------------------------------------------------------------------
// Every cudaMalloc / cudaFree of the library goes through a per-device cache of whole allocations keyed by size:
// a freed block is kept and handed out again for a request of (nearly) the same size.  The end-to-end paths
// (cgb_cg_solve_host, cgb_dist_set_matrix: a matrix, its work vectors and their scratch per call) then cost no
// driver allocation after the first call — cudaFree of a multi-GB block alone is milliseconds.  Semantics are
// cudaFree's: the device is synchronised before a block changes hands.  Blocks are never split, so a cached
// block is still a legal argument of cudaIpcGetMemHandle.  CGB_ALLOC_CACHE=0 disables it, CGB_ALLOC_CACHE_MAX_GB
// bounds what is kept (default 64), cgb_trim() returns everything to the driver; an allocation the driver refuses
// is retried once after a trim.
cudaError_t cached_malloc(void **p, size_t bytes);
cudaError_t cached_free(void *p);
cudaError_t cached_host_alloc(void **p, size_t bytes);
cudaError_t cached_host_free(void *p);
void cache_trim();
// host <-> device copies on the context's stream; pageable host memory is staged through pinned buffers (core.cu)
cudaError_t copy_h2d(cgb_ctx *ctx, void *d_dst, const void *h_src, size_t bytes);
cudaError_t copy_d2h(cgb_ctx *ctx, void *h_dst, const void *d_src, size_t bytes);

#ifdef __CUDACC__
// ---- device helpers ------------------------------------------------------------------------
// Streaming (read-once) loads: bypass L1 allocation so the gathered vector keeps the cache.
__device__ __forceinline__ double ld_stream_f64(const double *p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int ld_stream_s32(const int *p) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_stream_f64x2(const double *p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ int4 ld_stream_s32x4(const int *p) {
  int4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the CTA in a fixed order; result valid in thread 0.  smem: >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double *smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads(); // protect smem reuse across consecutive calls
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    const int nwarps = (blockDim.x + 31) >> 5;
    v = lane < nwarps ? smem[lane] : 0.0;
    v = warp_sum(v);
  }
  return v;
}

// Fixed-order sum of `count` per-CTA partials by one CTA; result valid in thread 0.
__device__ __forceinline__ double reduce_partials(const volatile double *partials, int count, double *smem) {
  double s = 0.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) s += partials[i];
  return block_sum(s, smem);
}

// Ticket pattern: publish this CTA's partial(s); returns true in ALL threads of the CTA that
// arrives last (its view of every partial is then complete).
__device__ __forceinline__ bool publish_and_elect(unsigned int *ticket, int nblocks) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == (unsigned int)(nblocks - 1));
    if (s_last) *ticket = 0u; // re-arm for the next launch
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0;
}
#endif // __CUDACC__

} // namespace cgb

#ifndef CGB_NO_ALLOC_MACROS
#define cudaMalloc(p, bytes) ::cgb::cached_malloc((void **)(p), (bytes))
#define cudaFree(p) ::cgb::cached_free((void *)(p))
#endif
