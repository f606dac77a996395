// cluster_sync_bench.cu — what the pieces of the cluster-resident CG loop's reductions cost on this GPU (cycles,
// clock64 in thread 0 of CTA 0, mean over 2,000 repetitions).  Build and run:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/cluster_sync_bench tools/cluster_sync_bench.cu && build/cluster_sync_bench
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>

namespace cg = cooperative_groups;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, int rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}

constexpr int kReps = 2000;

// mode 0: cluster.sync() only          1: __syncthreads() only           2: warp_sum (5 dependent shuffle+add)
//      3: fp64 division                4: block tree (warp_sum, smem, sync, warp_sum)
//      5: full exchange with barrier.cluster (what cluster.cuh does)
//      6: full exchange with an mbarrier, only warp 0 waits, the CTA follows through __syncthreads
__global__ void bench_kernel(int mode, long long *out, double *sink) {
  cg::cluster_group cl = cg::this_cluster();
  __shared__ double red[32], part[32], sc[2];
  __shared__ uint64_t bar;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, rank = (int)cl.block_rank(), ncta = (int)cl.num_blocks();
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(ncta));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cl.sync();
  double v = 1.0 + t * 1e-3, acc = 0.0;
  const long long t0 = clock64();
  for (int it = 0; it < kReps; it++) {
    if (mode == 0) {
      cl.sync();
    } else if (mode == 1) {
      __syncthreads();
    } else if (mode == 2) {
      v = warp_sum(v) + 1.0;
    } else if (mode == 3) {
      v = 1.0 + 1.0 / (v + 3.0);
    } else if (mode == 4) {
      double s = warp_sum(v);
      if (lane == 0) red[warp] = s;
      __syncthreads();
      if (warp == 0) {
        s = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.0;
        s = warp_sum(s);
        if (lane == 0) sc[0] = s;
      }
      __syncthreads();
      v = sc[0] * 1e-9 + 1.0;
    } else if (mode == 5) {
      double s = warp_sum(v);
      if (lane == 0) red[warp] = s;
      __syncthreads();
      if (warp == 0) {
        s = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.0;
        s = warp_sum(s);
        s = __shfl_sync(0xffffffffu, s, 0);
        if (lane < ncta) cl.map_shared_rank(part, lane)[rank] = s;
      }
      cl.sync();
      if (t < 32) {
        double x = lane < ncta ? part[lane] : 0.0;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (t == 0) sc[0] = 1.0 / x;
      }
      __syncthreads();
      v = sc[0] + 1.0;
    } else {
      double s = warp_sum(v);
      if (lane == 0) red[warp] = s;
      __syncthreads();
      if (warp == 0) {
        s = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.0;
        s = warp_sum(s);
        s = __shfl_sync(0xffffffffu, s, 0);
        if (lane < ncta) {
          const uint32_t slot = mapa(smem_u32(part + rank), lane), rbar = mapa(smem_u32(&bar), lane);
          asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(slot), "d"(s) : "memory");
          asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(rbar) : "memory");
        }
        uint32_t ok;
        do {
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                       : "=r"(ok)
                       : "r"(smem_u32(&bar)), "r"((uint32_t)(it & 1))
                       : "memory");
        } while (!ok);
        double x = lane < ncta ? part[lane] : 0.0;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (t == 0) sc[0] = 1.0 / x;
      }
      __syncthreads();
      v = sc[0] + 1.0;
      // the slots are rewritten next round: nobody may run ahead by more than the barrier allows (two phases)
    }
    acc += v;
  }
  const long long t1 = clock64();
  if (rank == 0 && t == 0) out[0] = (t1 - t0) / kReps;
  if (acc == 123.456) sink[0] = acc;
  cl.sync();
}

int main() {
  long long *out;
  double *sink;
  cudaMallocManaged(&out, sizeof(long long));
  cudaMalloc(&sink, sizeof(double));
  const char *names[] = {"cluster.sync", "__syncthreads", "warp_sum (fp64, 5 levels)", "fp64 division (+2 adds)", "block tree (2 x warp_sum, 2 x sync)",
                         "exchange, barrier.cluster", "exchange, mbarrier (warp 0 waits)"};
  cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int ncta : {16, 8}) {
    for (int threads : {1024, 256, 32}) {
      for (int mode = 0; mode < 7; mode++) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ncta);
        cfg.blockDim = dim3(threads);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = ncta;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, bench_kernel, mode, out, sink);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf("cluster %2d x %4d threads  %-38s  failed: %s\n", ncta, threads, names[mode], cudaGetErrorString(e));
          cudaGetLastError();
          continue;
        }
        printf("cluster %2d x %4d threads  %-38s  %6lld cycles\n", ncta, threads, names[mode], out[0]);
      }
    }
  }
  return 0;
}
