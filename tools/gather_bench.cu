// gather_bench.cu — what the L2 -> SM path of a B200 sustains on random 8-byte gathers.
//
// The SELL engine on BASELINE config C5 (random SPD, n = 10^7) is bound by exactly this: 197 M gathers of one double
// per SpMV out of a 40 MB column panel of p that lives in L2, each pulling a 32-byte sector across the L2 -> L1
// path.  This probe measures that ceiling in isolation: an index stream (coalesced, streaming loads, evict-first)
// and G independent gathers in flight per thread from a vector of `mb` MB, nothing else.
//   gather_bench [vector MB = 40] [gathers = 2^28] [reps = 5]
// prints one JSON object: gathers/s, the L2 -> SM sector rate (32 B per gather) and the DRAM-side index stream.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                  \
  do {                                                                                         \
    cudaError_t e_ = (x);                                                                      \
    if (e_ != cudaSuccess) {                                                                   \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                                 \
      return 1;                                                                                \
    }                                                                                          \
  } while (0)

__global__ void fill_idx(int *idx, int64_t n, uint32_t range) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t z = (uint64_t)(i + 1) * 0x9E3779B97F4A7C15ull;
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    idx[i] = (int)(z % range);
  }
}

template <int G>
__global__ void __launch_bounds__(256) gather_kernel(const int *__restrict__ idx, const double *__restrict__ x, int64_t n, double *out) {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i + (G - 1) * nthr < n; i += G * nthr) {
    int c[G];
#pragma unroll
    for (int g = 0; g < G; g++)
      asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(c[g]) : "l"(idx + i + g * nthr), "l"(pol));
    double v[G];
#pragma unroll
    for (int g = 0; g < G; g++) v[g] = __ldg(x + c[g]);
#pragma unroll
    for (int g = 0; g < G; g++) acc += v[g];
  }
  if (acc == 123.456) out[0] = acc; // keep the loads
}

template <int G>
static int run(const int *idx, const double *x, int64_t n, double *out, int sms, int reps, double *best_ms) {
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gather_kernel<G>, 256, 0));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  *best_ms = 1e30;
  for (int r = 0; r < reps + 1; r++) {
    CK(cudaEventRecord(a));
    gather_kernel<G><<<sms * occ, 256>>>(idx, x, n, out);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (r > 0 && ms < *best_ms) *best_ms = ms;
  }
  return 0;
}

int main(int argc, char **argv) {
  const double mb = argc > 1 ? atof(argv[1]) : 40.0;
  const int64_t n = argc > 2 ? atoll(argv[2]) : ((int64_t)1 << 28);
  const int reps = argc > 3 ? atoi(argv[3]) : 5;
  const uint32_t range = (uint32_t)(mb * 1e6 / 8.0);
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  int *idx = nullptr;
  double *x = nullptr, *out = nullptr;
  CK(cudaMalloc(&idx, sizeof(int) * (size_t)n));
  CK(cudaMalloc(&x, sizeof(double) * (size_t)range));
  CK(cudaMalloc(&out, 8));
  CK(cudaMemset(x, 0, sizeof(double) * (size_t)range));
  fill_idx<<<prop.multiProcessorCount * 8, 256>>>(idx, n, range);
  CK(cudaDeviceSynchronize());
  double ms4 = 0, ms8 = 0, ms16 = 0;
  if (run<4>(idx, x, n, out, prop.multiProcessorCount, reps, &ms4)) return 1;
  if (run<8>(idx, x, n, out, prop.multiProcessorCount, reps, &ms8)) return 1;
  if (run<16>(idx, x, n, out, prop.multiProcessorCount, reps, &ms16)) return 1;
  const double best = ms4 < ms8 ? (ms4 < ms16 ? ms4 : ms16) : (ms8 < ms16 ? ms8 : ms16);
  printf("{\"vector_MB\": %.1f, \"gathers\": %lld, \"ms_4_in_flight\": %.4f, \"ms_8_in_flight\": %.4f, \"ms_16_in_flight\": %.4f, "
         "\"gathers_per_s\": %.4e, \"l2_to_sm_sector_GBps\": %.1f, \"index_stream_GBps\": %.1f, \"device\": \"%s\"}\n",
         mb, (long long)n, ms4, ms8, ms16, (double)n / (best * 1e-3), 32.0 * (double)n / (best * 1e6), 4.0 * (double)n / (best * 1e6),
         prop.name);
  return 0;
}
