// static_solve.cu — cgb_cg_solve_static: the one-kernel CG loop (persistent.cuh) for callers that must not
// allocate.  Nothing in this file calls cudaMalloc, cudaFree, cudaMallocHost, an event or stream constructor, or a
// library routine that allocates.
//
// Who needs this.  cgb_cg_solve keeps its work vectors (r, two p buffers, A p), the solver state and the reduction
// partials in device memory that the matrix handle allocates on first use.  Some hosts of this library cannot do
// that:
//   * a simulation or framework that reserves (nearly) all of the GPU's memory for its own pool when it starts:
//     a later cudaMalloc from a plugin simply fails with out-of-memory;
//   * a time-step loop with a latency budget: cudaMalloc and cudaFree synchronise the whole device and take the
//     driver's allocation lock, so a solve that allocates stalls every other stream of the process;
//   * deployments whose rules forbid driver allocations after start-up (fixed memory budgets per client under
//     MPS, qualified real-time loops).
// For them the work area is part of the library image: a file-scope __device__ array, present on every GPU that
// loads libcgb200.so from the moment the module is loaded until the process exits (CGB_STATIC_SCRATCH_BYTES,
// about 4.3 MB per GPU), sized for systems of up to CGB_STATIC_MAX_ROWS rows -- the range in which cgb_cg_solve
// itself picks the one-kernel loop.  The caller passes only its inputs (CSR arrays, b) and its output (x).
//
// Arithmetic: the same kernels as cgb_cg_solve with CGB_LOOP_PERSISTENT (cg_init_kernel, cg_persistent_kernel),
// always on the direct-load row engine; stop rule and trip cap are the reference's (src/CG.hpp:396-404, :436).
#include <algorithm>
#include <mutex>

#include "launch.cuh"
#include "persistent.cuh"

using namespace cgb;

namespace {

constexpr int64_t kStaticMaxRows = CGB_STATIC_MAX_ROWS;
// layout of the work area, in doubles; every part starts on a 256-byte boundary
constexpr int64_t kOffState = 0;                                  // CgState, padded to 256 bytes
constexpr int64_t kOffPartials = 32;                              // 2 * kMaxPartials
constexpr int64_t kOffR = kOffPartials + 2 * (int64_t)kMaxPartials;
constexpr int64_t kOffP0 = kOffR + kStaticMaxRows;
constexpr int64_t kOffP1 = kOffP0 + kStaticMaxRows;
constexpr int64_t kOffAp = kOffP1 + kStaticMaxRows;
constexpr int64_t kScratchDoubles = kOffAp + kStaticMaxRows;
static_assert(sizeof(CgState) <= 32 * sizeof(double), "CgState must fit its 256-byte slot");
static_assert(kStaticMaxRows % 32 == 0, "parts of the work area stay 256-byte aligned");
static_assert(kScratchDoubles * sizeof(double) == CGB_STATIC_SCRATCH_BYTES, "header and layout disagree");

// The work area.  In HOST code this name is only a handle for cudaGetSymbolAddress: it is not a device address,
// and a kernel handed `g_static_scratch` from the host would fault (or, where the GPU can reach host memory,
// quietly use the wrong memory).  The address every launch below uses comes from static_setup().
static __device__ __align__(256) double g_static_scratch[kScratchDoubles];

struct DeviceSlot {
  bool ready = false;
  double *base = nullptr; // device address of g_static_scratch on this device
  int sm_count = 0;
  int coop = 0;
};
constexpr int kMaxDevices = 64;
DeviceSlot g_slots[kMaxDevices];
std::mutex g_setup_mutex;

// Once per device, with that device current: resolve the device address of the work area and read the two
// device attributes the launch needs.  Allocates nothing.
int static_setup(int dev, DeviceSlot **out) {
  if (dev < 0 || dev >= kMaxDevices) return fail(CGB_ERR_INVALID, "cgb_cg_solve_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  DeviceSlot &s = g_slots[dev];
  if (!s.ready) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
    if (e != cudaSuccess || !p)
      return fail(CGB_ERR_CUDA, "cgb_cg_solve_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
    int sms = 0, coop = 0;
    CGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CGB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    s.base = p;
    s.sm_count = sms;
    s.coop = coop;
    s.ready = true;
  }
  *out = &s;
  return CGB_OK;
}

// lanes per row of the direct-load engine: the rule of csr_plan (csr.cu)
int lanes_for(int64_t n, int64_t nnz) {
  const double mean = n > 0 ? (double)nnz / (double)n : 1.0;
  int lanes = 2;
  while (lanes < 32 && (double)lanes < mean) lanes <<= 1;
  return lanes;
}

} // namespace

extern "C" int cgb_cg_solve_static(void *stream, int64_t n, int64_t nnz, const double *d_vals, const int32_t *d_cols,
                                   const int32_t *d_rowptr, const double *d_b, double *d_x, double abs_tol,
                                   int64_t max_iter, cgb_cg_stats *stats) {
  if (!d_rowptr || (nnz > 0 && (!d_vals || !d_cols))) return fail(CGB_ERR_STATE, "No Matrix given"); // src/CG.hpp:270-272
  if (!d_b) return fail(CGB_ERR_STATE, "No right hand side to solve for");                              // src/CG.hpp:266-268
  CGB_REQUIRE(d_x, "cgb_cg_solve_static: x is NULL");
  CGB_REQUIRE(n >= 0 && nnz >= 0 && nnz <= 0x7fffffffLL, "cgb_cg_solve_static: bad size");
  if (n > kStaticMaxRows)
    return fail(CGB_ERR_UNSUPPORTED,
                "cgb_cg_solve_static: %lld rows do not fit the built-in work area (CGB_STATIC_MAX_ROWS = %lld); "
                "use cgb_cg_solve, which allocates its work vectors",
                (long long)n, (long long)kStaticMaxRows);
  CGB_REQUIRE(((uintptr_t)d_x & 15) == 0, "cgb_cg_solve_static: x must be 16-byte aligned");
  if (stats) memset(stats, 0, sizeof(*stats));
  if (max_iter < 0) max_iter = n + 1; // reference cap (src/CG.hpp:436)
  if (n == 0) return CGB_OK;

  int dev = -1;
  CGB_CUDA(cudaGetDevice(&dev));
  DeviceSlot *slot = nullptr;
  CGB_TRY(static_setup(dev, &slot)); // first call on this device: resolves the work area's address
  if (!slot->coop) return fail(CGB_ERR_UNSUPPORTED, "cgb_cg_solve_static: device %d has no cooperative launch", dev);
  cudaStream_t s = (cudaStream_t)stream;
  double *base = slot->base;
  CgState *st = reinterpret_cast<CgState *>(base + kOffState);
  double *partials = base + kOffPartials, *r = base + kOffR, *p0 = base + kOffP0, *p1 = base + kOffP1, *Ap = base + kOffAp;

  // The array keeps whatever the last call left.  What must start from a known value: the state (its ticket
  // counter in particular: a call that failed half-way may have left it armed) and both p buffers (the first
  // trip reads p_old with beta = 0, and 0 * NaN is NaN).  r, A p and the partials are written before they are read.
  CgState init;
  memset(&init, 0, sizeof(init));
  init.tol = abs_tol;
  init.max_iter = max_iter;
  init.done = 1; // K0 arms the loop
  CGB_CUDA(cudaMemcpyAsync(st, &init, sizeof(CgState), cudaMemcpyHostToDevice, s)); // pageable source: staged before the call returns
  CGB_CUDA(cudaMemsetAsync(p0, 0, sizeof(double) * (size_t)n, s));
  CGB_CUDA(cudaMemsetAsync(p1, 0, sizeof(double) * (size_t)n, s));

  CsrView V{};
  V.n = n;
  V.nnz = nnz;
  V.rowptr = d_rowptr;
  V.cols = d_cols;
  V.vals = d_vals;
  const int lanes = lanes_for(n, nnz);
  const int64_t rows_per_cta = (int64_t)(kBlock / lanes) * 2;
  const int64_t need = (n + rows_per_cta - 1) / rows_per_cta;
  int rc = CGB_OK;
  dispatch_lanes(lanes, kEngVec, [&](auto Lc, auto) {
    constexpr int L = decltype(Lc)::value;
    // K0: r = b - A x, rr = r.r (src/CG.hpp:303-357)
    const int grid0 = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)slot->sm_count * 8, need));
    cg_init_kernel<L, kEngVec><<<grid0, kBlock, 0, s>>>(V, (const double *)d_x, d_b, r, st, partials, (double *)nullptr);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      rc = fail(CGB_ERR_CUDA, "cgb_cg_solve_static: cg_init_kernel: %s", cudaGetErrorString(e));
      return;
    }
    auto kernel = cg_persistent_kernel<L>;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kBlock, 0) != cudaSuccess || occ < 1) {
      rc = fail(CGB_ERR_CUDA, "cgb_cg_solve_static: occupancy query failed");
      return;
    }
    int grid = (int)std::min<int64_t>((int64_t)occ * slot->sm_count, need); // all CTAs must be co-resident
    grid = std::max(1, std::min(grid, kMaxPartials));
    double *xp = d_x, *hist = nullptr;
    void *args[] = {&V, &xp, &r, &p0, &p1, &Ap, &st, &partials, &hist};
    e = cudaLaunchCooperativeKernel((void *)kernel, dim3(grid), dim3(kBlock), args, 0, s);
    if (e != cudaSuccess) rc = fail(CGB_ERR_CUDA, "cgb_cg_solve_static: cudaLaunchCooperativeKernel: %s", cudaGetErrorString(e));
  });
  CGB_TRY(rc);
  CgState fin;
  CGB_CUDA(cudaMemcpyAsync(&fin, st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  if (stats) {
    stats->iterations = fin.iter;
    stats->rr = fin.rr;
    stats->rr0 = fin.rr0;
    stats->converged = fin.done;
    stats->nan_stop = fin.nan_stop;
    stats->kernel_launches = 2;
    stats->loop_mode = CGB_LOOP_PERSISTENT;
    stats->format = CGB_FMT_CSR;
  }
  return CGB_OK;
}
