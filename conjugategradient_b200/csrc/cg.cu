// cg.cu — the CG solve: host orchestration of the gated iteration kernels.
// Interface replaced: CGSolver::CG::solve (reference src/CG.hpp:255-454).  The reference submits
// 12 device operations and one wait_and_throw per iteration; here a batch of iterations is one
// CUDA-graph launch and the stop flag lives on the device, polled once per batch.
#include "launch.cuh"
#include "cluster.cuh"
#include "persistent.cuh"

#include <algorithm>
#include <vector>

using namespace cgb;

namespace cgb {

// L2 persisting access-policy window over the gathered vector p (north_star):
// the matrix stream is evict-first, vector passes are streaming, the window keeps p's lines resident
// between the p-update that writes them and the SpMV gathers that read them (several times each).
static cudaAccessPolicyWindow p_window(cgb_ctx *ctx, const double *p, int64_t n) {
  cudaAccessPolicyWindow w;
  memset(&w, 0, sizeof(w));
  size_t bytes = sizeof(double) * (size_t)n;
  if (bytes > ctx->l2_window_max) bytes = ctx->l2_window_max;
  w.base_ptr = (void *)p;
  w.num_bytes = bytes;
  const double ratio = bytes ? (double)ctx->l2_persist_set / (double)bytes : 0.0;
  w.hitRatio = (float)(ratio > 1.0 ? 1.0 : ratio);
  if (const char *e = getenv("CGB_L2_HITRATIO")) w.hitRatio = (float)std::min(1.0, std::max(0.0, atof(e))); // A/B runs (profiles/r02_l2window.md)
  w.hitProp = cudaAccessPropertyPersisting;
  w.missProp = cudaAccessPropertyStreaming;
  return w;
}

struct LoopCfg {
  bool fused;
  bool l2_window;
  bool time_kernels;
  SpmvLaunch sl;
  int grid_vec;
  int grid_pup;
  bool jacobi;     // Jacobi-preconditioned variants of the vector kernels
  bool replace;    // residual replacement at the end of every batch
  const double *b; // right-hand side (replacement only)
  bool vec_tma;    // TMA-staged p-update / r-update kernels (vec_tma.cuh)
};

// Enqueue one CG iteration on `s`.  When `timed`, bracket the kernels with external event records
// (legal inside stream capture; they become event-record nodes of the graph).
static int enqueue_iteration(cgb_csr *A, double *x, const LoopCfg &c, cudaStream_t s, bool timed) {
  Workspace &w = A->ws;
  const CsrView V = csr_view_final(A, w.Ap);
  auto mark = [&](int i) -> cudaError_t {
    return timed ? cudaEventRecordWithFlags(w.ev[i], s, cudaEventRecordExternal) : cudaSuccess;
  };
  CGB_CUDA(mark(0));
  if (c.jacobi) cg_pupdate_jacobi_kernel<<<c.grid_pup, kBlock, 0, s>>>(A->n, w.r, w.dinv, w.p0, x, w.st);
  else if (!c.fused && c.vec_tma)
    cg_pupdate_tma_kernel<<<A->grid_pup_tma, kTmaThreads, vec_tma_smem_bytes(3), s>>>(A->n, w.r, w.p0, x, w.st);
  else if (!c.fused) cg_pupdate_kernel<<<c.grid_pup, kBlock, 0, s>>>(A->n, w.r, w.p0, x, w.st);
  CGB_CUDA(mark(1));
  launch_panel_partials(A, c.sl, s, w.p0, w.Ap, w.st); // column panels 0 .. P-2: Ap accumulates
  dispatch_lanes(c.sl.lanes, c.sl.eng, [&](auto Lc, auto Tc) {
    constexpr int L = decltype(Lc)::value;
    constexpr int T = decltype(Tc)::value;
    if (c.fused)
      launch_kernel(cg_spmv_kernel<L, T, true>, c.sl, s, V, (const double *)w.r, w.p0, w.p1, w.Ap, x, w.st, w.partials,
                    (double *)nullptr, (const double *)nullptr);
    else
      launch_kernel(cg_spmv_kernel<L, T, false>, c.sl, s, V, (const double *)w.r, w.p0, w.p1, w.Ap, x, w.st, w.partials,
                    (double *)nullptr, (const double *)nullptr);
  });
  CGB_CUDA(mark(2));
  if (c.jacobi) cg_update_jacobi_kernel<<<c.grid_vec, kBlock, 0, s>>>(A->n, w.r, w.Ap, w.dinv, w.st, w.partials, w.hist);
  else if (c.vec_tma)
    cg_update_tma_kernel<<<A->grid_vec_tma, kTmaThreads, vec_tma_smem_bytes(2), s>>>(A->n, w.r, w.Ap, w.st, w.partials, w.hist);
  else cg_update_kernel<<<c.grid_vec, kBlock, 0, s>>>(A->n, w.r, w.Ap, w.st, w.partials, w.hist, (const double *)nullptr,
                                                      (double *)nullptr);
  CGB_CUDA(mark(3));
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

// End of a batch with residual replacement: flush the pending x update, r = b - A x, rr = r.r, alpha = 0.
static int enqueue_replacement(cgb_csr *A, double *x, const LoopCfg &c, cudaStream_t s) {
  Workspace &w = A->ws;
  cg_xflush_kernel<<<c.grid_pup, kBlock, 0, s>>>(A->n, x, w.p0, w.p1, c.fused ? 1 : 0, w.st, 1);
  launch_panel_partials(A, c.sl, s, x, w.r, w.st);
  dispatch_lanes(c.sl.lanes, c.sl.eng, [&](auto Lc, auto Tc) {
    launch_kernel(cg_replace_kernel<decltype(Lc)::value, decltype(Tc)::value>, c.sl, s, csr_view_final(A, w.r),
                  (const double *)x, c.b, w.r, w.st, w.partials);
  });
  if (c.jacobi) cg_rho_kernel<<<c.grid_vec, kBlock, 0, s>>>(A->n, w.r, w.dinv, w.st, w.partials, 1);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

static int build_graph(cgb_csr *A, double *x, const LoopCfg &c, int batch, uint64_t key) {
  Workspace &w = A->ws;
  cudaStream_t s = A->ctx->stream;
  if (w.graph && w.graph_key == key && w.graph_x == x && (!c.replace || w.graph_b == c.b)) return CGB_OK;
  if (w.graph) {
    cudaGraphExecDestroy(w.graph);
    w.graph = nullptr;
  }
  cudaGraph_t g = nullptr;
  CGB_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  int rc = CGB_OK;
  for (int i = 0; i < batch && rc == CGB_OK; i++) rc = enqueue_iteration(A, x, c, s, c.time_kernels && i == 0);
  if (rc == CGB_OK && c.replace) rc = enqueue_replacement(A, x, c, s);
  cudaError_t e = cudaStreamEndCapture(s, &g);
  if (rc != CGB_OK) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
  if (c.l2_window) { // kernel nodes carry the access-policy window themselves
    size_t nn = 0;
    cudaGraphGetNodes(g, nullptr, &nn);
    std::vector<cudaGraphNode_t> nodes(nn);
    cudaGraphGetNodes(g, nodes.data(), &nn);
    cudaKernelNodeAttrValue v;
    memset(&v, 0, sizeof(v));
    v.accessPolicyWindow = p_window(A->ctx, w.p0, A->n);
    // CGB_L2_WINDOW_MODE=k1: only the SpMV nodes (the gathers) carry the window; default: every kernel of the trip
    const char *wm = getenv("CGB_L2_WINDOW_MODE");
    const bool k1_only = wm && !strcmp(wm, "k1");
    for (size_t i = 0; i < nn; i++) {
      cudaGraphNodeType t;
      if (cudaGraphNodeGetType(nodes[i], &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) continue;
      if (k1_only) {
        cudaKernelNodeParams kp;
        if (cudaGraphKernelNodeGetParams(nodes[i], &kp) != cudaSuccess || kp.blockDim.x != (unsigned)c.sl.block || kp.gridDim.x != (unsigned)c.sl.grid) continue;
      }
      cudaGraphKernelNodeSetAttribute(nodes[i], cudaKernelNodeAttributeAccessPolicyWindow, &v);
    }
    cudaGetLastError();
  }
  e = cudaGraphInstantiate(&w.graph, g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) {
    w.graph = nullptr;
    return fail(CGB_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
  }
  w.graph_key = key;
  w.graph_x = x;
  w.graph_b = c.b;
  return CGB_OK;
}

// The loop inside one thread-block cluster (cluster.cuh) when the matrix and the vectors fit its shared memory.
// *ran = false (and CGB_OK) when they do not, or the device cannot co-schedule such a cluster: the caller falls
// back to the cooperative kernel.  CGB_CLUSTER=0 disables it, CGB_CLUSTER_CTAS picks the cluster size (A/B runs).
template <int RPT, int TPB>
int cluster_launch(cgb_ctx *ctx, cgb_csr *A, double *d_x, int ncta, int R, int W, size_t smem, bool *ran) {
  auto kernel = cg_cluster_kernel<RPT, TPB>;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
      (ncta > 8 && cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)) {
    cudaGetLastError();
    return CGB_OK;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ncta);
  cfg.blockDim = dim3(TPB);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = ncta;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int clusters = 0;
  if (cudaOccupancyMaxActiveClusters(&clusters, kernel, &cfg) != cudaSuccess || clusters < 1) {
    cudaGetLastError();
    return CGB_OK;
  }
  Workspace &w = A->ws;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, csr_view(A), d_x, (const double *)w.r, w.st, w.hist, R, W);
  if (e != cudaSuccess) { // a launch-time refusal (e.g. no 16 free SMs in one GPC under MPS): the caller falls back
    cudaGetLastError();
    return CGB_OK;
  }
  *ran = true;
  return CGB_OK;
}

int cluster_try_solve(cgb_ctx *ctx, cgb_csr *A, double *d_x, bool *ran) {
  *ran = false;
  if (const char *e = getenv("CGB_CLUSTER"))
    if (atoi(e) == 0) return CGB_OK;
  if (A->n_cols != A->n) return CGB_OK;
  const int W = A->max_row;
  if (W < 1 || W > 64) return CGB_OK;
  int sizes[3] = {kClusterMaxCtas, 8, 0};
  if (const char *e = getenv("CGB_CLUSTER_CTAS")) sizes[0] = std::max(1, std::min(kClusterMaxCtas, atoi(e))), sizes[1] = 0;
  for (int i = 0; sizes[i] > 0 && !*ran; i++) {
    const int ncta = sizes[i];
    int R = (int)((A->n + ncta - 1) / ncta);
    R = (R + 31) & ~31;
    const size_t smem = cluster_smem_bytes(R, W);
    if (R > 65536 || smem > kClusterSmemMax) continue;
    // threads per CTA (C1, 1,024 rows per CTA: 3.4 us per trip with 512 threads, 3.6 with 1,024, 4.1 with 256 — fewer
    // warps make the block reductions and the barriers cheaper, up to the point where a thread's rows serialise;
    // 256 rows per CTA: 2.9 us with 256 threads, 3.3 with 1,024); CGB_CLUSTER_THREADS overrides (A/B runs)
    int threads = R <= 256 ? 256 : R <= 1024 ? 512 : 1024;
    if (const char *e = getenv("CGB_CLUSTER_THREADS")) threads = atoi(e) >= 1024 ? 1024 : atoi(e) >= 512 ? 512 : 256;
    const int rpt = (R + threads - 1) / threads;
    if (rpt > 4) continue;
#define CGB_CL_LAUNCH(T)                                                                                                 \
  (rpt == 1 ? cluster_launch<1, T>(ctx, A, d_x, ncta, R, W, smem, ran)                                                   \
            : rpt == 2 ? cluster_launch<2, T>(ctx, A, d_x, ncta, R, W, smem, ran) : cluster_launch<4, T>(ctx, A, d_x, ncta, R, W, smem, ran))
    CGB_TRY(threads == 256 ? CGB_CL_LAUNCH(256) : threads == 512 ? CGB_CL_LAUNCH(512) : CGB_CL_LAUNCH(1024));
#undef CGB_CL_LAUNCH
  }
  return CGB_OK;
}

} // namespace cgb

extern "C" {

void cgb_cg_default_opts(cgb_cg_opts *o) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->loop_mode = CGB_LOOP_AUTO;
  o->check_every = 0;
  o->fuse_p_update = 0;
}

int cgb_cg_solve(cgb_ctx *ctx, cgb_csr *A, const double *d_b, double *d_x, double abs_tol, int64_t max_iter,
                 const cgb_cg_opts *opts_in, cgb_cg_stats *stats) {
  CGB_REQUIRE(ctx, "cgb_cg_solve: ctx is NULL");
  if (!A) return fail(CGB_ERR_STATE, "No Matrix given");                    // src/CG.hpp:270-272
  if (!d_b) return fail(CGB_ERR_STATE, "No right hand side to solve for");  // src/CG.hpp:266-268
  CGB_REQUIRE(d_x, "cgb_cg_solve: x is NULL");
  CGB_REQUIRE(A->ctx == ctx, "cgb_cg_solve: matrix belongs to another context");
  CGB_REQUIRE(((uintptr_t)d_x & 15) == 0, "cgb_cg_solve: x must be 16-byte aligned (128-bit vector accesses)");
  if (A->n_cols > A->n)
    return fail(CGB_ERR_INVALID, "cgb_cg_solve: the matrix is not square (%lld rows, column ids up to %lld): rows without entries were "
                "compacted away (CGB_COO_COMPACT_ROWS, the reference reader's dialect)", (long long)A->n, (long long)A->n_cols);
  cgb_cg_opts o;
  if (opts_in) o = *opts_in; else cgb_cg_default_opts(&o);
  if (stats) memset(stats, 0, sizeof(*stats));
  const int64_t n = A->n;
  if (max_iter < 0) max_iter = n + 1; // reference cap (src/CG.hpp:436)
  if (n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;

  int mode = o.loop_mode;
  int batch = o.check_every > 0 ? o.check_every : 32;
  const bool replace = o.residual_replace_every > 0;
  if (replace) batch = o.residual_replace_every; // a batch ends with the replacement
  if ((int64_t)batch > max_iter) batch = (int)std::max<int64_t>(1, max_iter);
  const int64_t hist_cap = o.record_history ? max_iter + 1 : 0;
  CGB_TRY(csr_prepare(A));
  // One cooperative kernel for the whole loop (persistent.cuh) when asked for, or, under AUTO, when the
  // system is small enough that latency rules (CGB_PERSISTENT_MAX_N rows); it serves CSR rows only.
  int64_t persistent_max_n = 131072; // measured cross-over on B200: 2D 5-pt 256^2 still wins (89k vs 77k it/s), 512^2 loses
  if (const char *e = getenv("CGB_PERSISTENT_MAX_N")) persistent_max_n = atoll(e);
  const bool jacobi = o.jacobi != 0;
  const bool can_persist = !replace && !jacobi && ctx->coop_ok && (csr_engine(A) == kEngTma || csr_engine(A) == kEngVec) && !o.time_kernels;
  // The same loop inside one thread-block cluster (cluster.cuh) when the system fits its shared memory.
  const bool want_cluster = mode == CGB_LOOP_CLUSTER || (mode == CGB_LOOP_AUTO && n <= persistent_max_n);
  if (mode == CGB_LOOP_CLUSTER && !can_persist)
    return fail(CGB_ERR_UNSUPPORTED, "CGB_LOOP_CLUSTER serves plain CG on the CSR engines only");
  const bool persistent = can_persist && (mode == CGB_LOOP_PERSISTENT || mode == CGB_LOOP_CLUSTER ||
                                          (mode == CGB_LOOP_AUTO && n <= persistent_max_n));
  const bool cluster_only = mode == CGB_LOOP_CLUSTER;
  if (mode == CGB_LOOP_AUTO || mode == CGB_LOOP_PERSISTENT || mode == CGB_LOOP_CLUSTER)
    mode = persistent ? CGB_LOOP_PERSISTENT : CGB_LOOP_GRAPH;
  const bool fused = (o.fuse_p_update != 0 && csr_panels(A) == 0 && !jacobi) || persistent; // the persistent kernel forms p in the gather (ping-pong buffers)
  CGB_TRY(workspace_ensure(A, n, fused, hist_cap));
  Workspace &w = A->ws;
  // L2 window: explicit option, or CGB_L2_WINDOW=1/0; only meaningful without the fused p-update (p ping-pongs there)
  bool l2w = o.l2_persist_p != 0;
  if (const char *e = getenv("CGB_L2_WINDOW")) l2w = atoi(e) != 0;
  l2w = l2w && !fused && ctx->l2_persist_max > 0 && ctx->l2_window_max > 0;
  if (l2w && ctx->l2_persist_set == 0) {
    // carve-out: the device maximum, or CGB_L2_CARVE_MB (A/B runs: a carve-out sized to the window instead)
    size_t carve = ctx->l2_persist_max;
    if (const char *e = getenv("CGB_L2_CARVE_MB")) carve = std::min(carve, (size_t)(atof(e) * 1048576.0));
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess) ctx->l2_persist_set = carve;
    else {
      cudaGetLastError();
      l2w = false;
    }
  }
  if (jacobi && !w.dinv) { // 1 / diag(A), once per matrix
    int *d_bad = nullptr;
    CGB_CUDA(cudaMalloc(&w.dinv, sizeof(double) * (size_t)n));
    CGB_CUDA(cudaMalloc(&d_bad, sizeof(int)));
    CGB_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), s));
    diag_inv_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, A->rowptr, A->cols, A->vals, w.dinv, d_bad);
    int bad = 0;
    cudaError_t e = cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_bad);
    if (e != cudaSuccess || bad) {
      cudaFree(w.dinv);
      w.dinv = nullptr;
      if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "diag_inv_kernel: %s", cudaGetErrorString(e));
      return fail(CGB_ERR_UNSUPPORTED, "Jacobi preconditioner: a row has no non-zero diagonal entry");
    }
  }
  LoopCfg cfg{fused, l2w, o.time_kernels != 0, spmv_launch_cfg(A), A->grid_vec, A->grid_pup, jacobi, replace, d_b,
              A->vec_tma && !jacobi && ((uintptr_t)d_x & 15) == 0};
  const char *l2mode = getenv("CGB_L2_WINDOW_MODE");
  if (cfg.vec_tma) { // (static kernels: this translation unit's own instances need their shared-memory limit raised)
    CGB_CUDA(cudaFuncSetAttribute(cg_pupdate_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vec_tma_smem_bytes(3)));
    CGB_CUDA(cudaFuncSetAttribute(cg_update_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vec_tma_smem_bytes(2)));
  }
  if (l2w && !(l2mode && !strcmp(l2mode, "k1"))) { // (k1: the window rides on the SpMV graph nodes only)
    cudaStreamAttrValue v;
    memset(&v, 0, sizeof(v));
    v.accessPolicyWindow = p_window(ctx, w.p0, n);
    CGB_CUDA(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v));
  }
  struct WindowGuard { // the window never outlives the solve
    cudaStream_t s;
    bool on;
    ~WindowGuard() {
      if (!on) return;
      cudaStreamAttrValue v;
      memset(&v, 0, sizeof(v));
      cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v);
      cudaCtxResetPersistingL2Cache();
    }
  } wguard{s, l2w};
  if (cfg.time_kernels)
    for (auto &e : w.ev)
      if (!e) CGB_CUDA(cudaEventCreate(&e));

  cudaEvent_t t0, t1;
  CGB_CUDA(cudaEventCreate(&t0));
  CGB_CUDA(cudaEventCreate(&t1));
  struct EvGuard {
    cudaEvent_t a, b;
    ~EvGuard() {
      cudaEventDestroy(a);
      cudaEventDestroy(b);
    }
  } guard{t0, t1};

  CGB_CUDA(cudaEventRecord(t0, s));
  // state + work vectors (reference: src/CG.hpp:276-301)
  CgState init;
  memset(&init, 0, sizeof(init));
  init.tol = abs_tol;
  init.max_iter = max_iter;
  init.record_history = o.record_history ? 1 : 0;
  init.hist_cap = w.hist_cap;
  init.precond = jacobi ? 1 : 0;
  init.done = 1; // K0 arms the loop; until then nothing may run
  w.h_st[0] = init;
  CGB_CUDA(cudaMemcpyAsync(w.st, &w.h_st[0], sizeof(CgState), cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemsetAsync(w.p0, 0, sizeof(double) * (size_t)n, s));
  if (fused) CGB_CUDA(cudaMemsetAsync(w.p1, 0, sizeof(double) * (size_t)n, s));
  const int64_t extra = csr_panels(A) > 0 ? (int64_t)csr_panels(A) - 1 : 0; // partial launches per SpMV-shaped pass
  launch_panel_partials(A, cfg.sl, s, d_x, w.r, nullptr); // r doubles as the accumulator of K0
  dispatch_lanes(cfg.sl.lanes, cfg.sl.eng, [&](auto Lc, auto Tc) {
    launch_kernel(cg_init_kernel<decltype(Lc)::value, decltype(Tc)::value>, cfg.sl, s, csr_view_final(A, w.r),
                  (const double *)d_x, d_b, w.r, w.st, w.partials, w.hist);
  });
  if (jacobi) cg_rho_kernel<<<cfg.grid_vec, kBlock, 0, s>>>(n, w.r, w.dinv, w.st, w.partials, 0);
  CGB_CUDA(cudaGetLastError());
  int64_t launches = 1 + extra + (jacobi ? 1 : 0);

  const int per_iter = (fused ? 2 : 3) + (int)extra;
  const int per_replace = 2 + (int)extra + (jacobi ? 1 : 0);
  double t_p = 0, t_s = 0, t_u = 0;
  int64_t samples = 0;
  bool in_cluster = false;
  if (persistent && want_cluster) CGB_TRY(cluster_try_solve(ctx, A, d_x, &in_cluster));
  if (cluster_only && !in_cluster)
    return fail(CGB_ERR_UNSUPPORTED, "CGB_LOOP_CLUSTER: matrix and vectors do not fit one cluster's shared memory (or the device refused the cluster launch)");
  if (in_cluster) {
    launches += 1;
    mode = CGB_LOOP_CLUSTER;
  }
  if (persistent && !in_cluster) {
    int rc = CGB_OK;
    dispatch_lanes(A->lanes, kEngVec, [&](auto Lc, auto) {
      constexpr int L = decltype(Lc)::value;
      auto kernel = cg_persistent_kernel<L>;
      int occ = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kBlock, 0) != cudaSuccess || occ < 1) {
        rc = fail(CGB_ERR_CUDA, "cg_persistent_kernel: occupancy query failed");
        return;
      }
      // as many CTAs as one pass over the rows needs, never more than can be co-resident
      const int64_t rows_per_cta = (int64_t)(kBlock / L) * 2;
      int grid = (int)std::min<int64_t>((int64_t)occ * ctx->sm_count, (n + rows_per_cta - 1) / rows_per_cta);
      if (const char *e = getenv("CGB_PERSISTENT_CTAS")) grid = std::min(grid, std::max(1, atoi(e)));
      if (grid < 1) grid = 1;
      if (grid > kMaxPartials) grid = kMaxPartials;
      CsrView V = csr_view(A);
      double *xp = d_x, *rp = w.r, *p0 = w.p0, *p1 = w.p1, *ap = w.Ap, *parts = w.partials, *hist = w.hist;
      CgState *stp = w.st;
      void *args[] = {&V, &xp, &rp, &p0, &p1, &ap, &stp, &parts, &hist};
      cudaError_t e = cudaLaunchCooperativeKernel((void *)kernel, dim3(grid), dim3(kBlock), args, 0, s);
      if (e != cudaSuccess) rc = fail(CGB_ERR_CUDA, "cudaLaunchCooperativeKernel: %s", cudaGetErrorString(e));
    });
    CGB_TRY(rc);
    launches += 1;
  }
  if (mode == CGB_LOOP_GRAPH) {
    const uint64_t key = ((uint64_t)batch << 32) ^ ((uint64_t)cfg.sl.lanes << 8) ^ ((uint64_t)fused << 1) ^
                         (uint64_t)cfg.time_kernels ^ ((uint64_t)cfg.sl.grid << 16) ^ ((uint64_t)cfg.sl.eng << 2) ^
                         ((uint64_t)cfg.l2_window << 4) ^ ((uint64_t)replace << 5) ^ ((uint64_t)jacobi << 6) ^ ((uint64_t)cfg.vec_tma << 7) ^ ((uint64_t)csr_panels(A) << 48);
    CGB_TRY(build_graph(A, d_x, cfg, batch, key));
  }
  // Batches are issued one ahead of the poll so the device never waits for the host; kernels of an
  // overshooting batch return at once because the state says done.
  cudaEvent_t polled[2];
  CGB_CUDA(cudaEventCreateWithFlags(&polled[0], cudaEventDisableTiming));
  CGB_CUDA(cudaEventCreateWithFlags(&polled[1], cudaEventDisableTiming));
  EvGuard guard2{polled[0], polled[1]};
  const bool pipelined = !cfg.time_kernels;
  int64_t issued = 0; // iterations issued so far
  int slot = 0;
  bool finished = persistent; // the cooperative kernel has run the whole loop already
  bool have_prev = false;
  bool prev_active_at_start = true;
  while (!finished) {
    // issue one batch
    if (mode == CGB_LOOP_GRAPH) {
      CGB_CUDA(cudaGraphLaunch(w.graph, s));
    } else {
      for (int i = 0; i < batch; i++) CGB_TRY(enqueue_iteration(A, d_x, cfg, s, cfg.time_kernels && i == 0));
      if (replace) CGB_TRY(enqueue_replacement(A, d_x, cfg, s));
    }
    launches += (int64_t)batch * per_iter + (replace ? per_replace : 0);
    issued += batch;
    CGB_CUDA(cudaMemcpyAsync(&w.h_st[slot], w.st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaEventRecord(polled[slot], s));
    int check = slot;
    if (pipelined) {
      if (!have_prev) { // nothing to look at yet: go issue the second batch
        have_prev = true;
        slot ^= 1;
        if (issued < max_iter) continue;
      } else {
        check = slot ^ 1;
      }
    }
    CGB_CUDA(cudaEventSynchronize(polled[check]));
    const CgState &h = w.h_st[check];
    if (ctx->progress) ctx->progress((int64_t)h.iter, n, h.rr, ctx->progress_user); // reference: percentage line, src/CG.hpp:428-433
    if (cfg.time_kernels) {
      if (prev_active_at_start) {
        float a = 0, b = 0, c = 0;
        CGB_CUDA(cudaEventElapsedTime(&a, w.ev[0], w.ev[1]));
        CGB_CUDA(cudaEventElapsedTime(&b, w.ev[1], w.ev[2]));
        CGB_CUDA(cudaEventElapsedTime(&c, w.ev[2], w.ev[3]));
        t_p += a; t_s += b; t_u += c;
        samples++;
      }
      prev_active_at_start = !(h.done || h.iter >= h.max_iter);
    }
    if (h.done || h.iter >= h.max_iter) finished = true;
    slot ^= 1;
  }
  // the x update of the last trip (deferred: see cg_kernels.cuh), then drain and take the final state
  if (!persistent) {
    cg_xflush_kernel<<<cfg.grid_pup, kBlock, 0, s>>>(n, d_x, w.p0, w.p1, fused ? 1 : 0, w.st, 0);
    CGB_CUDA(cudaGetLastError());
    launches += 1;
  }
  CGB_CUDA(cudaMemcpyAsync(&w.h_st[0], w.st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaEventRecord(t1, s));
  CGB_CUDA(cudaEventSynchronize(t1));
  const CgState &f = w.h_st[0];
  w.hist_count = o.record_history ? std::min<int64_t>(f.iter + 1, w.hist_cap) : 0;
  if (stats) {
    float ms = 0.f;
    CGB_CUDA(cudaEventElapsedTime(&ms, t0, t1));
    stats->iterations = f.iter;
    stats->rr = f.rr;
    stats->rr0 = f.rr0;
    stats->converged = f.done;
    stats->nan_stop = f.nan_stop;
    stats->solve_ms = ms;
    stats->timed_samples = samples;
    if (samples > 0) {
      stats->pupdate_ms = t_p / samples;
      stats->spmv_ms = t_s / samples;
      stats->update_ms = t_u / samples;
    }
    stats->kernel_launches = launches;
    stats->loop_mode = mode;
    stats->format = cfg.sl.eng == kEngSell ? CGB_FMT_SELL : cfg.sl.eng == kEngDict ? CGB_FMT_DICT : CGB_FMT_CSR;
  }
  return CGB_OK;
}

int cgb_cg_solve_host(cgb_ctx *ctx, int64_t n, int64_t nnz, const double *h_vals, const int32_t *h_cols,
                      const int32_t *h_rowptr, const double *h_b, const double *h_x0, double *h_x, double abs_tol,
                      int64_t max_iter, const cgb_cg_opts *opts, cgb_cg_stats *stats) {
  CGB_REQUIRE(ctx && h_x, "cgb_cg_solve_host: NULL argument");
  if (!h_b) return fail(CGB_ERR_STATE, "No right hand side to solve for");
  if (!h_rowptr) return fail(CGB_ERR_STATE, "No Matrix given");
  ScopedDevice sd(ctx->device);
  cgb_csr *A = nullptr;
  CGB_TRY(cgb_csr_create(ctx, n, nnz, h_vals, h_cols, h_rowptr, &A));
  double *d_b = nullptr, *d_x = nullptr;
  int rc = cgb_malloc(ctx, sizeof(double) * (size_t)n, (void **)&d_b);
  if (rc == CGB_OK) rc = cgb_malloc(ctx, sizeof(double) * (size_t)n, (void **)&d_x);
  if (rc == CGB_OK) rc = cgb_memcpy_h2d(ctx, d_b, h_b, sizeof(double) * (size_t)n, 0);
  if (rc == CGB_OK)
    rc = h_x0 ? cgb_memcpy_h2d(ctx, d_x, h_x0, sizeof(double) * (size_t)n, 0)
              : cgb_memset(ctx, d_x, 0, sizeof(double) * (size_t)n); // x defaults to zeros (src/CG.hpp:291-297)
  if (rc == CGB_OK) rc = cgb_cg_solve(ctx, A, d_b, d_x, abs_tol, max_iter, opts, stats);
  if (rc == CGB_OK) rc = cgb_memcpy_d2h(ctx, h_x, d_x, sizeof(double) * (size_t)n, 1);
  cgb_free(ctx, d_b);
  cgb_free(ctx, d_x);
  cgb_csr_destroy(A);
  return rc;
}

int cgb_cg_history(cgb_ctx *ctx, cgb_csr *A, double *h_out, int64_t capacity, int64_t *count) {
  CGB_REQUIRE(ctx && A && count, "cgb_cg_history: NULL argument");
  *count = A->ws.hist_count;
  if (!h_out || capacity <= 0 || A->ws.hist_count == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  const int64_t m = std::min<int64_t>(capacity, A->ws.hist_count);
  CGB_CUDA(cudaMemcpyAsync(h_out, A->ws.hist, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

} // extern "C"
