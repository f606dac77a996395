/* cgb200.h — C-ABI of the B200-native Conjugate Gradient hot path.
 *
 * This is the drop-in boundary: plain C types, opaque handles, int status returns (0 = ok),
 * no exceptions, no torch types.  The C++ headers in this directory (CG.hpp,
 * LinearAlgebraTypes.hpp, VectorOperations.hpp — same names and classes as the reference's
 * src/ headers) are thin wrappers over these entry points; INTEGRATION.md shows the binding.
 *
 * The reference has no FFI of its own: its boundary is the header-only template API in
 * namespace CGSolver.  Each entry point below cites the reference interface it replaces
 * (paths relative to the reference tree).
 *
 * Every pointer named d_* is a DEVICE pointer obtained from cgb_malloc (or owned by a handle);
 * every pointer named h_* is a HOST pointer.  All arithmetic is IEEE fp64 on the GPU; there is
 * no CPU fallback: without a usable sm_100 device every compute call fails with CGB_ERR_CUDA.
 */
#ifndef CGB200_H
#define CGB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CGB_VERSION 100

enum {
  CGB_OK = 0,
  CGB_ERR_INVALID = 1, /* bad argument */
  CGB_ERR_CUDA = 2,    /* CUDA runtime error, see cgb_last_error() */
  CGB_ERR_NCCL = 3,    /* NCCL error / NCCL not loadable */
  CGB_ERR_STATE = 4,   /* missing matrix / right-hand side (src/CG.hpp:266-272) */
  CGB_ERR_IO = 5,      /* Matrix Market file problems */
  CGB_ERR_UNSUPPORTED = 6 /* the matrix does not fit the requested storage format (CGB_FMT_DICT) */
};

typedef struct cgb_ctx cgb_ctx;       /* device + stream; replaces acpp::sycl::queue (src/CG.hpp:61,70-77) */
typedef struct cgb_csr cgb_csr;       /* device CSR matrix; replaces CGSolver::Matrix (src/LinearAlgebraTypes.hpp:57-132) */
typedef struct cgb_event cgb_event;   /* replaces acpp::sycl::event */
typedef struct cgb_dist cgb_dist;     /* row-partitioned multi-GPU solver state (new) */

/* Thread-local text of the last failure in the calling thread ("" if none). */
const char *cgb_last_error(void);
int cgb_version(void);
/* Number of visible CUDA devices (0 when there is none; never fails). */
int cgb_device_count(void);

/* ---- context ------------------------------------------------------------------------------
 * replaces: default-constructed asycl::queue (src/CG.hpp:72), queue::wait_and_throw (:561-578) */
int cgb_ctx_create(int device, cgb_ctx **out);
int cgb_ctx_destroy(cgb_ctx *ctx);
int cgb_ctx_sync(cgb_ctx *ctx);
/* cudaStream_t of the context, for callers that record their own events on it. */
void *cgb_ctx_stream(cgb_ctx *ctx);
int cgb_ctx_device(cgb_ctx *ctx);
int cgb_ctx_sm_count(cgb_ctx *ctx);

/* Progress of a running solve, for the reference's Debuglevel::Verbose percentage line (src/CG.hpp:428-433: it
 * prints counter / N every 100 trips).  The callback runs on the calling host thread each time the solve polls the
 * device-side state (once per batch of check_every trips); the one-kernel loops for small systems never poll.
 * fn == NULL removes it. */
typedef void (*cgb_progress_fn)(int64_t trips, int64_t n, double rr, void *user);
int cgb_ctx_set_progress(cgb_ctx *ctx, cgb_progress_fn fn, void *user);

/* ---- raw device memory ---------------------------------------------------------------------
 * replaces: asycl::malloc_device / free / queue.copy / queue.fill
 * (src/LinearAlgebraTypes.hpp:43-49,109-119,160-183,213-224).  Allocations are 256-byte aligned.
 * Copies are ordered on the context stream; the *_sync flag waits for completion. */
int cgb_malloc(cgb_ctx *ctx, size_t bytes, void **d_out);
int cgb_free(cgb_ctx *ctx, void *d_ptr);
/* Device memory freed by the library (cgb_free, cgb_csr_destroy, work vectors, multi-GPU blocks) is kept in a
 * per-device cache of whole allocations and handed out again for requests of the same size, so repeated
 * host-buffer solves (cgb_cg_solve_host, cgb_dist_set_matrix) cost no driver allocation after the first; the
 * reference allocates afresh on every setMatrix / solve (src/LinearAlgebraTypes.hpp:109-119, src/CG.hpp:276-301).
 * So cgb_free RETAINS memory: parked blocks are invisible to the other allocators of the process (NCCL, cub, torch,
 * cupy), and only the library's own failed allocation trims them.  At most a quarter of the device's memory and never
 * more than 32 GB stay parked per device (CGB_ALLOC_CACHE_MAX_GB overrides the bound, CGB_ALLOC_CACHE=0 disables the
 * cache); cgb_trim returns the cached blocks to the driver -- call it before handing the GPU to another allocator.
 * cgb_dist_create trims before it creates the NCCL communicator. */
int cgb_trim(void);
int cgb_memcpy_h2d(cgb_ctx *ctx, void *d_dst, const void *h_src, size_t bytes, int sync);
int cgb_memcpy_d2h(cgb_ctx *ctx, void *h_dst, const void *d_src, size_t bytes, int sync);
int cgb_memcpy_d2d(cgb_ctx *ctx, void *d_dst, const void *d_src, size_t bytes);
/* Direction inferred from the pointers (unified addressing), like a SYCL USM queue.copy. */
int cgb_memcpy(cgb_ctx *ctx, void *dst, const void *src, size_t bytes, int sync);
int cgb_memset(cgb_ctx *ctx, void *d_ptr, int byte, size_t bytes);
int cgb_fill_f64(cgb_ctx *ctx, double *d_ptr, double value, size_t n);
/* Page-locked host staging (pinned H2D/D2H run at PCIe speed). */
int cgb_host_alloc(size_t bytes, void **h_out);
int cgb_host_free(void *h_ptr);

/* ---- events --------------------------------------------------------------------------------
 * replaces: asycl::event returned by every VectorOperations method (src/VectorOperations.hpp).
 * The stream is in-order, so every depends_on() list of the reference is satisfied trivially. */
int cgb_event_record(cgb_ctx *ctx, cgb_event **out);
int cgb_event_sync(cgb_event *ev);
int cgb_event_elapsed_ms(cgb_event *start, cgb_event *stop, double *ms);
int cgb_event_destroy(cgb_event *ev);

/* ---- CSR matrix ----------------------------------------------------------------------------
 * replaces: Matrix<DT>::init (src/LinearAlgebraTypes.hpp:101-121) — three device allocations and
 * three H2D copies.  n = number of rows, rowptr has n+1 entries, 32-bit indices as in the
 * reference (std::vector<int>).  The handle also owns the SpMV launch plan (row tiles). */
int cgb_csr_create(cgb_ctx *ctx, int64_t n, int64_t nnz, const double *h_vals, const int32_t *h_cols,
                   const int32_t *h_rowptr, cgb_csr **out);
/* Adopt arrays already resident on the device (ownership is NOT taken unless take_ownership). */
int cgb_csr_from_device(cgb_ctx *ctx, int64_t n, int64_t nnz, double *d_vals, int32_t *d_cols,
                        int32_t *d_rowptr, int take_ownership, cgb_csr **out);
int cgb_csr_destroy(cgb_csr *A);
int64_t cgb_csr_n(const cgb_csr *A);
int64_t cgb_csr_nnz(const cgb_csr *A);
double *cgb_csr_vals(cgb_csr *A);    /* Matrix::data_ptr()    (:74) */
int32_t *cgb_csr_cols(cgb_csr *A);   /* Matrix::columns_ptr() (:80) */
int32_t *cgb_csr_rowptr(cgb_csr *A); /* Matrix::rows_ptr()    (:86) */
/* Build the SELL-C-sigma shadow copy used by the SpMV when format == CGB_FMT_SELL
 * (C = 32 rows per slice; sigma = sorting window in rows, 1 = no permutation). */
int cgb_csr_build_sell(cgb_csr *A, int sigma);
/* Stored entries of the SELL copy including padding (0 before cgb_csr_build_sell). */
int64_t cgb_csr_sell_padded(const cgb_csr *A);
/* Selects the storage the SpMV-shaped kernels stream.  CGB_FMT_AUTO and CGB_FMT_CSR use the caller's CSR
 * arrays as they are (TMA-staged tiles); CGB_FMT_AUTO switches to the SELL copy when the row lengths are
 * irregular (tiles that overflow a TMA stage, or a long tail); CGB_FMT_SELL builds the SELL copy on first
 * use (sigma = 32768) unless cgb_csr_build_sell was called with another window. */
enum { CGB_FMT_AUTO = 0, CGB_FMT_CSR = 1, CGB_FMT_SELL = 2, CGB_FMT_DICT = 3 };
/* CGB_FMT_DICT (explicit option, never chosen by CGB_FMT_AUTO): row-pattern dictionary copy.  A matrix from a
 * constant-coefficient stencil repeats a handful of row patterns (the sequence of (column - row, value) pairs of a
 * row: 27 for a 7-point Laplacian on a box); with at most 1024 distinct patterns holding at most 3072 entries in
 * total, the matrix is stored as one 16-bit pattern id per row plus the pattern table: 2 bytes per ROW instead of
 * CSR's 12 per non-zero.  Lossless (every row is verified against its pattern); row sums keep the CSR entry order,
 * so results are bit-identical to the one-lane-per-row CSR engines.  cgb_csr_set_format returns CGB_ERR_UNSUPPORTED
 * (and leaves the format unchanged) when the matrix does not fit.  cgb_csr_dict_bytes: size of the copy (0: not
 * built); cgb_csr_dict_patterns: distinct row patterns found. */
int64_t cgb_csr_dict_bytes(const cgb_csr *A);
int cgb_csr_dict_patterns(const cgb_csr *A);
int cgb_csr_set_format(cgb_csr *A, int format);
int cgb_csr_get_format(const cgb_csr *A);
/* Column panels of the SELL engine: A = [A_0 | ... | A_{P-1}] split by column range, one pass per panel
 * accumulating into the output, so that the part of the gathered vector a pass touches stays in L2 (random
 * sparsity patterns whose vector exceeds L2: BASELINE config C5).  0 = automatic (by the vector's size; the
 * default), 1 = never split, P = exactly P panels.  Takes effect at the next SpMV-shaped call.  The row sum
 * becomes (A_0 x + A_1 x) + ... : a different summation order from the reference's left-to-right row
 * (src/VectorOperations.hpp:455-462), inside north_star's 1e-10 tolerance, bitwise reproducible. */
int cgb_csr_set_panels(cgb_csr *A, int panels);
/* Panels in use after the last SpMV-shaped call (0: the matrix is streamed in one piece). */
int cgb_csr_get_panels(const cgb_csr *A);

/* On-device generation of the synthetic stencils of SURVEY.md §8(d) (identical arrays to
 * conjugategradient_b200/generators.py): kind 0 = 2D 5-point (m x m), 1 = 3D 7-point,
 * 2 = 3D 27-point on an m x m x mz grid.  Rows [row_begin, row_end) of the global matrix are
 * generated (a contiguous row block, for the row-partitioned multi-GPU case); column indices
 * stay GLOBAL.  Pass row_begin = 0, row_end = -1 for the whole matrix. */
int cgb_gen_stencil(cgb_ctx *ctx, int kind, int64_t m, int64_t mz, int64_t row_begin, int64_t row_end,
                    cgb_csr **out);

/* On-device generation of the random SPD matrix with power-law row lengths:
 * strictly-lower degree of row i = min(d_max, floor(d_min * u^(-1/(gamma-1)))) capped at i, distinct
 * columns (one per stratum of [0, i)), values in (-1, 0), mirrored, a_ii = 1 + sum |a_ij|.
 * Deterministic in (n, seed, d_min, d_max, gamma).  Rows [row_begin, row_end) with GLOBAL columns
 * (row_begin = 0, row_end = -1: the whole matrix). */
int cgb_gen_random_spd(cgb_ctx *ctx, int64_t n, uint64_t seed, int d_min, int d_max, double gamma, int64_t row_begin,
                       int64_t row_end, cgb_csr **out);

/* ---- per-op kernels ------------------------------------------------------------------------
 * Scalar operands are DEVICE pointers dereferenced at kernel run time, and reductions ADD INTO
 * *d_result (no initialisation), exactly like the reference's sycl::reduction without
 * initialize_to_identity.  All are alias-safe for the in-place uses in src/CG.hpp:390,392,418. */
/* y = A x                      replaces VectorOperations::spmv (src/VectorOperations.hpp:438-466) */
int cgb_spmv(cgb_ctx *ctx, cgb_csr *A, const double *d_x, double *d_y);
/* *d_result += sum x[i]*y[i]   replaces dot_product_trivial/_optimised/dot_product (:287-309,110-285) */
int cgb_dot_acc(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, double *d_result);
/* *d_result += sum x[i]^2      replaces norm (:311-331) */
int cgb_norm2_acc(cgb_ctx *ctx, int64_t n, const double *d_x, double *d_result);
/* out = (*d_a) x + (*d_b) y    replaces saxpby (:349-367) */
int cgb_axpby(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, const double *d_a,
              const double *d_b, double *d_out);
/* out = x + sign * (*d_b) y, sign = +1: sapbx (:410-428), sign = -1: sambx (:380-397) */
int cgb_xpby(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, const double *d_b, int sign,
             double *d_out);
/* One device scalar: *d = value (Scalar::init, src/LinearAlgebraTypes.hpp:216-224), and read-back. */
int cgb_scalar_set(cgb_ctx *ctx, double *d_scalar, double value);
int cgb_scalar_get(cgb_ctx *ctx, const double *d_scalar, double *h_value);

/* ---- the CG solve --------------------------------------------------------------------------
 * replaces: CG::solve (src/CG.hpp:255-454) with identical stop rule: ABSOLUTE tolerance on the
 * recurrence residual, checked against r.r from the START of the iteration (one-iteration lag,
 * :396-404), NaN stops (:401), at most max_iter loop trips (reference: N+1, :436). */
enum { CGB_LOOP_AUTO = 0, CGB_LOOP_GRAPH = 1, CGB_LOOP_PERSISTENT = 2, CGB_LOOP_STEPWISE = 3, CGB_LOOP_CLUSTER = 4 };

typedef struct cgb_cg_opts {
  int loop_mode;        /* CGB_LOOP_*: CUDA-graph batches, one persistent cooperative kernel, one thread-block
                           cluster with the matrix and the vectors resident in its shared memory (systems of
                           about 1.5 MB: CGB_ERR_UNSUPPORTED when it does not fit), or one launch per kernel
                           with a host check every check_every iterations.  AUTO: cluster when it fits, else
                           the cooperative kernel up to 131,072 rows, else graph batches */
  int check_every;      /* iterations per graph launch / between host polls (0 = default 32) */
  int fuse_p_update;    /* 1: p = r + beta p is formed inside the SpMV gather (2 kernels per
                           iteration, double-buffered p); 0: separate p-update kernel (3 kernels) */
  int time_kernels;     /* 1: bracket the kernels of one iteration per batch with CUDA events */
  int l2_persist_p;     /* 1: L2 persisting access-policy window over the gathered vector */
  int record_history;   /* 1: keep r.r after every iteration (read with cgb_cg_history) */
  int residual_replace_every; /* k > 0: every k iterations r is recomputed as b - A x (true residual) and rr
                           from it, so the stop rule cannot run ahead of the real residual the way the
                           reference's recurrence does at tight tolerances; one extra
                           SpMV per k iterations; single-GPU graph / stepwise loops.  0 (default): classical
                           recurrence, as the reference */
  int jacobi;           /* 1: Jacobi-preconditioned CG (z = D^-1 r formed on the fly from 1/diag(A); alpha and beta
                           from r.z; the stop rule stays on |r|).  A different algorithm from the reference's:
                           iteration counts are not comparable with it.  Single-GPU graph / stepwise loops.
                           CGB_ERR_UNSUPPORTED when a row has no non-zero diagonal entry.  0 (default): plain CG */
  int reserved[4];
} cgb_cg_opts;

typedef struct cgb_cg_stats {
  int64_t iterations;   /* loop trips executed (== reference loop trips) */
  double rr;            /* recurrence r.r after the last trip */
  double rr0;           /* r0.r0 */
  int converged;        /* 1 if the stop test fired (includes NaN stop) */
  int nan_stop;         /* 1 if the stop was caused by NaN */
  double solve_ms;      /* CUDA-event time of the whole solve on the context stream */
  double spmv_ms;       /* mean duration of the SpMV(+p.Ap) kernel (time_kernels only, else 0) */
  double update_ms;     /* mean duration of the x/r update(+r.r) kernel */
  double pupdate_ms;    /* mean duration of the p-update kernel (0 when fused) */
  int64_t timed_samples;
  int64_t kernel_launches; /* kernels launched by this solve */
  int loop_mode;        /* mode actually used */
  int format;           /* CGB_FMT_* actually used */
  double reserved[4];
} cgb_cg_stats;

void cgb_cg_default_opts(cgb_cg_opts *opts);
/* Device-resident solve.  d_x: in = initial guess, out = solution (CG::setInital / x zeros,
 * src/CG.hpp:215-219,291-297).  max_iter < 0 means the reference cap n+1.  Work vectors are
 * cached inside the csr handle and reused across solves.  d_x must be 16-byte aligned (every cgb_malloc'ed
 * pointer is; an odd-offset sub-vector is refused with CGB_ERR_INVALID): the update kernels use 128-bit accesses. */
int cgb_cg_solve(cgb_ctx *ctx, cgb_csr *A, const double *d_b, double *d_x, double abs_tol, int64_t max_iter,
                 const cgb_cg_opts *opts, cgb_cg_stats *stats);
/* Host-buffer solve: what Tester.cpp does between setMatrix and extract (test/Tester.cpp:36-49):
 * uploads CSR + b (+ x0 if non-NULL), solves, downloads x.  Used for the end-to-end timing. */
int cgb_cg_solve_host(cgb_ctx *ctx, int64_t n, int64_t nnz, const double *h_vals, const int32_t *h_cols,
                      const int32_t *h_rowptr, const double *h_b, const double *h_x0, double *h_x,
                      double abs_tol, int64_t max_iter, const cgb_cg_opts *opts, cgb_cg_stats *stats);
/* r.r history of the last solve on A with record_history: h_out[k] = r.r after k trips. */
int cgb_cg_history(cgb_ctx *ctx, cgb_csr *A, double *h_out, int64_t capacity, int64_t *count);

/* ---- a solve that allocates nothing ---------------------------------------------------------------
 * No counterpart in the reference (its solve allocates every work vector afresh, src/CG.hpp:276-301); added for
 * hosts that cannot let this library allocate device memory: applications that hand (nearly) all of the GPU's
 * memory to their own pool at start-up, so that a later cudaMalloc from a plugin fails; time-step loops that cannot
 * afford the device-wide synchronisation of cudaMalloc / cudaFree; deployments that forbid driver allocations after
 * initialisation.  cgb_cg_solve keeps r, p, A p, the solver state and the reduction partials in memory it allocates
 * on first use; this entry point keeps them in a fixed work area inside the library image instead: a __device__
 * array of CGB_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits.  It makes
 * no call to cudaMalloc, cudaFree or any allocating routine, creates no stream or event, and needs no cgb_ctx and
 * no cgb_csr: the caller passes its own device arrays (CSR with 32-bit indices, b, x; x in = initial guess,
 * out = solution, 16-byte aligned) and a cudaStream_t of the CURRENT device (NULL = the default stream).
 *   - Limit: n <= CGB_STATIC_MAX_ROWS (the range where cgb_cg_solve itself uses the one-kernel loop); a larger
 *     system is refused with CGB_ERR_UNSUPPORTED and a message that names the limit.  nnz is not limited by the
 *     work area (the matrix is the caller's).
 *   - The first call on each device looks up the work area's device address (once per device, with that device
 *     current).  Make that first call during start-up, before any stream capture begins.
 *   - One work area per device, shared by all calls, never cleared between them: each call resets what must start
 *     from zero itself.  Two calls that overlap on the same GPU (two host threads, two streams) race on it and give
 *     wrong results; run them one after the other.  The call returns after the solve has finished (it reads the
 *     final state back), so calls made from one thread never overlap.
 *   - Same kernels, stop rule and trip cap as cgb_cg_solve with CGB_LOOP_PERSISTENT (max_iter < 0: n + 1).
 *     stats (may be NULL): iterations, rr, rr0, converged, nan_stop, kernel_launches; the timing fields stay 0. */
#define CGB_STATIC_MAX_ROWS 131072
#define CGB_STATIC_SCRATCH_BYTES 4260096 /* (4 * CGB_STATIC_MAX_ROWS + 8224) doubles */
int cgb_cg_solve_static(void *stream, int64_t n, int64_t nnz, const double *d_vals, const int32_t *d_cols,
                        const int32_t *d_rowptr, const double *d_b, double *d_x, double abs_tol, int64_t max_iter,
                        cgb_cg_stats *stats);

/* replaces: CG::accuracy (src/CG.hpp:463-515): *rr = sum (b - A x)^2, *xx = sum x^2;
 * accuracy() = |rr / xx|. */
int cgb_residual(cgb_ctx *ctx, cgb_csr *A, const double *d_x, const double *d_b, double *rr, double *xx);
/* replaces: CG::memoryFootprint (src/CG.hpp:555-558) */
size_t cgb_memory_footprint(int64_t n, int64_t nnz);

/* ---- Matrix Market input -------------------------------------------------------------------
 * replaces: read_file (test/mm_reader.cpp:154-171) with the same dialect: banner ignored, the
 * line after the banner discarded, '%' lines skipped, size line unused, every off-diagonal
 * mirrored, entries sorted by (row, column), a row offset per distinct row id. */
typedef struct cgb_mm cgb_mm;
int cgb_mm_read(const char *path, cgb_mm **out);
int64_t cgb_mm_n(const cgb_mm *m);
int64_t cgb_mm_nnz(const cgb_mm *m);
const double *cgb_mm_vals(const cgb_mm *m);
const int32_t *cgb_mm_cols(const cgb_mm *m);
const int32_t *cgb_mm_rowptr(const cgb_mm *m);
int cgb_mm_free(cgb_mm *m);

/* ---- triplets (COO) -> device CSR ---------------------------------------------------------------
 * replaces: the CSR assembly inside the reference reader (test/mm_reader.cpp:45-107: vector of tuples, mirror,
 * std::sort, row walk) and the element-by-element HiFlow3 bridge (src/CG.hpp:104-151) with one upload and a
 * radix sort + scans on the GPU.  Output: rows ascending, columns ascending inside a row. */
enum {
  CGB_COO_MIRROR = 1,         /* add (j, i, v) for every off-diagonal (i, j, v): one stored triangle -> full matrix */
  CGB_COO_SUM_DUPLICATES = 2, /* equal (i, j) are summed in input order (finite-element assembly); else kept as they are */
  CGB_COO_COMPACT_ROWS = 4,   /* a row offset per DISTINCT row id: empty rows vanish, n = number of distinct rows */
  CGB_COO_ONE_BASED = 8,      /* indices start at 1 */
  CGB_COO_REFERENCE_DIALECT = 1 | 4 | 8 /* what read_file does to a Matrix Market body (test/mm_reader.cpp:62-104) */
};
/* n_rows: matrix dimension, or -1 = largest index + 1 (must be -1 with CGB_COO_COMPACT_ROWS). */
int cgb_csr_from_triplets(cgb_ctx *ctx, int64_t n_rows, int64_t count, const int32_t *h_rows, const int32_t *h_cols,
                          const double *h_vals, int flags, cgb_csr **out);
int cgb_csr_from_triplets_device(cgb_ctx *ctx, int64_t n_rows, int64_t count, const int32_t *d_rows, const int32_t *d_cols,
                                 const double *d_vals, int flags, cgb_csr **out);
/* Matrix Market file -> device CSR without a host CSR in between: the body is parsed on all host cores
 * (same dialect as cgb_mm_read / the reference's read_file), assembled on the GPU.  Same arrays as
 * cgb_mm_read + cgb_csr_create on well-formed files (no duplicate coordinates). */
int cgb_csr_from_mm_file(cgb_ctx *ctx, const char *path, cgb_csr **out);

/* ---- row-partitioned multi-GPU (new; the reference is single-device) -------------------------
 * One process per GPU.  Rank r owns the contiguous global rows [row_begin, row_end); its local
 * CSR (cgb_csr with n = row_end-row_begin rows and GLOBAL column indices) is split into interior
 * and halo parts.  Default transport: peer memory over NVLink (cudaIpc windows) inside the library's
 * own kernels -- boundary entries of p are stored straight into the owners' halo buffers, the two
 * scalars per iteration are all-reduced through per-rank slots and flags; for stencil-like blocks the
 * whole solve is ONE persistent cooperative kernel per rank.  NCCL carries the set-up exchanges and
 * is the fallback transport (CGB_DIST_TRANSPORT=nccl).  All cgb_dist_* calls are collective: every
 * rank makes the same calls in the same order.  A peer that does not answer within
 * CGB_PEER_TIMEOUT_MS (default 30000) makes the solve return CGB_ERR_NCCL and the handle unusable
 * (CGB_ERR_STATE afterwards): destroy it on every rank.
 * Two rules follow from the solve being collective.  (1) Every rank passes the same cgb_cg_opts: the first solve of a
 * matrix with a given option set checks this across the ranks (CGB_ERR_INVALID on a mismatch) and remembers the
 * loop it chose; later solves with that option set skip the check and its host barrier, so a rank that changes its
 * options alone afterwards waits in a collective the others never enter.  (2) Without that barrier the ranks must
 * enter a repeated solve within CGB_PEER_TIMEOUT_MS of each other: an application with long, uneven host phases
 * between solves (file output on one rank) raises the limit or synchronises its ranks itself before the call. */
#define CGB_NCCL_ID_BYTES 128
int cgb_nccl_unique_id(void *h_id128);
/* row_starts: nranks+1 global row offsets (same on every rank). */
int cgb_dist_create(cgb_ctx *ctx, int rank, int nranks, const void *h_id128, const int64_t *row_starts,
                    cgb_csr *A_local_global_cols, cgb_dist **out);
/* Replace the matrix of an existing handle (same communicator): re-runs the split and halo plan. */
int cgb_dist_set_matrix(cgb_dist *d, const int64_t *row_starts, cgb_csr *A_local_global_cols);
int cgb_dist_destroy(cgb_dist *d);
/* d_b, d_x: local row block.  stats->iterations etc. are global (identical on all ranks). */
int cgb_dist_cg_solve(cgb_dist *d, const double *d_b, double *d_x, double abs_tol, int64_t max_iter,
                      const cgb_cg_opts *opts, cgb_cg_stats *stats);
int cgb_dist_residual(cgb_dist *d, const double *d_x, const double *d_b, double *rr, double *xx);
int64_t cgb_dist_halo_count(const cgb_dist *d);
/* How the gathered vector travels for the current matrix: 0 = halo exchange (only the boundary entries the rank's
 * rows touch: stencils), 1 = all-gather (every rank stores its whole slice into every rank's full-length copy over
 * NVLink; chosen when a rank needs most of the remote entries anyway: random matrices).  CGB_DIST_ALLGATHER=0/1 forces it. */
int cgb_dist_exchange_mode(const cgb_dist *d);

/* Host-only halo plan (no GPU needed; used by the CPU tests of the partitioning logic):
 * given a local row block with GLOBAL column indices, returns the sorted distinct remote
 * columns (the receive list) grouped by owning rank.  h_recv_cols may be NULL to query the count. */
int cgb_plan_halo(int rank, int nranks, const int64_t *row_starts, int64_t nnz_local, const int32_t *h_cols,
                  int32_t *h_recv_cols, int64_t capacity, int64_t *recv_count, int64_t *recv_per_rank);

#ifdef __cplusplus
}
#endif
#endif /* CGB200_H */
