// dist_persistent.cuh — the row-partitioned CG loop as ONE cooperative kernel per rank (included by dist.cu).
//
// The launch-per-kernel loop of dist_impl.cuh pays five kernel boundaries and two last-CTA all-reduce round trips
// per trip (about 60 us of fixed cost at 8 GPUs: VERDICT r1, weak item 5).  Here a rank runs its whole solve in one
// kernel, the way persistent.cuh does on one GPU; a trip is
//
//   phase P   x += alpha_prev p (deferred) ; p = r + beta p                          (src/CG.hpp:390, :418)
//   -- grid barrier --
//   pack      boundary entries of p stored straight into the owners' halo buffers (NVLink), then this rank's
//             sequence flag raised in every destination window by the CTA that finishes the pack last
//   phase S   Ap = A_loc p through the TMA-staged CSR engine, partial p.Ap             (:374-379)
//             -- the halo transfer is in flight meanwhile --
//   -- grid barrier --
//   phase H   wait for the halo flags; Ap[row] += A_halo[row,:] . halo, partial p.Ap += p[row] * that
//   -- reduction: per-CTA partials -> the last CTA to arrive sums them in fixed order and stores the rank's total
//      into slot [rank] of EVERY rank's window; ALL CTAs of all ranks then wait for the P flags in their own window
//      and add the P slots in rank order.  One step is the grid barrier and the all-reduce at once, every thread
//      ends with the same bits, nothing is broadcast through memory. --          alpha = rr / p.Ap   (:385-386)
//   phase U   r -= alpha Ap, partial r.r                                              (:392, :406)
//   -- reduction (same) --   stop rule on the OLD rr (:396-404), beta = rr'/rr (:414), ++trip (:436)
//
// Sequence numbers (PeerSeq) continue those of the set-up kernels and are written back at the end, so the
// residual / set-up kernels that follow see a consistent window.  Slots are double-buffered on the parity of the
// reduction number: a rank can be at most one reduction ahead of the slowest (finishing reduction s needs every
// rank's contribution to s, which a rank only sends once ALL its CTAs have finished reading reduction s-1).  The
// same argument makes the single-buffered halo safe: nobody packs trip k+1 before everyone is past phase H of trip k.
// Data written by other CTAs inside the kernel (p, Ap) is read with plain (weak) loads after an acquire at the
// barrier, never through the non-coherent path.  Every wait carries a time limit and watches the state's
// comm_error flag, so one failing wait takes every CTA of the rank out of the loop instead of hanging the GPU.
#pragma once

#include "vec_tma.cuh"

#ifndef CGB_DIST_VEC_TMA
#define CGB_DIST_VEC_TMA 1 // the two vector passes through TMA-staged tiles (vec_tma.cuh); 0: grid-stride loops of 16-byte loads
#endif
#ifndef CGB_DIST_UNROLL_P
#define CGB_DIST_UNROLL_P 1 // independent 16-byte loads per array and thread in the p-update phase (C4 on 2 GPUs: 1 -> 537 it/s, 2 -> 518)
#endif
#ifndef CGB_DIST_UNROLL_U
#define CGB_DIST_UNROLL_U 1 // ... in the r-update phase (4 measured together with the 2 above)
#endif

namespace {

struct DistBar {                 // device words of the persistent loop, zeroed by the host before every launch
  unsigned long long arrive;     // grid barrier: arrivals so far (monotone)
  unsigned long long gen;        // grid barrier: generations completed
  unsigned long long pack;       // CTAs that finished packing (monotone)
  unsigned long long red;        // reduction ticket (monotone)
  unsigned long long phase_ns[10]; // CGB_DIST_TIMING: time CTA 0 spent in P, barrier, pack, S, (barrier +) halo flag wait, halo rows, all-reduce, U, all-reduce (sums over the trips)
};

struct DistLoopArgs {
  CsrView A;                     // A_loc: all local rows, owned columns
  int64_t nh;                    // rows that touch remote columns
  const int *h_rowptr, *h_cols;  // A_halo (columns index the halo buffer)
  const double *h_vals;
  const int *halo_rows;
  const int4 *hplan;             // halo rows grouped by the CTA whose tiles hold them: {row, first entry, end, 1} (nullptr: grid-stride
  const int *hoff;               // pass behind a grid barrier); hoff[b] .. hoff[b+1] = CTA b's share
  const double *halo;
  int64_t S;                     // entries this rank sends
  const int *send_idx;
  PeerSend ps;
  PeerRecv pr;
  PeerCtx pc;
  double *x, *r, *p, *Ap;
  CgState *st;
  double *partials;
  double *hist;
  DistBar *bar;
  long long timeout_cycles;
  int vec_tma;                   // the vector passes through TMA-staged tiles (long vectors) or grid-stride loops (short ones)
  int phase_timing;              // CGB_DIST_TIMING: CTA 0 stamps the phases (globaltimer)
  int side_pack;                 // the pack's fence / ticket / flags on the producer lane behind its first tiles (1) or inline before the tiles (0, A/B)
  unsigned jitter_ns;            // CGB_DIST_JITTER_NS (tests): every CTA sleeps a pseudo-random time below this before each phase
};

struct GatherWeak { // p is rewritten every trip by other CTAs of this very kernel: plain ld.global, ordered by the barrier's acquire
  const double *x;
  // an ordinary load: the kernel also WRITES p through another pointer, so the compiler cannot promote it to the
  // non-coherent path (checked in the SASS: LDG.E.64, no .CONSTANT), and it stays free to schedule the gathers
  // (an asm volatile with a memory clobber here cost the tile loop its overlap of shared-memory reads and gathers)
  __device__ __forceinline__ double operator()(int c) const { return x[c]; }
  struct At {
    const double *xr;
    __device__ __forceinline__ double operator()(int off) const { return xr[off]; }
  };
  __device__ __forceinline__ At at(int64_t row) const { return At{x + row}; }
};

__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Spin until *flag >= want (SYS = flag written by a peer GPU).  False on time-out or when another CTA reported one.
// The poll is a RELAXED load and the acquire comes once, after it succeeds: an acquire load per poll would
// invalidate the SM's L1 on every spin (CCTL.IVALL), under the feet of the co-resident CTAs that are still gathering.
template <bool SYS>
__device__ __forceinline__ unsigned long long dist_poll(const unsigned long long *p) {
  unsigned long long v;
  if (SYS) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  else asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
template <bool SYS>
__device__ __forceinline__ void dist_acquire_fence() {
  if (SYS) asm volatile("fence.acq_rel.sys;" ::: "memory");
  else asm volatile("fence.acq_rel.gpu;" ::: "memory");
}
template <bool SYS>
__device__ __forceinline__ bool dist_wait(const unsigned long long *flag, unsigned long long want, CgState *st, long long limit) {
  if (dist_poll<SYS>(flag) < want) {
    const long long t0 = clock64();
    int spins = 0;
    while (dist_poll<SYS>(flag) < want) {
      if ((++spins & 63) == 0) {
        if (*(volatile int *)&st->comm_error) return false;
        if (clock64() - t0 > limit) {
          *(volatile int *)&st->comm_error = 1;
          __threadfence();
          return false;
        }
      }
    }
  }
  // acquire = one acquire LOAD of the flag (it can only have grown).  A fence here would also have to drain this
  // thread's own outstanding stores system-wide (peer stores over NVLink): measured 30 us per trip slower on C2 at 2 GPUs.
  (void)(SYS ? ld_acquire_sys_u64(flag) : ld_acquire_gpu_u64(flag));
  return true;
}

// All threads of all CTAs; `gen` is the caller's count of barriers passed so far.  False: leave the loop.
__device__ __forceinline__ bool dist_grid_barrier(DistBar *bar, unsigned long long &gen, CgState *st, long long limit) {
  __shared__ int s_ok;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned long long t = atomicAdd(&bar->arrive, 1ull) + 1ull;
    bool ok = true;
    if (t == (gen + 1ull) * gridDim.x) st_release_gpu_u64(&bar->gen, gen + 1ull);
    else ok = dist_wait<false>(&bar->gen, gen + 1ull, st, limit);
    __threadfence();
    s_ok = ok ? 1 : 0;
  }
  gen += 1ull;
  __syncthreads();
  return s_ok != 0;
}

// Sum of `v` over all threads of all CTAs of ALL ranks, the same bits in every thread.  `nred` counts the reductions of
// this launch, `seq` the all-reduces on the window since it was created.  *ok is cleared on a failed wait.
__device__ __forceinline__ double dist_allreduce(double v, const DistLoopArgs &a, PeerWin *const *win, unsigned long long &nred,
                                                 unsigned long long &seq, double *red, double *slots, bool *ok) {
  __shared__ int s_last, s_okr;
  __shared__ double s_total;
  v = block_sum(v, red);
  if (threadIdx.x == 0) {
    *(volatile double *)&a.partials[blockIdx.x] = v;
    __threadfence();
    const unsigned long long t = atomicAdd(&a.bar->red, 1ull) + 1ull;
    s_last = (t == (nred + 1ull) * gridDim.x);
    s_okr = 1;
  }
  nred += 1ull;
  seq += 1ull;
  const int par = (int)(seq & 1ull);
  __syncthreads();
  if (s_last) { // this CTA saw every partial of the rank: fixed-order sum, then one store per destination window
    __threadfence();
    const double total = reduce_partials(a.partials, gridDim.x, red);
    if (threadIdx.x == 0) s_total = total;
    __syncthreads();
    if (threadIdx.x < a.pc.nranks) {
      PeerWin *w = win[threadIdx.x];
      st_relaxed_sys_f64(&w->slot[par][a.pc.rank], s_total);
      st_release_sys_u64(&w->sflag[par][a.pc.rank], seq);
    }
  }
  if (threadIdx.x < a.pc.nranks) {
    const PeerWin *me = win[a.pc.rank];
    if (dist_wait<true>(&me->sflag[par][threadIdx.x], seq, a.st, a.timeout_cycles)) slots[threadIdx.x] = ld_relaxed_sys_f64(&me->slot[par][threadIdx.x]);
    else {
      slots[threadIdx.x] = 0.0;
      s_okr = 0;
    }
  }
  __syncthreads();
  double sum = 0.0;
  for (int q = 0; q < a.pc.nranks; q++) sum += slots[q]; // rank order: identical on every rank
  if (!s_okr) *ok = false;
  __syncthreads(); // slots / s_last are reused by the next call
  return sum;
}

// The vector phases run on the SpMV's launch shape (3 CTAs of 288 threads per SM, less than half the threads the
// stand-alone vector kernels use): every thread keeps UNROLL independent 16-byte loads per array in flight instead.
__device__ __forceinline__ void dist_pupdate_body(int64_t n, const double *r, double *p, double *x, double alpha, double beta) {
  constexpr int UNROLL = CGB_DIST_UNROLL_P;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = n >> 1;
  const double2 *r2 = reinterpret_cast<const double2 *>(r);
  double2 *p2 = reinterpret_cast<double2 *>(p);
  double2 *x2 = reinterpret_cast<double2 *>(x);
  int64_t i = tid;
  for (; i + (UNROLL - 1) * nthr < n2; i += UNROLL * nthr) {
    double2 rv[UNROLL], pv[UNROLL], xv[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
      rv[u] = r2[i + u * nthr];
      pv[u] = p2[i + u * nthr];
      xv[u] = x2[i + u * nthr];
    }
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
      xv[u].x = fma(alpha, pv[u].x, xv[u].x);
      xv[u].y = fma(alpha, pv[u].y, xv[u].y);
      pv[u].x = fma(beta, pv[u].x, rv[u].x);
      pv[u].y = fma(beta, pv[u].y, rv[u].y);
      x2[i + u * nthr] = xv[u];
      p2[i + u * nthr] = pv[u];
    }
  }
  for (; i < n2; i += nthr) {
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

// Loads of data that every trip reads again and nothing in the kernel writes (halo plan, halo part of the matrix):
// non-coherent path, L2 evict-last hint so the matrix stream does not push them out between trips.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ int4 ld_keep_int4(const int4 *p, uint64_t pol) {
  int4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.s32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ double ld_keep_f64(const double *p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ int ld_keep_s32(const int *p, uint64_t pol) {
  int v;
  asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}

// Second half of the halo pack, run by the TMA engine's producer lane once its first tiles are in flight: order the
// CTA's peer stores (they precede the CTA barrier that opened the tile loop, so this thread's fence is cumulative over
// them), take the ticket, and in the CTA that finishes last raise this rank's flag in every destination window.
struct PackFinish {
  bool active;
  unsigned long long target, seq;
  DistBar *bar;
  PeerWin *const *win;
  const PeerSend *ps;
  int rank, nranks;
  __device__ __forceinline__ void operator()() const {
    if (!active) return;
    __threadfence_system();
    const unsigned long long t = atomicAdd(&bar->pack, 1ull) + 1ull;
    if (t != target) return;
    __threadfence_system(); // acquire side of the ticket: every CTA's stores precede the flags
    for (int q = 0; q < nranks; q++)
      if (ps->soff[q + 1] > ps->soff[q]) st_release_sys_u64(&win[q]->hflag[rank], seq);
  }
};

__device__ __forceinline__ unsigned long long dist_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// CTA 0, thread 0: add the time since the last stamp to phase `k`
#define DIST_STAMP(k)                                                   \
  if (a.phase_timing && blockIdx.x == 0 && threadIdx.x == 0) {          \
    const unsigned long long now_ = dist_now_ns();                      \
    a.bar->phase_ns[k] += now_ - t_phase;                               \
    t_phase = now_;                                                     \
  }

// Stress hook for the flag protocols (tests/test_dist_gpu.py, variant "jitter"): CTAs and ranks drift apart at random.
__device__ __forceinline__ void dist_jitter(const DistLoopArgs &a, unsigned salt) {
  if (a.jitter_ns == 0) return;
  unsigned h = (blockIdx.x + 1u) * 2654435761u ^ (salt + 1u) * 40503u ^ (unsigned)(a.pc.rank + 1) * 2246822519u;
  h ^= h >> 15;
  h *= 2246822519u;
  h ^= h >> 13;
  if (threadIdx.x == 0) __nanosleep(h % a.jitter_ns);
  __syncthreads();
}

// r -= alpha Ap, this thread's share of r.r.  No __restrict__ / const here: Ap is written by other CTAs of this very kernel,
// so its loads must stay on the coherent path (cg_update_body's qualifiers would allow ld.global.nc).
__device__ __forceinline__ double dist_update_body(int64_t n, double *r, double *Ap, double alpha) {
  constexpr int UNROLL = CGB_DIST_UNROLL_U;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = n >> 1;
  double2 *r2 = reinterpret_cast<double2 *>(r);
  double2 *q2 = reinterpret_cast<double2 *>(Ap);
  double acc = 0.0;
  int64_t i = tid;
  // the order in which a thread adds its squares is fixed by (n, grid): results stay bitwise reproducible
  for (; i + (UNROLL - 1) * nthr < n2; i += UNROLL * nthr) {
    double2 rv[UNROLL], qv[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
      rv[u] = r2[i + u * nthr];
      qv[u] = q2[i + u * nthr];
    }
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
      rv[u].x = fma(-alpha, qv[u].x, rv[u].x);
      rv[u].y = fma(-alpha, qv[u].y, rv[u].y);
      r2[i + u * nthr] = rv[u];
      acc = fma(rv[u].x, rv[u].x, acc);
      acc = fma(rv[u].y, rv[u].y, acc);
    }
  }
  for (; i < n2; i += nthr) {
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

template <int L>
__global__ void __launch_bounds__(kTmaThreads, 3) cg_dist_persistent_kernel(const __grid_constant__ DistLoopArgs a) {
  extern __shared__ __align__(128) unsigned char cgb_dyn_smem[];
  __shared__ double red[32];
  __shared__ double slots[kMaxRanks];
  __shared__ PeerWin *win[kMaxRanks];
  __shared__ int s_flag;
  CgState *st = a.st;
  if (threadIdx.x < a.pc.nranks) win[threadIdx.x] = a.pc.win[threadIdx.x];
  __syncthreads();
  // the set-up kernels armed the state; every thread carries its own copy of the scalars (same slots, same order => same bits)
  double rr = st->rr, alpha = 0.0, beta = 0.0;
  const double tol = st->tol;
  const long long max_iter = st->max_iter, hist_cap = st->hist_cap;
  const int record = st->record_history;
  long long it = 0;
  int done = st->done, nan_stop = 0;
  unsigned long long gen = 0, nred = 0, npack = 0;
  unsigned long long red_seq = a.pc.seq->red, halo_seq = a.pc.seq->halo;
  const int64_t n = a.A.n;
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  bool ok = true;
  unsigned long long t_phase = a.phase_timing ? dist_now_ns() : 0ull;
  while (!done && it < max_iter) {
    // ---- phase P ----
    dist_jitter(a, 4u * (unsigned)it);
    if (CGB_DIST_VEC_TMA && a.vec_tma) vec_tma_pupdate(n, a.r, a.p, a.x, alpha, beta, cgb_dyn_smem);
    else dist_pupdate_body(n, a.r, a.p, a.x, alpha, beta);
    DIST_STAMP(0)
    if (!dist_grid_barrier(a.bar, gen, st, a.timeout_cycles)) { ok = false; break; }
    DIST_STAMP(1)
    dist_jitter(a, 4u * (unsigned)it + 1u);
    // ---- pack: boundary entries of p to their owners, flags by the CTA that finishes last ----
    halo_seq += 1ull;
    bool pack_cta;
    unsigned long long pack_target;
    {
      // only as many CTAs as the send list needs (4 entries per thread) take part in the pack and its ticket; the
      // others go straight to their tiles
      const int64_t per_cta = (int64_t)blockDim.x * 4;
      const int pack_ctas = (int)min((int64_t)gridDim.x, (a.S + per_cta - 1) / per_cta);
      pack_cta = (int)blockIdx.x < pack_ctas;
      pack_target = npack + (unsigned long long)pack_ctas;
      if (pack_cta)
        for (int64_t k = gtid; k < a.S; k += (int64_t)pack_ctas * blockDim.x) {
          int q = 0;
          while (k >= a.ps.soff[q + 1]) q++;
          a.ps.dst[q][k - a.ps.soff[q]] = a.p[a.send_idx[k]];
        }
      npack += (unsigned long long)pack_ctas;
      if (!a.side_pack && pack_cta) { // A/B: finish the pack here, on the critical path
        __syncthreads();
        if (threadIdx.x == 0) PackFinish{true, pack_target, halo_seq, a.bar, win, &a.ps, a.pc.rank, a.pc.nranks}();
      }
      // the stores are on their way; the fence that orders them, the ticket and the flags run on an idle lane of the
      // producer warp while the tiles stream (PackFinish below), behind the CTA barrier that opens the tile loop
    }
    DIST_STAMP(2)
    // ---- phase S: local rows through the TMA tile engine ----
    double a0 = 0.0, a1 = 0.0;
    {
      GatherWeak g{a.p};
      EpiDotP e{a.Ap, a.p};
      const PackFinish side{pack_cta && a.side_pack, pack_target, halo_seq, a.bar, win, &a.ps, a.pc.rank, a.pc.nranks};
      spmv_rows_tma<L>(a.A.n, a.A.nnz, a.A.rowptr, a.A.cols, a.A.vals, g, e, blockIdx.x, gridDim.x, a0, a1, cgb_dyn_smem, side);
      spmv_tma_teardown(cgb_dyn_smem, kTmaConsumers / L);
    }
    DIST_STAMP(3)
    // A CTA that has the halo plan adds the halo part to the rows of ITS OWN tiles: Ap there was stored by this CTA's
    // consumers (visible after the teardown's CTA barrier), so no grid barrier stands between the two phases and a CTA
    // that is early starts on its halo rows while others still stream tiles.
    const bool own_rows = a.hplan != nullptr || a.nh == 0;
    if (!own_rows && !dist_grid_barrier(a.bar, gen, st, a.timeout_cycles)) { ok = false; break; }
    dist_jitter(a, 4u * (unsigned)it + 2u);
    // ---- phase H: halo rows once the neighbours' entries are here ----
    if (a.nh > 0) {
      if (threadIdx.x == 0) s_flag = 1;
      __syncthreads();
      if (threadIdx.x < a.pc.nranks && a.pr.from[threadIdx.x]) {
        if (!dist_wait<true>(&win[a.pc.rank]->hflag[threadIdx.x], halo_seq, st, a.timeout_cycles)) s_flag = 0;
      }
      __syncthreads();
      if (!s_flag) { ok = false; break; }
      DIST_STAMP(4)
      // HU rows per thread at a time.  The pass is a chain of dependent round trips (plan entry -> column and value ->
      // halo entry) over a few rows per thread: the plan entries of the next round are fetched while this one runs,
      // the first entry of every row (the only one of a 7-point slab) is issued for all HU rows before any is used,
      // and plan and matrix part carry an L2 evict-last hint (a few MB that every trip reads again).
      constexpr int HU = 2;
      if (own_rows) {
        const uint64_t keep = l2_policy_evict_last();
        const int64_t h1 = (int64_t)a.hoff[blockIdx.x + 1], hs = (int64_t)blockDim.x;
        int64_t k0 = (int64_t)a.hoff[blockIdx.x] + threadIdx.x;
        int4 e[HU];
#pragma unroll
        for (int u = 0; u < HU; u++) e[u] = k0 + u * hs < h1 ? ld_keep_int4(a.hplan + k0 + u * hs, keep) : make_int4(0, 0, 0, 0);
        for (; k0 < h1; k0 += HU * hs) {
          int4 nx[HU];
#pragma unroll
          for (int u = 0; u < HU; u++) {
            const int64_t kn = k0 + (HU + u) * hs;
            nx[u] = kn < h1 ? ld_keep_int4(a.hplan + kn, keep) : make_int4(0, 0, 0, 0);
          }
          double v0[HU];
          int c0[HU];
#pragma unroll
          for (int u = 0; u < HU; u++) {
            const bool ne = e[u].y < e[u].z;
            v0[u] = ne ? ld_keep_f64(a.h_vals + e[u].y, keep) : 0.0;
            c0[u] = ne ? ld_keep_s32(a.h_cols + e[u].y, keep) : 0;
          }
          double s4[HU];
#pragma unroll
          for (int u = 0; u < HU; u++) s4[u] = e[u].y < e[u].z ? fma(v0[u], __ldcg(a.halo + c0[u]), 0.0) : 0.0;
#pragma unroll
          for (int u = 0; u < HU; u++)
            for (int j = e[u].y + 1; j < e[u].z; j++) s4[u] = fma(ld_keep_f64(a.h_vals + j, keep), __ldcg(a.halo + ld_keep_s32(a.h_cols + j, keep)), s4[u]);
#pragma unroll
          for (int u = 0; u < HU; u++)
            if (e[u].w) {
              a.Ap[e[u].x] += s4[u];
              a0 = fma(a.p[e[u].x], s4[u], a0);
            }
#pragma unroll
          for (int u = 0; u < HU; u++) e[u] = nx[u];
        }
      } else {
        for (int64_t k0 = gtid; k0 < a.nh; k0 += HU * nthr) {
          int jb[HU], je[HU], row[HU];
#pragma unroll
          for (int u = 0; u < HU; u++) {
            const int64_t k = k0 + u * nthr;
            const bool in = k < a.nh;
            jb[u] = in ? a.h_rowptr[k] : 0;
            je[u] = in ? a.h_rowptr[k + 1] : 0;
            row[u] = in ? a.halo_rows[k] : 0;
          }
          double s4[HU] = {};
#pragma unroll
          for (int u = 0; u < HU; u++)
            for (int j = jb[u]; j < je[u]; j++) s4[u] = fma(a.h_vals[j], __ldcg(a.halo + a.h_cols[j]), s4[u]);
#pragma unroll
          for (int u = 0; u < HU; u++)
            if (k0 + u * nthr < a.nh) {
              a.Ap[row[u]] += s4[u];
              a0 = fma(a.p[row[u]], s4[u], a0);
            }
        }
      }
    }
    DIST_STAMP(5)
    const double pAp = dist_allreduce(a0, a, win, nred, red_seq, red, slots, &ok);
    if (!ok) break;
    DIST_STAMP(6)
    alpha = rr / pAp; // src/CG.hpp:385-386
    // ---- phase U ----
    dist_jitter(a, 4u * (unsigned)it + 3u);
    const double acc = (CGB_DIST_VEC_TMA && a.vec_tma) ? vec_tma_update(n, a.r, a.Ap, alpha, cgb_dyn_smem) : dist_update_body(n, a.r, a.Ap, alpha);
    DIST_STAMP(7)
    const double rr_new = dist_allreduce(acc, a, win, nred, red_seq, red, slots, &ok);
    if (!ok) break;
    DIST_STAMP(8)
    if (isnan(rr) || sqrt(rr) <= tol) { // the rr from BEFORE this trip's update (one-trip lag, src/CG.hpp:396-404)
      done = 1;
      nan_stop = isnan(rr) ? 1 : 0;
    }
    beta = rr_new / rr;
    rr = rr_new;
    it++;
    if (record && it < hist_cap && blockIdx.x == 0 && threadIdx.x == 0) a.hist[it] = rr_new;
  }
  // the x update of the last trip (its p is complete: every CTA passed the barrier after phase P)
  if (it > 0 && ok) {
    const int64_t n2 = n >> 1;
    double2 *x2 = reinterpret_cast<double2 *>(a.x);
    const double2 *p2 = reinterpret_cast<const double2 *>(a.p);
    for (int64_t i = gtid; i < n2; i += nthr) {
      double2 xv = x2[i];
      const double2 pv = p2[i];
      xv.x = fma(alpha, pv.x, xv.x);
      xv.y = fma(alpha, pv.y, xv.y);
      x2[i] = xv;
    }
    if ((n & 1) && gtid == 0) a.x[n - 1] = fma(alpha, a.p[n - 1], a.x[n - 1]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->rr = rr;
    st->alpha = alpha;
    st->beta = beta;
    st->iter = it;
    st->done = ok ? done : 1;
    st->nan_stop = nan_stop;
    if (!ok) st->comm_error = 1;
    a.pc.seq->red = red_seq;
    a.pc.seq->halo = halo_seq;
  }
}

// Host side: is the persistent loop applicable on THIS rank (the ranks then agree collectively), and its launch.
bool dist_persistent_eligible(const cgb_dist *d, const cgb_cg_opts &o) {
  if (const char *e = getenv("CGB_DIST_PERSISTENT"))
    if (atoi(e) == 0) return false;
  if (!d->peer || !d->ctx->coop_ok || d->allgather) return false;
  if (o.loop_mode != CGB_LOOP_AUTO && o.loop_mode != CGB_LOOP_PERSISTENT) return false;
  if (o.time_kernels || o.jacobi || o.residual_replace_every > 0 || o.fuse_p_update) return false;
  const cgb_csr *A = d->Aloc;
  if (A->n <= 0 || csr_engine(A) != kEngTma) return false;
  // the halo rows are served by a plain row-per-thread pass: fine for stencil slabs (a few entries per boundary row), not
  // for blocks whose halo part is a real share of the matrix (random matrices: those keep the launch-per-kernel loop)
  if (d->Ahalo && d->Ahalo->nnz * 2 > A->nnz) return false;
  return true;
}

// ---- halo rows grouped by owning CTA (tile t of the SpMV engine belongs to CTA t % grid) ----
__global__ void hplan_key_kernel(int64_t nh, const int *halo_rows, int tile_rows, int grid, int *key, int *val) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nh) return;
  key[k] = (int)((halo_rows[k] / tile_rows) % grid);
  val[k] = (int)k;
}
__global__ void hplan_off_kernel(int64_t nh, const int *key_sorted, int grid, int *hoff) { // hoff[b] = first k with key >= b
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > grid) return;
  int64_t lo = 0, hi = nh;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (key_sorted[mid] < b) lo = mid + 1; else hi = mid;
  }
  hoff[b] = (int)lo;
}

__global__ void hplan_fill_kernel(int64_t nh, const int *perm, const int *halo_rows, const int *h_rowptr, int4 *plan) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nh) return;
  const int k = perm[j];
  plan[j] = make_int4(halo_rows[k], h_rowptr[k], h_rowptr[k + 1], 1);
}

struct HplanBuf { // freed on scope exit
  void *p = nullptr;
  ~HplanBuf() { cudaFree(p); }
  template <class T> T *as() { return static_cast<T *>(p); }
};

int dist_halo_plan(cgb_dist *d, int grid, int tile_rows) {
  if (d->hplan && d->hplan_grid == grid && d->hplan_tile == tile_rows) return CGB_OK;
  cudaStream_t s = d->ctx->stream;
  cudaFree(d->hplan);
  cudaFree(d->hoff);
  d->hplan = nullptr;
  d->hoff = nullptr;
  d->hplan_grid = d->hplan_tile = 0;
  const int64_t nh = d->nh;
  if (nh <= 0) return CGB_OK;
  CGB_CUDA(cudaMalloc(&d->hplan, sizeof(int4) * (size_t)nh));
  CGB_CUDA(cudaMalloc(&d->hoff, sizeof(int32_t) * (size_t)(grid + 1)));
  HplanBuf key, key_s, val, perm, tmp;
  CGB_CUDA(cudaMalloc(&perm.p, sizeof(int) * (size_t)nh));
  CGB_CUDA(cudaMalloc(&key.p, sizeof(int) * (size_t)nh));
  CGB_CUDA(cudaMalloc(&key_s.p, sizeof(int) * (size_t)nh));
  CGB_CUDA(cudaMalloc(&val.p, sizeof(int) * (size_t)nh));
  hplan_key_kernel<<<(unsigned)((nh + 255) / 256), 256, 0, s>>>(nh, d->halo_rows, tile_rows, grid, key.as<int>(), val.as<int>());
  CGB_CUDA(cudaGetLastError());
  int bits = 1;
  while ((1 << bits) < grid) bits++;
  size_t t = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, t, key.as<int>(), key_s.as<int>(), val.as<int>(), perm.as<int>(), (int)nh, 0, bits, s);
  CGB_CUDA(cudaMalloc(&tmp.p, t + 16));
  cub::DeviceRadixSort::SortPairs(tmp.p, t, key.as<int>(), key_s.as<int>(), val.as<int>(), perm.as<int>(), (int)nh, 0, bits, s); // stable
  hplan_off_kernel<<<(unsigned)((grid + 1 + 255) / 256), 256, 0, s>>>(nh, key_s.as<int>(), grid, d->hoff);
  hplan_fill_kernel<<<(unsigned)((nh + 255) / 256), 256, 0, s>>>(nh, perm.as<int>(), d->halo_rows, d->Ahalo->rowptr, d->hplan);
  CGB_CUDA(cudaGetLastError());
  CGB_CUDA(cudaStreamSynchronize(s));
  d->hplan_grid = grid;
  d->hplan_tile = tile_rows;
  return CGB_OK;
}

template <int L>
int dist_persistent_launch_L(cgb_dist *d, double *d_x, long long timeout_cycles) {
  cgb_csr *A = d->Aloc;
  Workspace &w = A->ws;
  cudaStream_t s = d->ctx->stream;
  auto kernel = cg_dist_persistent_kernel<L>;
  const size_t smem = tma_smem_bytes(kTmaConsumers / L);
  CGB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kTmaThreads, smem));
  if (occ < 1) return fail(CGB_ERR_CUDA, "cg_dist_persistent_kernel: does not fit an SM");
  int grid = std::min(occ * d->ctx->sm_count, kMaxPartials);
  if (const char *e = getenv("CGB_DIST_PERSISTENT_CTAS")) grid = std::max(1, std::min(grid, atoi(e)));
  if (!d->bar) CGB_CUDA(cudaMalloc(&d->bar, sizeof(DistBar)));
  DistLoopArgs a;
  memset(&a, 0, sizeof(a));
  a.A = csr_view(A);
  a.nh = d->nh;
  if (d->Ahalo) {
    a.h_rowptr = d->Ahalo->rowptr;
    a.h_cols = d->Ahalo->cols;
    a.h_vals = d->Ahalo->vals;
  }
  a.halo_rows = d->halo_rows;
  a.halo = d->halo;
  bool own_rows = true; // CGB_DIST_OWN_HALO=0: the halo rows as a grid-stride pass behind a grid barrier (A/B)
  if (const char *e = getenv("CGB_DIST_OWN_HALO")) own_rows = atoi(e) != 0;
  if (own_rows && d->nh > 0) {
    CGB_TRY(dist_halo_plan(d, grid, kTmaConsumers / L));
    a.hplan = d->hplan;
    a.hoff = d->hoff;
  }
  a.S = d->S;
  a.send_idx = d->send_idx;
  a.ps = d->send;
  a.pr = d->recv;
  a.pc = d->pctx();
  a.x = d_x;
  a.r = w.r;
  a.p = w.p0;
  a.Ap = w.Ap;
  a.st = w.st;
  a.partials = w.partials;
  a.hist = w.hist;
  a.bar = static_cast<DistBar *>(d->bar);
  a.timeout_cycles = timeout_cycles;
  // TMA-staged vector passes want >= 8 tiles per CTA (C4 on 2 GPUs: 570 against 547 it/s; C2 on 2 GPUs, 2 tiles per
  // CTA: 15.7k against 17.6k); CGB_VEC_TMA=0/1 forces it
  a.vec_tma = A->n >= (int64_t)grid * 8 * kVecTile;
  if (const char *e = getenv("CGB_VEC_TMA")) a.vec_tma = atoi(e) != 0;
  a.phase_timing = getenv("CGB_DIST_TIMING") != nullptr;
  a.side_pack = 1;
  if (const char *e = getenv("CGB_DIST_SIDE_PACK")) a.side_pack = atoi(e) != 0;
  if (const char *e = getenv("CGB_DIST_JITTER_NS")) a.jitter_ns = (unsigned)std::max(0, atoi(e));
  CGB_CUDA(cudaMemsetAsync(d->bar, 0, sizeof(DistBar), s));
  void *args[] = {&a};
  CGB_CUDA(cudaLaunchCooperativeKernel((void *)kernel, dim3(grid), dim3(kTmaThreads), args, smem, s));
  return CGB_OK;
}

int dist_persistent_launch(cgb_dist *d, double *d_x, long long timeout_cycles) {
  switch (d->Aloc->lanes_tma) {
  case 1: return dist_persistent_launch_L<1>(d, d_x, timeout_cycles);
  case 2: return dist_persistent_launch_L<2>(d, d_x, timeout_cycles);
  case 4: return dist_persistent_launch_L<4>(d, d_x, timeout_cycles);
  case 8: return dist_persistent_launch_L<8>(d, d_x, timeout_cycles);
  case 16: return dist_persistent_launch_L<16>(d, d_x, timeout_cycles);
  default: return dist_persistent_launch_L<32>(d, d_x, timeout_cycles);
  }
}

} // namespace
