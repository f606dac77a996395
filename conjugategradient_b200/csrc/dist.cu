// dist.cu — row-partitioned multi-GPU CG (new: the reference is single-device).
//
// One process per GPU.  Rank r owns global rows [row_starts[r], row_starts[r+1]).  Its row block
// is split once, on the device, into
//   A_loc  : all local rows, only the columns it owns (indices rebased to the block), and
//   A_halo : only the rows that touch remote columns, indices into a compact halo buffer.
// Per iteration (peer transport, the default; peer.cuh): p-update -> pack: the boundary entries of p
// are stored into the owners' halo buffers over NVLink and the owners' flags raised -> A_loc SpMV
// (the transfer overlaps it) -> halo part: wait for the flags, add the A_halo correction to Ap and
// p.Ap -> all-reduce p.Ap through per-rank slots and flags -> r update + r.r, all-reduced the same
// way, stop rule.  No library collective in the loop.  Three loops run this (chosen collectively):
//   dist_persistent.cuh  the whole solve as ONE cooperative kernel per rank (stencil-like blocks)
//   all-gather mode      blocks coupled to most of the vector keep their global columns; every rank
//                        stores its slice of p into every rank's full-length copy (dist_setup_allgather)
//   dist_enqueue_iteration  five launches per trip inside a CUDA graph (everything else, NCCL transport)
// NCCL transport (fallback when cudaIpc peer mapping is unavailable, CGB_DIST_TRANSPORT=nccl, and
// the comparison point): ncclSend/Recv on a side stream + two ncclAllReduce + a state kernel.
// NCCL is resolved with dlopen so single-GPU users never need it; it always carries the set-up
// exchanges (receive lists, IPC handles).
#include "launch.cuh"
#include "peer.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <vector>

using namespace cgb;

// ---- NCCL through dlopen ------------------------------------------------------------------------
namespace {
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int nccl_load() {
  if (g_nccl.handle) return CGB_OK;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(CGB_ERR_NCCL, "cannot load libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                        \
  *(void **)(&g_nccl.field) = dlsym(h, name);                                   \
  if (!g_nccl.field) return fail(CGB_ERR_NCCL, "libnccl lacks symbol %s", name)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(AllGather, "ncclAllGather");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.handle = h;
  return CGB_OK;
}

#define CGB_NCCL(expr)                                                                                   \
  do {                                                                                                   \
    ncclResult_t _r = (expr);                                                                            \
    if (_r != ncclSuccess)                                                                               \
      return fail(CGB_ERR_NCCL, "%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(_r), __FILE__, __LINE__); \
  } while (0)
} // namespace

namespace cgb {
constexpr int64_t kHaloEngineMinNnz = 1 << 20; // below this the one-pass halo kernel is used
struct PeerSend { // where the entries of the packed send list go: slice q -> dst[q][0 .. soff[q+1]-soff[q])
  long long soff[kMaxRanks + 1];
  double *dst[kMaxRanks];
};
struct PeerRecv {
  int from[kMaxRanks]; // 1 where this rank receives halo entries from that rank
};
struct PeerGather { // all-gather mode: where this rank's slice goes in every rank's full-length copy (dst[rank] = local copy)
  double *dst[kMaxRanks];
};
} // namespace cgb

struct cgb_dist {
  cgb_ctx *ctx = nullptr;
  int rank = 0, nranks = 1;
  ncclComm_t comm = nullptr;
  std::vector<int64_t> row_starts;
  int64_t n = 0;        // local rows
  cgb_csr *Aloc = nullptr;
  cgb_csr *Ahalo = nullptr;   // nh rows, columns index the halo buffer
  int32_t *halo_rows = nullptr; // nh local row ids
  int64_t nh = 0;
  int64_t H = 0;              // halo entries received
  int64_t S = 0;              // entries sent
  std::vector<int64_t> rcount, roff, scount, soff;
  int32_t *send_idx = nullptr; // local indices to pack, grouped by destination rank
  double *send_buf = nullptr;
  double *halo = nullptr;
  double *scal = nullptr;      // [0] pAp local, [1] pAp global, [2] rr local, [3] rr global, [4..5] residual
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_pack = nullptr, ev_halo = nullptr;
  // peer transport (peer.cuh)
  bool peer = false;
  size_t halo_bytes = 0;                // size of the halo allocation (IPC-exported)
  PeerWin *win = nullptr;               // this rank's window
  std::vector<PeerWin *> win_of;        // all ranks' windows as mapped here (win_of[rank] == win)
  PeerWin **d_win = nullptr;            // the same, on the device
  PeerSeq *seq = nullptr;
  std::vector<double *> halo_of;        // peers' halo buffers as mapped here (nullptr where nothing is sent)
  PeerSend send;                        // destination table of the pack kernel
  PeerRecv recv;                        // which ranks this rank receives from
  // all-gather mode (dense coupling: most remote entries are needed): the block keeps its GLOBAL columns and
  // gathers from a full-length copy of the vector that every rank fills with its slice over NVLink
  bool allgather = false;
  double *pfull = nullptr;              // n_global doubles (IPC-exported)
  size_t pfull_bytes = 0;
  std::vector<double *> pfull_of;       // peers' copies as mapped here
  PeerGather gather;
  void *bar = nullptr;                  // DistBar of the persistent loop (dist_persistent.cuh)
  // persistent loop: halo rows grouped by the CTA whose SpMV tiles hold them, stable order; one int4 per row
  // {local row, first entry, one past the last entry in A_halo, 1}; hoff[b] .. hoff[b+1]: CTA b's share.  Valid for
  // this grid size and tile height.
  int4 *hplan = nullptr;
  int32_t *hoff = nullptr;
  int hplan_grid = 0, hplan_tile = 0;
  int loop_key = 0, loop_persistent = 0; // the collective loop decision of the last solve's option set (0: none yet for this matrix)
  bool dead = false;                    // a peer wait timed out: windows and sequence numbers are inconsistent across ranks
  PeerCtx pctx() const { return PeerCtx{rank, nranks, d_win, seq}; }
};

namespace {

// ---- setup kernels (one-off) ----------------------------------------------------------------------
__global__ void count_split_kernel(int64_t n, const int *rowptr, const int *cols, int lo, int hi, int *cnt_loc,
                                   int *cnt_rem) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int l = 0, r = 0;
  for (int j = rowptr[i]; j < rowptr[i + 1]; j++) {
    const int c = cols[j];
    if (c >= lo && c < hi) l++; else r++;
  }
  cnt_loc[i] = l;
  cnt_rem[i] = r;
}

struct IsRemote {
  int lo, hi;
  __host__ __device__ bool operator()(const int &c) const { return c < lo || c >= hi; }
};
struct NonZero {
  __host__ __device__ bool operator()(const int &c) const { return c != 0; }
};

__device__ __forceinline__ int lower_bound_dev(const int *a, int n, int key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// halo_slot[i] = index of row i among the rows that have remote entries (valid where cnt_rem[i] > 0)
__global__ void split_fill_kernel(int64_t n, const int *rowptr, const int *cols, const double *vals, int lo, int hi,
                                  const int *rowptr_loc, int *cols_loc, double *vals_loc, const int *cnt_rem,
                                  const int *halo_slot, const int *rowptr_halo, int *cols_halo, double *vals_halo,
                                  int *halo_rows, const int *recv_cols, int H) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int ol = rowptr_loc[i];
  int oh = 0;
  const bool has = cnt_rem[i] > 0;
  if (has) {
    const int slot = halo_slot[i];
    oh = rowptr_halo[slot];
    halo_rows[slot] = (int)i;
  }
  for (int j = rowptr[i]; j < rowptr[i + 1]; j++) {
    const int c = cols[j];
    if (c >= lo && c < hi) {
      cols_loc[ol] = c - lo;
      vals_loc[ol++] = vals[j];
    } else {
      cols_halo[oh] = lower_bound_dev(recv_cols, H, c);
      vals_halo[oh++] = vals[j];
    }
  }
}

__global__ void rebase_kernel(int *idx, int64_t n, int lo) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) idx[i] -= lo;
}

// ---- per-iteration kernels ------------------------------------------------------------------------
__global__ void pack_kernel(int64_t S, const int *__restrict__ send_idx, const double *__restrict__ p,
                            double *__restrict__ send_buf, const CgState *st, int gated) {
  if (gated && !cg_active(st)) return;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = tid; k < S; k += nthr) send_buf[k] = p[send_idx[k]];
}

// y[halo_rows[k]] += sum_j vals[j] * halo[cols[j]]; optionally the p.Ap correction:
// out[0] = base[0] + sum_k p[row] * (halo part of row).  One CTA (the halo part is tiny).
__global__ void __launch_bounds__(kBlock) halo_spmv_kernel(int64_t nh, const int *__restrict__ rowptr,
                                                            const int *__restrict__ cols,
                                                            const double *__restrict__ vals,
                                                            const int *__restrict__ halo_rows,
                                                            const double *__restrict__ halo, double *y,
                                                            const double *p, const double *base, double *out,
                                                            const CgState *st, int gated, double sign) {
  if (gated && !cg_active(st)) return;
  __shared__ double red[32];
  double acc = 0.0;
  for (int64_t k = threadIdx.x; k < nh; k += blockDim.x) {
    double s = 0.0;
    for (int j = rowptr[k]; j < rowptr[k + 1]; j++) s = fma(vals[j], halo[cols[j]], s);
    const int row = halo_rows[k];
    y[row] += sign * s;
    if (p) acc = fma(p[row], s, acc);
  }
  if (out) {
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) out[0] = (base ? base[0] : 0.0) + acc;
  }
}

// Multi-CTA flavour for large halos: partial sums per CTA, last CTA finalises.
__global__ void __launch_bounds__(kBlock) halo_spmv_multi_kernel(int64_t nh, const int *__restrict__ rowptr,
                                                                  const int *__restrict__ cols,
                                                                  const double *__restrict__ vals,
                                                                  const int *__restrict__ halo_rows,
                                                                  const double *__restrict__ halo, double *y,
                                                                  const double *p, const double *base, double *out,
                                                                  CgState *st, int gated, double *partials) {
  if (gated && !cg_active(st)) return;
  __shared__ double red[32];
  double acc = 0.0;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = tid; k < nh; k += nthr) {
    double s = 0.0;
    for (int j = rowptr[k]; j < rowptr[k + 1]; j++) s = fma(vals[j], halo[cols[j]], s);
    const int row = halo_rows[k];
    y[row] += s;
    if (p) acc = fma(p[row], s, acc);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double s = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0 && out) out[0] = (base ? base[0] : 0.0) + s;
  }
}

// r = b - t, partial r.r -> *local_out            (dist flavour of K0, src/CG.hpp:324-341)
__global__ void __launch_bounds__(kBlock) dist_init_kernel(int64_t n, const double *__restrict__ b,
                                                            const double *__restrict__ t, double *__restrict__ r,
                                                            CgState *st, double *partials, double *local_out) {
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int64_t i = tid; i < n; i += nthr) {
    const double ri = b[i] - t[i];
    r[i] = ri;
    acc = fma(ri, ri, acc);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double s = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) *local_out = s;
  }
}

__global__ void dist_arm_kernel(CgState *st, const double *rr_global, double *hist) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const double rr = *rr_global;
  st->rr = rr;
  st->rr0 = rr;
  st->pAp = 0.0;
  st->alpha = 0.0;
  st->beta = 0.0;
  st->iter = 0;
  st->done = 0;
  st->nan_stop = 0;
  if (st->record_history && st->hist_cap > 0) hist[0] = rr;
}

// sum (b - t)^2 and sum x^2 -> out[0], out[1] (local)        (src/CG.hpp:487-498)
__global__ void __launch_bounds__(kBlock) dist_resid_kernel(int64_t n, const double *__restrict__ b,
                                                             const double *__restrict__ t,
                                                             const double *__restrict__ x, CgState *st,
                                                             double *partials, double *out) {
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double a0 = 0.0, a1 = 0.0;
  for (int64_t i = tid; i < n; i += nthr) {
    const double a = b[i] - t[i];
    a0 = fma(a, a, a0);
    a1 = fma(x[i], x[i], a1);
  }
  a0 = block_sum(a0, red);
  a1 = block_sum(a1, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = a0;
    partials[kMaxPartials + blockIdx.x] = a1;
  }
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double s0 = reduce_partials(partials, gridDim.x, red);
    const double s1 = reduce_partials(partials + kMaxPartials, gridDim.x, red);
    if (threadIdx.x == 0) {
      out[0] = s0;
      out[1] = s1;
    }
  }
}

// ---- peer transport kernels (peer.cuh) ---------------------------------------------------------------
__device__ __forceinline__ void peer_fail(CgState *st) { // a wait timed out: stop the loop, tell the host
  st->comm_error = 1;
  st->done = 1;
}

// Boundary entries of `vec` go straight into the owners' halo buffers (NVLink stores); the last CTA
// raises this rank's flag in every destination window.
__global__ void __launch_bounds__(kBlock) peer_pack_kernel(int64_t S, const int *__restrict__ send_idx,
                                                            const double *__restrict__ vec, PeerSend ps, PeerCtx pc,
                                                            CgState *st, int gated) {
  if (gated && !cg_active(st)) return;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = tid; k < S; k += nthr) {
    int q = 0;
    while (k >= ps.soff[q + 1]) q++;
    ps.dst[q][k - ps.soff[q]] = vec[send_idx[k]];
  }
  __threadfence_system();
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const unsigned long long s = pc.seq->halo + 1;
    const int q = threadIdx.x;
    if (q < pc.nranks && ps.soff[q + 1] > ps.soff[q]) {
      __threadfence_system();
      st_release_sys_u64(&pc.win[q]->hflag[pc.rank], s);
    }
    __syncthreads();
    if (threadIdx.x == 0) pc.seq->halo = s;
  }
}

// y[halo_rows[k]] += A_halo[k,:] . halo after the halos have arrived.
// MODE 1 (CG loop): also p.Ap = base + correction, all-reduced over the ranks, alpha = rr / p.Ap.
// MODE 0 (set-up, residual): correction only.
template <int MODE>
__global__ void __launch_bounds__(kBlock) peer_halo_kernel(int64_t nh, const int *__restrict__ rowptr,
                                                            const int *__restrict__ cols,
                                                            const double *__restrict__ vals,
                                                            const int *__restrict__ halo_rows, const double *halo,
                                                            double *y, const double *p, const double *base, PeerCtx pc,
                                                            PeerRecv pr, CgState *st, int gated, double *partials) {
  if (gated && !cg_active(st)) return;
  __shared__ double red[32];
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < pc.nranks && pr.from[threadIdx.x]) {
    if (!peer_wait(&pc.win[pc.rank]->hflag[threadIdx.x], pc.seq->halo)) s_ok = 0;
  }
  __syncthreads();
  const bool ok_wait = s_ok != 0;
  double acc = 0.0;
  if (ok_wait) {
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = tid; k < nh; k += nthr) {
      double s = 0.0;
      for (int j = rowptr[k]; j < rowptr[k + 1]; j++) s = fma(vals[j], __ldcg(halo + cols[j]), s);
      const int row = halo_rows[k];
      y[row] += s;
      if (MODE == 1) acc = fma(p[row], s, acc);
    }
  } else if (threadIdx.x == 0) {
    peer_fail(st);
  }
  if (MODE == 1) {
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    if (publish_and_elect(&st->ticket, gridDim.x)) {
      const double corr = reduce_partials(partials, gridDim.x, red);
      bool ok = true;
      const double pAp = peer_allreduce(pc, base[0] + corr, &ok);
      if (threadIdx.x == 0) {
        if (!ok) peer_fail(st);
        st->pAp = pAp;
        st->alpha = st->rr / pAp; // src/CG.hpp:385-386
      }
    }
  }
}

// Same job as peer_halo_kernel, for halo parts that are a real share of the matrix (irregular matrices:
// at 8 ranks 7/8 of a random row block's entries are remote): A_halo goes through the regular row
// engines (TMA-staged CSR tiles or SELL slices) with the halo buffer as the gathered vector.
struct GatherCg { // coherent at L2: the halo buffer is written by peers until the flags are up
  const double *x;
  __device__ __forceinline__ double operator()(int c) const { return __ldcg(x + c); }
  struct At {
    const double *xr;
    __device__ __forceinline__ double operator()(int off) const { return __ldcg(xr + off); }
  };
  __device__ __forceinline__ At at(int64_t row) const { return At{x + row}; }
};
struct EpiHaloAdd { // y[rows[k]] += s ; acc0 += p[rows[k]] * s
  double *y;
  const double *p;
  const int *rows;
  __device__ __forceinline__ void row(int64_t k, double s, double &a0, double &) const {
    const int r = rows[k];
    y[r] += s;
    if (p) a0 = fma(p[r], s, a0);
  }
};

template <int L, int ENG, int MODE>
__global__ void CGB_BOUNDS(ENG) peer_halo_engine_kernel(CsrView H, const int *__restrict__ halo_rows, const double *halo,
                                                        double *y, const double *p, const double *base, PeerCtx pc, PeerRecv pr,
                                                        CgState *st, int gated, double *partials) {
  if (gated && !cg_active(st)) return;
  __shared__ double red[32];
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < pc.nranks && pr.from[threadIdx.x]) {
    if (!peer_wait(&pc.win[pc.rank]->hflag[threadIdx.x], pc.seq->halo)) s_ok = 0;
  }
  __syncthreads();
  double a0 = 0.0, a1 = 0.0;
  if (s_ok) { // uniform over the CTA: the engines synchronise inside
    GatherCg g{halo};
    EpiHaloAdd e{y, MODE == 1 ? p : nullptr, halo_rows};
    spmv_rows<L, ENG>(H, g, e, a0, a1);
  } else if (threadIdx.x == 0) {
    peer_fail(st);
  }
  if (MODE == 1) {
    a0 = block_sum(a0, red);
    if (threadIdx.x == 0) partials[blockIdx.x] = a0;
    if (publish_and_elect(&st->ticket, gridDim.x)) {
      const double corr = reduce_partials(partials, gridDim.x, red);
      bool ok = true;
      const double pAp = peer_allreduce(pc, base[0] + corr, &ok);
      if (threadIdx.x == 0) {
        if (!ok) peer_fail(st);
        st->pAp = pAp;
        st->alpha = st->rr / pAp;
      }
    }
  }
}

// K2 with the r.r all-reduce and the state advance in the last CTA.
__global__ void __launch_bounds__(kBlock) peer_update_kernel(int64_t n, double *__restrict__ r,
                                                              const double *__restrict__ Ap, CgState *st,
                                                              double *partials, double *hist, PeerCtx pc) {
  if (!cg_active(st)) return;
  __shared__ double red[32];
  double acc = cg_update_body(n, r, Ap, st->alpha);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double local = reduce_partials(partials, gridDim.x, red);
    bool ok = true;
    const double rr_new = peer_allreduce(pc, local, &ok);
    if (threadIdx.x == 0) {
      if (!ok) peer_fail(st);
      else cg_advance_state(st, rr_new, hist);
    }
  }
}

// r = b - t, rr = all-reduce(r.r), arm the state          (dist flavour of K0, src/CG.hpp:324-341)
__global__ void __launch_bounds__(kBlock) peer_init_kernel(int64_t n, const double *__restrict__ b,
                                                            const double *__restrict__ t, double *__restrict__ r,
                                                            CgState *st, double *partials, double *hist, PeerCtx pc) {
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int64_t i = tid; i < n; i += nthr) {
    const double ri = b[i] - t[i];
    r[i] = ri;
    acc = fma(ri, ri, acc);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double local = reduce_partials(partials, gridDim.x, red);
    bool ok = true;
    const double rr = peer_allreduce(pc, local, &ok);
    if (threadIdx.x == 0) {
      st->rr = rr;
      st->rr0 = rr;
      st->pAp = 0.0;
      st->alpha = 0.0;
      st->beta = 0.0;
      st->iter = 0;
      st->nan_stop = 0;
      if (st->record_history && st->hist_cap > 0) hist[0] = rr;
      if (ok) st->done = 0;
      else st->comm_error = 1; // stays disarmed
    }
  }
}

// sum (b - t)^2 and sum x^2, all-reduced -> out[0], out[1]; out[2] = 1 on a peer timeout
__global__ void __launch_bounds__(kBlock) peer_resid_kernel(int64_t n, const double *__restrict__ b,
                                                             const double *__restrict__ t,
                                                             const double *__restrict__ x, CgState *st,
                                                             double *partials, double *out, PeerCtx pc) {
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double a0 = 0.0, a1 = 0.0;
  for (int64_t i = tid; i < n; i += nthr) {
    const double a = b[i] - t[i];
    a0 = fma(a, a, a0);
    a1 = fma(x[i], x[i], a1);
  }
  a0 = block_sum(a0, red);
  a1 = block_sum(a1, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = a0;
    partials[kMaxPartials + blockIdx.x] = a1;
  }
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const double s0 = reduce_partials(partials, gridDim.x, red);
    const double s1 = reduce_partials(partials + kMaxPartials, gridDim.x, red);
    bool ok = true;
    const double g0 = peer_allreduce(pc, s0, &ok);
    const double g1 = peer_allreduce(pc, s1, &ok);
    if (threadIdx.x == 0) {
      out[0] = g0;
      out[1] = g1;
      out[2] = ok ? 0.0 : 1.0;
    }
  }
}

// All-gather mode: this rank's slice of `vec` goes into every rank's full-length copy (coalesced 16-byte NVLink stores),
// the last CTA raises this rank's flag in every window.  The slice starts at this rank's first row, which may be odd:
// the 16-byte stores then start one entry in (the buffers themselves are 2 MiB aligned, so all copies agree).
__global__ void __launch_bounds__(kBlock) peer_allgather_kernel(int64_t n, const double *__restrict__ vec, PeerGather pg,
                                                                 PeerCtx pc, CgState *st, int gated) {
  if (gated && !cg_active(st)) return;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t head = n > 0 ? (int64_t)((reinterpret_cast<uintptr_t>(pg.dst[pc.rank]) >> 3) & 1) : 0;
  const int64_t n2 = (n - head) >> 1;
  if (head == 0) {
    const double2 *v2 = reinterpret_cast<const double2 *>(vec);
    for (int64_t i = tid; i < n2; i += nthr) {
      const double2 v = v2[i];
      for (int q = 0; q < pc.nranks; q++) reinterpret_cast<double2 *>(pg.dst[q])[i] = v;
    }
  } else {
    for (int64_t i = tid; i < n2; i += nthr) {
      const double2 v = make_double2(vec[2 * i + 1], vec[2 * i + 2]);
      for (int q = 0; q < pc.nranks; q++) reinterpret_cast<double2 *>(pg.dst[q] + 1)[i] = v;
    }
  }
  if (tid == 0) { // the entry before the first aligned pair and the one after the last
    for (int q = 0; q < pc.nranks; q++) {
      if (head) pg.dst[q][0] = vec[0];
      if (head + 2 * n2 < n) pg.dst[q][n - 1] = vec[n - 1];
    }
  }
  __threadfence_system();
  if (publish_and_elect(&st->ticket, gridDim.x)) {
    const unsigned long long s = pc.seq->halo + 1;
    const int q = threadIdx.x;
    if (q < pc.nranks && q != pc.rank) {
      __threadfence_system();
      st_release_sys_u64(&pc.win[q]->hflag[pc.rank], s);
    }
    __syncthreads();
    if (threadIdx.x == 0) pc.seq->halo = s;
  }
}

// All-gather mode: one warp waits until every peer's slice has arrived; the SpMV launches behind it in the stream.
__global__ void peer_flags_wait_kernel(PeerCtx pc, PeerRecv pr, CgState *st, int gated) {
  if (gated && !cg_active(st)) return;
  bool ok = true;
  if (threadIdx.x < pc.nranks && pr.from[threadIdx.x]) ok = peer_wait(&pc.win[pc.rank]->hflag[threadIdx.x], pc.seq->halo);
  if (!ok) peer_fail(st);
}

template <class T>
int d2h_vec(cgb_ctx *ctx, std::vector<T> &out, const T *d, size_t n) {
  out.resize(n);
  if (n) CGB_CUDA(cudaMemcpyAsync(out.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int owner_of(const std::vector<int64_t> &rs, int64_t col) {
  return (int)(std::upper_bound(rs.begin(), rs.end(), col) - rs.begin()) - 1;
}

// Exchange p's boundary entries: pack on the main stream, send/recv on the comm stream.
// On return the comm stream has ev_halo recorded; the caller makes the main stream wait for it.
int halo_exchange(cgb_dist *d, const double *vec, int gated) {
  cudaStream_t s = d->ctx->stream, c = d->comm_stream;
  if (d->allgather) { // the whole slice to every rank, then wait for everybody else's
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((d->n / 2 + kBlock - 1) / kBlock, (int64_t)d->ctx->sm_count * 4));
    peer_allgather_kernel<<<grid, kBlock, 0, s>>>(d->n, vec, d->gather, d->pctx(), d->Aloc->ws.st, gated);
    peer_flags_wait_kernel<<<1, 32, 0, s>>>(d->pctx(), d->recv, d->Aloc->ws.st, gated);
    CGB_CUDA(cudaGetLastError());
    return CGB_OK;
  }
  if (d->peer) {
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((d->S + kBlock - 1) / kBlock, (int64_t)d->ctx->sm_count * 2));
    peer_pack_kernel<<<grid, kBlock, 0, s>>>(d->S, d->send_idx, vec, d->send, d->pctx(), d->Aloc->ws.st, gated);
    CGB_CUDA(cudaGetLastError());
    return CGB_OK;
  }
  if (d->S > 0) {
    const int grid = (int)std::min<int64_t>((d->S + kBlock - 1) / kBlock, (int64_t)d->ctx->sm_count * 2);
    pack_kernel<<<grid, kBlock, 0, s>>>(d->S, d->send_idx, vec, d->send_buf, d->Aloc->ws.st, gated);
    CGB_CUDA(cudaGetLastError());
  }
  CGB_CUDA(cudaEventRecord(d->ev_pack, s));
  CGB_CUDA(cudaStreamWaitEvent(c, d->ev_pack, 0));
  CGB_NCCL(g_nccl.GroupStart());
  for (int q = 0; q < d->nranks; q++) {
    if (d->scount[q] > 0) CGB_NCCL(g_nccl.Send(d->send_buf + d->soff[q], (size_t)d->scount[q], ncclDouble, q, d->comm, c));
    if (d->rcount[q] > 0) CGB_NCCL(g_nccl.Recv(d->halo + d->roff[q], (size_t)d->rcount[q], ncclDouble, q, d->comm, c));
  }
  CGB_NCCL(g_nccl.GroupEnd());
  CGB_CUDA(cudaEventRecord(d->ev_halo, c));
  return CGB_OK;
}

// peer transport: wait for the halos, add the A_halo part; mode 1 also all-reduces p.Ap and forms alpha
int peer_halo_apply(cgb_dist *d, double *y, const double *p, const double *base, int mode, int gated) {
  cudaStream_t s = d->ctx->stream;
  Workspace &w = d->Aloc->ws;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((d->nh + kBlock - 1) / kBlock, (int64_t)d->ctx->sm_count * 4));
  cgb_csr *H = d->Ahalo;
  if (!H) { // all-gather mode: no halo part; mode 1 still all-reduces p.Ap and forms alpha
    if (mode == 1)
      peer_halo_kernel<1><<<1, kBlock, 0, s>>>(0, nullptr, nullptr, nullptr, nullptr, nullptr, y, p, base, d->pctx(), d->recv, w.st,
                                               gated, w.partials);
    CGB_CUDA(cudaGetLastError());
    return CGB_OK;
  }
  int64_t engine_min = kHaloEngineMinNnz;
  if (const char *e = getenv("CGB_HALO_ENGINE_MIN")) engine_min = atoll(e);
  if (H->nnz >= engine_min && H->n > 0) { // a real share of the matrix: regular row engines over A_halo
    const SpmvLaunch lc = spmv_launch_cfg(H);
    const CsrView V = csr_view(H);
    const PeerCtx pc = d->pctx();
    const PeerRecv pr = d->recv;
    const int *hr = d->halo_rows;
    const double *hb = d->halo;
    CgState *stp = w.st;
    double *parts = w.partials;
    dispatch_lanes(lc.lanes, lc.eng, [&](auto Lc, auto Tc) {
      constexpr int LL = decltype(Lc)::value, EE = decltype(Tc)::value;
      if (mode == 1) launch_kernel(peer_halo_engine_kernel<LL, EE, 1>, lc, s, V, hr, hb, y, p, base, pc, pr, stp, gated, parts);
      else launch_kernel(peer_halo_engine_kernel<LL, EE, 0>, lc, s, V, hr, hb, y, p, base, pc, pr, stp, gated, parts);
    });
    CGB_CUDA(cudaGetLastError());
    return CGB_OK;
  }
  if (mode == 1)
    peer_halo_kernel<1><<<grid, kBlock, 0, s>>>(d->nh, H->rowptr, H->cols, H->vals, d->halo_rows, d->halo, y, p, base,
                                                d->pctx(), d->recv, w.st, gated, w.partials);
  else
    peer_halo_kernel<0><<<grid, kBlock, 0, s>>>(d->nh, H->rowptr, H->cols, H->vals, d->halo_rows, d->halo, y, p, base,
                                                d->pctx(), d->recv, w.st, gated, w.partials);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int halo_apply(cgb_dist *d, double *y, const double *p, const double *base, double *out, int gated) {
  if (d->peer) return peer_halo_apply(d, y, p, base, 0, gated);
  cudaStream_t s = d->ctx->stream;
  CGB_CUDA(cudaStreamWaitEvent(s, d->ev_halo, 0));
  Workspace &w = d->Aloc->ws;
  if (d->nh > 64 * kBlock) {
    const int grid = (int)std::min<int64_t>((d->nh + kBlock - 1) / kBlock, (int64_t)d->ctx->sm_count * 4);
    halo_spmv_multi_kernel<<<grid, kBlock, 0, s>>>(d->nh, d->Ahalo->rowptr, d->Ahalo->cols, d->Ahalo->vals,
                                                   d->halo_rows, d->halo, y, p, base, out, w.st, gated, w.partials);
  } else {
    halo_spmv_kernel<<<1, kBlock, 0, s>>>(d->nh, d->Ahalo ? d->Ahalo->rowptr : nullptr,
                                          d->Ahalo ? d->Ahalo->cols : nullptr, d->Ahalo ? d->Ahalo->vals : nullptr,
                                          d->halo_rows, d->halo, y, p, base, out, w.st, gated, 1.0);
  }
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

} // namespace

extern "C" {

int cgb_nccl_unique_id(void *h_id128) {
  CGB_REQUIRE(h_id128, "cgb_nccl_unique_id: NULL argument");
  CGB_TRY(nccl_load());
  ncclUniqueId id;
  CGB_NCCL(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == CGB_NCCL_ID_BYTES, "ncclUniqueId size");
  memcpy(h_id128, &id, sizeof(id));
  return CGB_OK;
}

int cgb_plan_halo(int rank, int nranks, const int64_t *row_starts, int64_t nnz_local, const int32_t *h_cols,
                  int32_t *h_recv_cols, int64_t capacity, int64_t *recv_count, int64_t *recv_per_rank) {
  CGB_REQUIRE(row_starts && recv_count && nranks >= 1 && rank >= 0 && rank < nranks && (nnz_local == 0 || h_cols),
              "cgb_plan_halo: bad argument");
  const int64_t lo = row_starts[rank], hi = row_starts[rank + 1];
  std::vector<int32_t> rem;
  for (int64_t j = 0; j < nnz_local; j++)
    if (h_cols[j] < lo || h_cols[j] >= hi) rem.push_back(h_cols[j]);
  std::sort(rem.begin(), rem.end());
  rem.erase(std::unique(rem.begin(), rem.end()), rem.end());
  *recv_count = (int64_t)rem.size();
  if (recv_per_rank) {
    std::vector<int64_t> rs(row_starts, row_starts + nranks + 1);
    for (int q = 0; q < nranks; q++) recv_per_rank[q] = 0;
    for (int32_t c : rem) {
      const int q = owner_of(rs, c);
      if (q < 0 || q >= nranks) return fail(CGB_ERR_INVALID, "cgb_plan_halo: column %d outside the global range", c);
      recv_per_rank[q]++;
    }
  }
  if (h_recv_cols) {
    CGB_REQUIRE(capacity >= (int64_t)rem.size(), "cgb_plan_halo: capacity too small");
    std::copy(rem.begin(), rem.end(), h_recv_cols);
  }
  return CGB_OK;
}

// every rank has unmapped the others' halo buffers before anyone frees its own (collective)
static void peer_release_halo(cgb_dist *d) {
  for (size_t q = 0; q < d->halo_of.size(); q++)
    if ((int)q != d->rank && d->halo_of[q]) cudaIpcCloseMemHandle(d->halo_of[q]);
  d->halo_of.clear();
  for (size_t q = 0; q < d->pfull_of.size(); q++)
    if ((int)q != d->rank && d->pfull_of[q]) cudaIpcCloseMemHandle(d->pfull_of[q]);
  d->pfull_of.clear();
  if (d->peer && d->comm && d->scal) { // barrier
    g_nccl.AllReduce(d->scal + 6, d->scal + 7, 1, ncclDouble, ncclSum, d->comm, d->ctx->stream);
    cudaStreamSynchronize(d->ctx->stream);
  }
}

static void dist_release_matrix(cgb_dist *d) {
  cudaStreamSynchronize(d->ctx->stream);
  if (d->comm_stream) cudaStreamSynchronize(d->comm_stream);
  if (d->Aloc) peer_release_halo(d);
  cgb_csr_destroy(d->Aloc);
  cgb_csr_destroy(d->Ahalo);
  cudaFree(d->halo_rows);
  cudaFree(d->send_idx);
  cudaFree(d->send_buf);
  cudaFree(d->halo);
  cudaFree(d->pfull);
  cudaFree(d->hplan);
  cudaFree(d->hoff);
  d->hplan = nullptr;
  d->hoff = nullptr;
  d->hplan_grid = d->hplan_tile = 0;
  d->loop_key = 0;
  d->pfull = nullptr;
  d->allgather = false;
  d->Aloc = d->Ahalo = nullptr;
  d->halo_rows = nullptr;
  d->send_idx = nullptr;
  d->send_buf = nullptr;
  d->halo = nullptr;
  d->nh = d->H = d->S = 0;
}

int cgb_dist_destroy(cgb_dist *d) {
  if (!d) return CGB_OK;
  ScopedDevice sd(d->ctx->device);
  dist_release_matrix(d);
  for (size_t q = 0; q < d->win_of.size(); q++)
    if ((int)q != d->rank && d->win_of[q]) cudaIpcCloseMemHandle(d->win_of[q]);
  if (d->peer && d->comm && d->scal) { // nobody frees its window while a peer still maps it
    g_nccl.AllReduce(d->scal + 6, d->scal + 7, 1, ncclDouble, ncclSum, d->comm, d->ctx->stream);
    cudaStreamSynchronize(d->ctx->stream);
  }
  cudaFree(d->win);
  cudaFree(d->bar);
  cudaFree(d->d_win);
  cudaFree(d->seq);
  if (d->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(d->comm);
  cudaFree(d->scal);
  if (d->ev_pack) cudaEventDestroy(d->ev_pack);
  if (d->ev_halo) cudaEventDestroy(d->ev_halo);
  if (d->comm_stream) cudaStreamDestroy(d->comm_stream);
  delete d;
  return CGB_OK;
}

int64_t cgb_dist_halo_count(const cgb_dist *d) { return d ? d->H : 0; }
int cgb_dist_exchange_mode(const cgb_dist *d) { return d && d->allgather ? 1 : 0; }

} // extern "C"

#include "dist_persistent.cuh"
#include "dist_impl.cuh"
