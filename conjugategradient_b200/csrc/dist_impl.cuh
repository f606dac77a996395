// dist_impl.cuh — second half of dist.cu: set-up (matrix split + halo plan) and the solve loop.
#pragma once

namespace {

__global__ void flag_kernel(int64_t n, const int *cnt, int *flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = cnt[i] > 0;
}

struct DevBuf { // frees on scope exit
  void *p = nullptr;
  ~DevBuf() { cudaFree(p); }
  template <class T> T *as() { return (T *)p; }
};

// CGB_DIST_TIMING=1: rank 0 prints the wall time of the set-up phases (stream-synchronised) to stderr
struct PhaseTimer {
  bool on;
  cudaStream_t s;
  std::chrono::steady_clock::time_point t;
  PhaseTimer(const cgb_dist *d) : on(getenv("CGB_DIST_TIMING") != nullptr && d->rank == 0), s(d->ctx->stream), t(std::chrono::steady_clock::now()) {}
  void mark(const char *what) {
    if (!on) return;
    cudaStreamSynchronize(s);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[cgb dist timing] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

int dev_alloc(DevBuf &b, size_t bytes) {
  CGB_CUDA(cudaMalloc(&b.p, bytes ? bytes : 16));
  return CGB_OK;
}

int read_int(cgb_ctx *ctx, const int *d, int *out) {
  CGB_CUDA(cudaMemcpyAsync(out, d, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}


// ---- peer transport set-up (peer.cuh) -------------------------------------------------------------
// All-gather of one cudaIpcMemHandle_t per rank (through NCCL; set-up only).
int gather_ipc_handles(cgb_dist *d, const cudaIpcMemHandle_t &mine, std::vector<cudaIpcMemHandle_t> &all) {
  const int P = d->nranks;
  cudaStream_t s = d->ctx->stream;
  DevBuf b_in, b_out;
  CGB_TRY(dev_alloc(b_in, sizeof(cudaIpcMemHandle_t)));
  CGB_TRY(dev_alloc(b_out, sizeof(cudaIpcMemHandle_t) * (size_t)P));
  CGB_CUDA(cudaMemcpyAsync(b_in.p, &mine, sizeof(mine), cudaMemcpyHostToDevice, s));
  CGB_NCCL(g_nccl.AllGather(b_in.p, b_out.p, sizeof(cudaIpcMemHandle_t), ncclChar, d->comm, s));
  all.resize(P);
  CGB_CUDA(cudaMemcpyAsync(all.data(), b_out.p, sizeof(cudaIpcMemHandle_t) * (size_t)P, cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  return CGB_OK;
}

// Every rank learns whether ALL ranks succeeded (min over ranks of `ok`).
int all_agree(cgb_dist *d, int ok, int *all_ok) {
  cudaStream_t s = d->ctx->stream;
  const double v = ok ? 1.0 : 0.0;
  double out = 0.0;
  CGB_CUDA(cudaMemcpyAsync(d->scal + 6, &v, sizeof(double), cudaMemcpyHostToDevice, s));
  CGB_NCCL(g_nccl.AllReduce(d->scal + 6, d->scal + 7, 1, ncclDouble, ncclMin, d->comm, s));
  CGB_CUDA(cudaMemcpyAsync(&out, d->scal + 7, sizeof(double), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  *all_ok = out > 0.5;
  return CGB_OK;
}

// Sums over the ranks of two host numbers (set-up decisions taken on global figures).
int all_sum2(cgb_dist *d, double a, double b, double *sa, double *sb) {
  cudaStream_t s = d->ctx->stream;
  const double v[2] = {a, b};
  double out[2] = {0.0, 0.0};
  CGB_CUDA(cudaMemcpyAsync(d->scal + 12, v, sizeof(v), cudaMemcpyHostToDevice, s));
  CGB_NCCL(g_nccl.AllReduce(d->scal + 12, d->scal + 14, 2, ncclDouble, ncclSum, d->comm, s));
  CGB_CUDA(cudaMemcpyAsync(out, d->scal + 14, sizeof(out), cudaMemcpyDeviceToHost, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  *sa = out[0];
  *sb = out[1];
  return CGB_OK;
}

// CGB_PEER_TIMEOUT_MS (default 30000) in SM clock cycles
long long peer_timeout_cycles(const cgb_ctx *ctx) {
  double ms = 30000.0;
  if (const char *e = getenv("CGB_PEER_TIMEOUT_MS")) ms = std::max(1.0, atof(e));
  static int khz_of[64] = {}; // asked once per device: the attribute goes through the driver, this runs once per solve
  int &khz = khz_of[ctx->device & 63];
  if (khz <= 0 && (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device) != cudaSuccess || khz <= 0)) khz = 2000000;
  return (long long)(ms * (double)khz);
}

constexpr size_t kIpcGranule = (size_t)2 << 20; // exported allocations are whole 2 MiB blocks

// Allocate this rank's window, map every peer's window.  Leaves d->peer = true only if it worked everywhere.
int peer_window_setup(cgb_dist *d) {
  d->peer = false;
  const char *env = getenv("CGB_DIST_TRANSPORT");
  int want = !(env && !strcmp(env, "nccl")) && d->nranks <= kMaxRanks;
  int ok = want;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok) {
    ok = cudaMalloc(&d->win, kIpcGranule) == cudaSuccess && cudaMemset(d->win, 0, kIpcGranule) == cudaSuccess &&
         cudaMalloc(&d->seq, sizeof(PeerSeq)) == cudaSuccess && cudaMemset(d->seq, 0, sizeof(PeerSeq)) == cudaSuccess &&
         cudaMalloc(&d->d_win, sizeof(PeerWin *) * kMaxRanks) == cudaSuccess &&
         cudaIpcGetMemHandle(&mine, d->win) == cudaSuccess;
    if (!ok) cudaGetLastError();
  }
  std::vector<cudaIpcMemHandle_t> all;
  CGB_TRY(gather_ipc_handles(d, mine, all)); // collective: every rank takes part, successful or not
  int everyone = 0;
  CGB_TRY(all_agree(d, ok, &everyone));
  d->win_of.assign(d->nranks, nullptr);
  if (everyone) {
    d->win_of[d->rank] = d->win;
    for (int q = 0; q < d->nranks && ok; q++) {
      if (q == d->rank) continue;
      void *ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
      } else {
        d->win_of[q] = (PeerWin *)ptr;
      }
    }
    CGB_TRY(all_agree(d, ok, &everyone));
  }
  if (everyone) {
    CGB_CUDA(cudaMemcpy(d->d_win, d->win_of.data(), sizeof(PeerWin *) * (size_t)d->nranks, cudaMemcpyHostToDevice));
    const long long limit = peer_timeout_cycles(d->ctx);
    CGB_CUDA(cudaMemcpyToSymbol(g_peer_timeout_cycles, &limit, sizeof(limit)));
    d->peer = true;
  } else {
    for (int q = 0; q < d->nranks; q++)
      if (q != d->rank && d->win_of[q]) cudaIpcCloseMemHandle(d->win_of[q]);
    d->win_of.clear();
    if (want && d->rank == 0 && getenv("CGB_VERBOSE"))
      fprintf(stderr, "cgb200: peer-memory transport unavailable, using NCCL send/recv + all-reduce\n");
  }
  return CGB_OK;
}

// After the halo buffer of the current matrix exists: map the peers' halo buffers this rank sends
// into and build the destination table of the pack kernel.  M = all-gathered receive counts
// (M[q*P + r] = entries rank q receives from rank r).
int peer_halo_setup(cgb_dist *d, const std::vector<int64_t> &M) {
  const int P = d->nranks;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  int ok = cudaIpcGetMemHandle(&mine, d->halo) == cudaSuccess;
  if (!ok) cudaGetLastError();
  std::vector<cudaIpcMemHandle_t> all;
  CGB_TRY(gather_ipc_handles(d, mine, all));
  d->halo_of.assign(P, nullptr);
  memset(&d->send, 0, sizeof(d->send));
  memset(&d->recv, 0, sizeof(d->recv));
  for (int q = 0; q < P; q++) {
    d->send.soff[q] = d->soff[q];
    d->recv.from[q] = d->rcount[q] > 0;
    if (q == d->rank || d->scount[q] == 0 || !ok) continue;
    void *ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      ok = 0;
      continue;
    }
    d->halo_of[q] = (double *)ptr;
    int64_t off = 0; // where this rank's entries start in q's halo buffer: q receives from ranks in rank order
    for (int r = 0; r < d->rank; r++) off += M[(size_t)q * P + r];
    d->send.dst[q] = d->halo_of[q] + off;
  }
  d->send.soff[P] = d->soff[P];
  for (int q = P + 1; q <= kMaxRanks; q++) d->send.soff[q] = d->soff[P];
  int everyone = 0;
  CGB_TRY(all_agree(d, ok, &everyone));
  if (!everyone) return fail(CGB_ERR_NCCL, "peer transport: mapping a peer's halo buffer failed (set CGB_DIST_TRANSPORT=nccl)");
  return CGB_OK;
}

// All-gather mode set-up: A_loc = a copy of the caller's block as it is (global columns), no halo part; a full-length
// buffer for the gathered vector, exported to the peers, and the peers' buffers mapped here.
int dist_setup_allgather(cgb_dist *d, cgb_csr *A, PhaseTimer &pt) {
  cgb_ctx *ctx = d->ctx;
  cudaStream_t s = ctx->stream;
  const int P = d->nranks;
  const int64_t n = A->n, n_global = d->row_starts[P], lo = d->row_starts[d->rank];
  DevBuf g_r, g_c, g_v;
  CGB_TRY(dev_alloc(g_r, sizeof(int32_t) * (size_t)(n + 1) + 64));
  CGB_TRY(dev_alloc(g_c, sizeof(int32_t) * (size_t)A->nnz + 64));
  CGB_TRY(dev_alloc(g_v, sizeof(double) * (size_t)A->nnz + 64));
  CGB_CUDA(cudaMemcpyAsync(g_r.p, A->rowptr, sizeof(int32_t) * (size_t)(n + 1), cudaMemcpyDeviceToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(g_c.p, A->cols, sizeof(int32_t) * (size_t)A->nnz, cudaMemcpyDeviceToDevice, s));
  CGB_CUDA(cudaMemcpyAsync(g_v.p, A->vals, sizeof(double) * (size_t)A->nnz, cudaMemcpyDeviceToDevice, s));
  CGB_TRY(cgb_csr_from_device(ctx, n, A->nnz, g_v.as<double>(), g_c.as<int32_t>(), g_r.as<int32_t>(), 1, &d->Aloc));
  g_r.p = g_c.p = g_v.p = nullptr;
  d->Aloc->n_cols = n_global; // the gathered vector is the full-length copy: column panels are sized from it
  if (A->format != CGB_FMT_AUTO) {
    const int rc = cgb_csr_set_format(d->Aloc, A->format);
    if (rc != CGB_OK && rc != CGB_ERR_UNSUPPORTED) return rc;
  }
  d->Ahalo = nullptr;
  d->nh = 0;
  d->S = 0;
  d->rcount.assign(P, 0);
  d->scount.assign(P, 0);
  d->roff.assign(P + 1, 0);
  d->soff.assign(P + 1, 0);
  memset(&d->send, 0, sizeof(d->send));
  memset(&d->recv, 0, sizeof(d->recv));
  for (int q = 0; q < P; q++) d->recv.from[q] = q != d->rank && d->row_starts[q + 1] > d->row_starts[q];
  d->pfull_bytes = ((sizeof(double) * (size_t)(n_global + 2) + kIpcGranule - 1) / kIpcGranule) * kIpcGranule;
  CGB_CUDA(cudaMalloc(&d->pfull, d->pfull_bytes));
  CGB_CUDA(cudaMemsetAsync(d->pfull, 0, d->pfull_bytes, s));
  CGB_CUDA(cudaStreamSynchronize(s));
  pt.mark("setup: all-gather block copy");
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  int ok = cudaIpcGetMemHandle(&mine, d->pfull) == cudaSuccess;
  if (!ok) cudaGetLastError();
  std::vector<cudaIpcMemHandle_t> all;
  CGB_TRY(gather_ipc_handles(d, mine, all));
  d->pfull_of.assign(P, nullptr);
  memset(&d->gather, 0, sizeof(d->gather));
  for (int q = 0; q < P; q++) {
    if (q == d->rank) {
      d->pfull_of[q] = d->pfull;
    } else if (ok) {
      void *ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
      } else {
        d->pfull_of[q] = (double *)ptr;
      }
    }
    if (d->pfull_of[q]) d->gather.dst[q] = d->pfull_of[q] + lo;
  }
  int everyone = 0;
  CGB_TRY(all_agree(d, ok, &everyone));
  if (!everyone) return fail(CGB_ERR_NCCL, "all-gather mode: mapping a peer's vector buffer failed (set CGB_DIST_ALLGATHER=0)");
  d->allgather = true;
  pt.mark("setup: all-gather mapping");
  return CGB_OK;
}

int dist_setup(cgb_dist *d, cgb_csr *A) {
  cgb_ctx *ctx = d->ctx;
  cudaStream_t s = ctx->stream;
  const int64_t n = A->n;
  const int lo = (int)d->row_starts[d->rank], hi = (int)d->row_starts[d->rank + 1];
  const int tb = 256;
  const unsigned gb = (unsigned)std::max<int64_t>(1, (n + tb - 1) / tb);
  PhaseTimer pt(d);
  DevBuf b_cl, b_cr, b_flag, b_slot, b_tmp, b_rem, b_rem_sorted, b_num, b_cnt_compact;
  CGB_TRY(dev_alloc(b_cl, sizeof(int) * (size_t)(n + 1)));
  CGB_TRY(dev_alloc(b_cr, sizeof(int) * (size_t)(n + 1)));
  CGB_TRY(dev_alloc(b_flag, sizeof(int) * (size_t)(n + 1)));
  CGB_TRY(dev_alloc(b_slot, sizeof(int) * (size_t)(n + 1)));
  CGB_TRY(dev_alloc(b_num, sizeof(int) * 4));
  int *cnt_loc = b_cl.as<int>(), *cnt_rem = b_cr.as<int>(), *flag = b_flag.as<int>(), *slot = b_slot.as<int>();
  int *d_num = b_num.as<int>();
  if (n > 0) {
    count_split_kernel<<<gb, tb, 0, s>>>(n, A->rowptr, A->cols, lo, hi, cnt_loc, cnt_rem);
    flag_kernel<<<gb, tb, 0, s>>>(n, cnt_rem, flag);
    CGB_CUDA(cudaGetLastError());
  }
  // temp storage big enough for every cub call below
  size_t tmp_bytes = 0, t;
  const int ni = (int)n, nz = (int)A->nnz;
  int *nullp = nullptr;
  cub::DeviceScan::InclusiveSum(nullptr, t, nullp, nullp, ni, s); tmp_bytes = std::max(tmp_bytes, t);
  cub::DeviceScan::ExclusiveSum(nullptr, t, nullp, nullp, ni, s); tmp_bytes = std::max(tmp_bytes, t);
  cub::DeviceSelect::If(nullptr, t, nullp, nullp, nullp, nz, IsRemote{lo, hi}, s); tmp_bytes = std::max(tmp_bytes, t);
  cub::DeviceSelect::If(nullptr, t, nullp, nullp, nullp, ni, NonZero{}, s); tmp_bytes = std::max(tmp_bytes, t);
  CGB_TRY(dev_alloc(b_tmp, tmp_bytes));

  pt.mark("setup: count local/remote");
  // A_loc row offsets
  DevBuf g_rl, g_rh, g_cl, g_vl, g_ch, g_vh; // own the new arrays until the two handles below take them over
  int32_t *rowptr_loc = nullptr;
  CGB_CUDA(cudaMalloc(&rowptr_loc, sizeof(int32_t) * (size_t)(n + 1) + 64));
  g_rl.p = rowptr_loc;
  CGB_CUDA(cudaMemsetAsync(rowptr_loc, 0, sizeof(int32_t), s));
  int nnz_loc = 0, nh = 0, nnz_halo = 0;
  if (n > 0) {
    t = tmp_bytes;
    cub::DeviceScan::InclusiveSum(b_tmp.p, t, cnt_loc, rowptr_loc + 1, ni, s);
    CGB_TRY(read_int(ctx, rowptr_loc + n, &nnz_loc));
    t = tmp_bytes;
    cub::DeviceScan::ExclusiveSum(b_tmp.p, t, flag, slot, ni, s);
    int last_slot = 0, last_flag = 0;
    CGB_TRY(read_int(ctx, slot + n - 1, &last_slot));
    CGB_TRY(read_int(ctx, flag + n - 1, &last_flag));
    nh = last_slot + last_flag;
  }
  nnz_halo = (int)A->nnz - nnz_loc;
  d->nh = nh;

  // A_halo row offsets: compact the non-zero remote counts (row order), then scan
  int32_t *rowptr_halo = nullptr;
  CGB_CUDA(cudaMalloc(&rowptr_halo, sizeof(int32_t) * (size_t)(nh + 1) + 64));
  g_rh.p = rowptr_halo;
  CGB_CUDA(cudaMemsetAsync(rowptr_halo, 0, sizeof(int32_t), s));
  if (nh > 0) {
    t = tmp_bytes;
    cub::DeviceSelect::If(b_tmp.p, t, cnt_rem, rowptr_halo + 1, d_num, ni, NonZero{}, s);
    t = tmp_bytes;
    cub::DeviceScan::InclusiveSum(b_tmp.p, t, rowptr_halo + 1, rowptr_halo + 1, nh, s);
  }

  pt.mark("setup: row offsets");
  // receive list: sorted distinct remote columns
  CGB_TRY(dev_alloc(b_rem, sizeof(int) * (size_t)(nnz_halo + 1)));
  CGB_TRY(dev_alloc(b_rem_sorted, sizeof(int) * (size_t)(nnz_halo + 1)));
  int32_t *recv_cols = nullptr;
  CGB_CUDA(cudaMalloc(&recv_cols, sizeof(int32_t) * (size_t)(nnz_halo + 1)));
  DevBuf recv_guard;
  recv_guard.p = recv_cols;
  int H = 0;
  if (nnz_halo > 0) {
    t = tmp_bytes;
    cub::DeviceSelect::If(b_tmp.p, t, A->cols, b_rem.as<int>(), d_num, nz, IsRemote{lo, hi}, s);
    size_t t2 = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, t2, b_rem.as<int>(), b_rem_sorted.as<int>(), nnz_halo, 0, 32, s);
    DevBuf b_tmp2;
    CGB_TRY(dev_alloc(b_tmp2, t2));
    cub::DeviceRadixSort::SortKeys(b_tmp2.p, t2, b_rem.as<int>(), b_rem_sorted.as<int>(), nnz_halo, 0, 32, s);
    size_t t3 = 0;
    cub::DeviceSelect::Unique(nullptr, t3, b_rem_sorted.as<int>(), recv_cols, d_num, nnz_halo, s);
    DevBuf b_tmp3;
    CGB_TRY(dev_alloc(b_tmp3, t3));
    cub::DeviceSelect::Unique(b_tmp3.p, t3, b_rem_sorted.as<int>(), recv_cols, d_num, nnz_halo, s);
    CGB_TRY(read_int(ctx, d_num, &H));
  }
  d->H = H;

  pt.mark("setup: receive list");
  // Dense coupling (a random matrix: nearly every remote entry of the vector is needed by every rank): keep the block
  // whole with its global columns and all-gather the vector instead of splitting off a halo part that would be most
  // of the matrix ("if the union of needed remote columns is >~ half of n, all-gather p").
  {
    const int64_t n_global = d->row_starts[d->nranks];
    // the figure is global (remote entries needed / remote entries there are, summed over the ranks): a rank with few
    // rows of its own needs less than half of the rest without that changing what the others have to receive
    double need = 0.0, there = 0.0;
    CGB_TRY(all_sum2(d, (double)H, (double)(n_global - n), &need, &there));
    int want = d->peer && there > 0.0 && need >= 0.5 * there;
    if (const char *e = getenv("CGB_DIST_ALLGATHER")) want = d->peer && atoi(e) != 0;
    int all = 0;
    CGB_TRY(all_agree(d, want, &all));
    if (all) return dist_setup_allgather(d, A, pt);
  }
  // fill both parts
  int32_t *cols_loc = nullptr, *cols_halo = nullptr;
  double *vals_loc = nullptr, *vals_halo = nullptr;
  CGB_CUDA(cudaMalloc(&cols_loc, sizeof(int32_t) * (size_t)nnz_loc + 64));
  g_cl.p = cols_loc;
  CGB_CUDA(cudaMalloc(&vals_loc, sizeof(double) * (size_t)nnz_loc + 64));
  g_vl.p = vals_loc;
  CGB_CUDA(cudaMalloc(&cols_halo, sizeof(int32_t) * (size_t)nnz_halo + 64));
  g_ch.p = cols_halo;
  CGB_CUDA(cudaMalloc(&vals_halo, sizeof(double) * (size_t)nnz_halo + 64));
  g_vh.p = vals_halo;
  CGB_CUDA(cudaMalloc(&d->halo_rows, sizeof(int32_t) * (size_t)(nh + 1)));
  if (n > 0) {
    split_fill_kernel<<<gb, tb, 0, s>>>(n, A->rowptr, A->cols, A->vals, lo, hi, rowptr_loc, cols_loc, vals_loc, cnt_rem,
                                        slot, rowptr_halo, cols_halo, vals_halo, d->halo_rows, recv_cols, H);
    CGB_CUDA(cudaGetLastError());
  }
  pt.mark("setup: fill local + halo");
  CGB_TRY(cgb_csr_from_device(ctx, n, nnz_loc, vals_loc, cols_loc, rowptr_loc, 1, &d->Aloc));
  g_rl.p = g_cl.p = g_vl.p = nullptr;
  CGB_TRY(cgb_csr_from_device(ctx, nh, nnz_halo, vals_halo, cols_halo, rowptr_halo, 1, &d->Ahalo));
  g_rh.p = g_ch.p = g_vh.p = nullptr;
  d->Aloc->panels_ok = d->Ahalo->panels_ok = false; // the distributed loop launches one kernel per block itself
  d->Ahalo->n_cols = H;
  // An explicit storage format on the caller's block carries over to its local part, which holds nearly all the
  // entries (the halo part keeps the automatic choice).  A block that does not fit the format stays as it is.
  if (A->format != CGB_FMT_AUTO) {
    const int rc = cgb_csr_set_format(d->Aloc, A->format);
    if (rc != CGB_OK && rc != CGB_ERR_UNSUPPORTED) return rc;
  }

  pt.mark("setup: handles + format");
  // who sends what: receive counts per owner, all-gathered into a matrix
  std::vector<int32_t> h_recv;
  CGB_TRY(d2h_vec(ctx, h_recv, recv_cols, (size_t)H));
  const int P = d->nranks;
  d->rcount.assign(P, 0);
  for (int32_t c : h_recv) {
    const int q = owner_of(d->row_starts, c);
    if (q < 0 || q >= P || q == d->rank) return fail(CGB_ERR_INVALID, "cgb_dist_create: column %d has no remote owner", c);
    d->rcount[q]++;
  }
  d->roff.assign(P + 1, 0);
  for (int q = 0; q < P; q++) d->roff[q + 1] = d->roff[q] + d->rcount[q];
  DevBuf b_counts, b_matrix;
  CGB_TRY(dev_alloc(b_counts, sizeof(int64_t) * (size_t)P));
  CGB_TRY(dev_alloc(b_matrix, sizeof(int64_t) * (size_t)P * P));
  CGB_CUDA(cudaMemcpyAsync(b_counts.p, d->rcount.data(), sizeof(int64_t) * (size_t)P, cudaMemcpyHostToDevice, s));
  CGB_NCCL(g_nccl.AllGather(b_counts.p, b_matrix.p, (size_t)P, ncclInt64, d->comm, s));
  std::vector<int64_t> M;
  CGB_TRY(d2h_vec(ctx, M, b_matrix.as<int64_t>(), (size_t)P * P));
  d->scount.assign(P, 0);
  for (int q = 0; q < P; q++) d->scount[q] = M[(size_t)q * P + d->rank];
  d->soff.assign(P + 1, 0);
  for (int q = 0; q < P; q++) d->soff[q + 1] = d->soff[q] + d->scount[q];
  d->S = d->soff[P];
  CGB_CUDA(cudaMalloc(&d->send_idx, sizeof(int32_t) * (size_t)(d->S + 1)));
  CGB_CUDA(cudaMalloc(&d->send_buf, sizeof(double) * (size_t)(d->S + 1)));
  d->halo_bytes = ((sizeof(double) * (size_t)(H + 1) + kIpcGranule - 1) / kIpcGranule) * kIpcGranule;
  CGB_CUDA(cudaMalloc(&d->halo, d->halo_bytes));
  CGB_CUDA(cudaMemsetAsync(d->halo, 0, d->halo_bytes, s));
  CGB_NCCL(g_nccl.GroupStart());
  for (int q = 0; q < P; q++) {
    if (d->rcount[q] > 0) CGB_NCCL(g_nccl.Send(recv_cols + d->roff[q], (size_t)d->rcount[q], ncclInt32, q, d->comm, s));
    if (d->scount[q] > 0) CGB_NCCL(g_nccl.Recv(d->send_idx + d->soff[q], (size_t)d->scount[q], ncclInt32, q, d->comm, s));
  }
  CGB_NCCL(g_nccl.GroupEnd());
  if (d->S > 0) {
    rebase_kernel<<<(unsigned)((d->S + tb - 1) / tb), tb, 0, s>>>(d->send_idx, d->S, lo);
    CGB_CUDA(cudaGetLastError());
  }
  CGB_CUDA(cudaStreamSynchronize(s));
  pt.mark("setup: send/receive plan");
  if (d->peer) CGB_TRY(peer_halo_setup(d, M));
  pt.mark("setup: peer halo mapping");
  return CGB_OK;
}

struct DistLoop {
  cudaGraphExec_t graph = nullptr;
};

int dist_enqueue_iteration(cgb_dist *d, double *x) {
  cgb_csr *A = d->Aloc;
  Workspace &w = A->ws;
  cudaStream_t s = d->ctx->stream;
  const CsrView V = csr_view(A);
  const SpmvLaunch lc = spmv_launch_cfg(A);
  cg_pupdate_kernel<<<A->grid_pup, kBlock, 0, s>>>(A->n, w.r, w.p0, x, w.st);
  CGB_TRY(halo_exchange(d, w.p0, 1));
  if (d->allgather) { // the block with its global columns over the full-length copy of p (column panels as on one GPU)
    launch_panel_partials(A, lc, s, d->pfull, w.Ap, w.st);
    const CsrView Vf = csr_view_final(A, w.Ap);
    const double *gather = d->pfull;
    dispatch_lanes(lc.lanes, lc.eng, [&](auto Lc, auto Tc) {
      launch_kernel(cg_spmv_kernel<decltype(Lc)::value, decltype(Tc)::value, false>, lc, s, Vf, (const double *)w.r, w.p0,
                    w.p1, w.Ap, x, w.st, w.partials, d->scal + 0, gather);
    });
  } else {
    dispatch_lanes(lc.lanes, lc.eng, [&](auto Lc, auto Tc) {
      launch_kernel(cg_spmv_kernel<decltype(Lc)::value, decltype(Tc)::value, false>, lc, s, V, (const double *)w.r, w.p0,
                    w.p1, w.Ap, x, w.st, w.partials, d->scal + 0, (const double *)nullptr);
    });
  }
  if (d->peer) { // halo correction + all-reduce(p.Ap) + alpha, then r update + all-reduce(r.r) + state advance
    CGB_TRY(peer_halo_apply(d, w.Ap, w.p0, d->scal + 0, 1, 1));
    peer_update_kernel<<<A->grid_vec, kBlock, 0, s>>>(A->n, w.r, w.Ap, w.st, w.partials, w.hist, d->pctx());
    CGB_CUDA(cudaGetLastError());
    return CGB_OK;
  }
  CGB_TRY(halo_apply(d, w.Ap, w.p0, d->scal + 0, d->scal + 0, 1));
  CGB_NCCL(g_nccl.AllReduce(d->scal + 0, d->scal + 1, 1, ncclDouble, ncclSum, d->comm, s));
  cg_update_kernel<<<A->grid_vec, kBlock, 0, s>>>(A->n, w.r, w.Ap, w.st, w.partials, w.hist, d->scal + 1, d->scal + 2);
  CGB_NCCL(g_nccl.AllReduce(d->scal + 2, d->scal + 3, 1, ncclDouble, ncclSum, d->comm, s));
  cg_advance_kernel<<<1, 32, 0, s>>>(w.st, d->scal + 3, w.hist);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

} // namespace

extern "C" {

int cgb_dist_create(cgb_ctx *ctx, int rank, int nranks, const void *h_id128, const int64_t *row_starts,
                    cgb_csr *A, cgb_dist **out) {
  CGB_REQUIRE(ctx && h_id128 && row_starts && A && out, "cgb_dist_create: NULL argument");
  CGB_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "cgb_dist_create: bad rank");
  CGB_REQUIRE(row_starts[rank + 1] - row_starts[rank] == A->n, "cgb_dist_create: row block size mismatch");
  CGB_REQUIRE(row_starts[nranks] < ((int64_t)1 << 31), "cgb_dist_create: global size must fit 32-bit columns");
  *out = nullptr;
  CGB_TRY(nccl_load());
  ScopedDevice sd(ctx->device);
  cgb_dist *d = new cgb_dist();
  d->ctx = ctx;
  d->rank = rank;
  d->nranks = nranks;
  d->n = A->n;
  d->row_starts.assign(row_starts, row_starts + nranks + 1);
  int rc = CGB_OK;
  do {
    ncclUniqueId id;
    memcpy(&id, h_id128, sizeof(id));
    cache_trim(); // NCCL allocates its buffers with the driver: blocks parked in the library's cache must not starve it
    ncclResult_t r = g_nccl.CommInitRank(&d->comm, nranks, id, rank);
    if (r != ncclSuccess) {
      rc = fail(CGB_ERR_NCCL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
      d->comm = nullptr;
      break;
    }
    cudaError_t e = cudaStreamCreateWithFlags(&d->comm_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->ev_pack, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->ev_halo, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc(&d->scal, sizeof(double) * 16);
    if (e == cudaSuccess) e = cudaMemsetAsync(d->scal, 0, sizeof(double) * 16, ctx->stream);
    if (e != cudaSuccess) {
      rc = fail(CGB_ERR_CUDA, "cgb_dist_create: %s", cudaGetErrorString(e));
      break;
    }
    rc = peer_window_setup(d);
    if (rc == CGB_OK) rc = dist_setup(d, A);
  } while (0);
  if (rc != CGB_OK) {
    std::string keep = cgb_last_error();
    d->peer = false; // no collective barriers while tearing down a half-built handle
    cgb_dist_destroy(d);
    set_error("%s", keep.c_str());
    return rc;
  }
  *out = d;
  return CGB_OK;
}

int cgb_dist_set_matrix(cgb_dist *d, const int64_t *row_starts, cgb_csr *A) {
  CGB_REQUIRE(d && row_starts && A, "cgb_dist_set_matrix: NULL argument");
  CGB_REQUIRE(row_starts[d->rank + 1] - row_starts[d->rank] == A->n, "cgb_dist_set_matrix: row block size mismatch");
  CGB_REQUIRE(row_starts[d->nranks] < ((int64_t)1 << 31), "cgb_dist_set_matrix: global size must fit 32-bit columns");
  ScopedDevice sd(d->ctx->device);
  PhaseTimer pt(d);
  dist_release_matrix(d);
  pt.mark("set_matrix: release old");
  d->n = A->n;
  d->row_starts.assign(row_starts, row_starts + d->nranks + 1);
  return dist_setup(d, A);
}

int cgb_dist_cg_solve(cgb_dist *d, const double *d_b, double *d_x, double abs_tol, int64_t max_iter,
                      const cgb_cg_opts *opts_in, cgb_cg_stats *stats) {
  CGB_REQUIRE(d, "cgb_dist_cg_solve: handle is NULL");
  if (!d->Aloc) return fail(CGB_ERR_STATE, "No Matrix given");
  if (!d_b) return fail(CGB_ERR_STATE, "No right hand side to solve for");
  CGB_REQUIRE(d_x, "cgb_dist_cg_solve: x is NULL");
  CGB_REQUIRE(((uintptr_t)d_x & 15) == 0, "cgb_dist_cg_solve: x must be 16-byte aligned (128-bit vector accesses)");
  if (d->dead) return fail(CGB_ERR_STATE, "cgb_dist_cg_solve: this handle saw a peer time-out; destroy it on every rank and create a new one");
  cgb_cg_opts o;
  if (opts_in) o = *opts_in; else cgb_cg_default_opts(&o);
  if (stats) memset(stats, 0, sizeof(*stats));
  cgb_ctx *ctx = d->ctx;
  cgb_csr *A = d->Aloc;
  const int64_t n = A->n, n_global = d->row_starts[d->nranks];
  if (max_iter < 0) max_iter = n_global + 1;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  int batch = o.check_every > 0 ? o.check_every : 32;
  if ((int64_t)batch > max_iter) batch = (int)std::max<int64_t>(1, max_iter);
  const int64_t hist_cap = o.record_history ? max_iter + 1 : 0;
  PhaseTimer pt(d);
  CGB_TRY(csr_prepare(A));
  if (d->Ahalo) CGB_TRY(csr_prepare(d->Ahalo));
  pt.mark("solve: prepare engines");
  CGB_TRY(workspace_ensure(A, n, false, hist_cap));
  Workspace &w = A->ws;
  const bool use_graph = o.loop_mode != CGB_LOOP_STEPWISE;
  // Collective decisions, and the barrier that keeps rank skew out of the in-kernel waits: the host work above
  // (engine preparation, allocations, lazy module loads) takes a different time on every rank, and the first peer
  // kernel below waits for its peers with a time limit.
  // (Collective once per matrix and option set: repeated solves of the same system re-use the answer, which keeps an
  // NCCL round trip and two host synchronisations out of short solves.)
  const int okey = 1 + o.loop_mode + 16 * (o.time_kernels != 0) + 32 * (o.jacobi != 0) + 64 * (o.residual_replace_every > 0) +
                   128 * (o.fuse_p_update != 0);
  int persistent = 0;
  if (d->loop_key == okey) {
    persistent = d->loop_persistent;
  } else {
    // every rank must have passed the same options: sum k = N k and sum k^2 = N k^2 only if all k are equal
    double sk = 0.0, sk2 = 0.0;
    CGB_TRY(all_sum2(d, (double)okey, (double)okey * okey, &sk, &sk2));
    if (sk != (double)d->nranks * okey || sk2 != (double)d->nranks * okey * okey)
      return fail(CGB_ERR_INVALID, "cgb_dist_cg_solve: the ranks passed different options (loop_mode, time_kernels, jacobi, "
                  "residual_replace_every, fuse_p_update must be the same on every rank)");
    CGB_TRY(all_agree(d, dist_persistent_eligible(d, o) ? 1 : 0, &persistent));
    d->loop_key = okey;
    d->loop_persistent = persistent;
  }
  if (o.loop_mode == CGB_LOOP_PERSISTENT && !persistent)
    return fail(CGB_ERR_UNSUPPORTED, "cgb_dist_cg_solve: CGB_LOOP_PERSISTENT needs the peer transport, the TMA row engine on every rank and a small halo part");

  cudaEvent_t t0, t1;
  CGB_CUDA(cudaEventCreate(&t0));
  CGB_CUDA(cudaEventCreate(&t1));
  struct G {
    cudaEvent_t a, b;
    ~G() { cudaEventDestroy(a); cudaEventDestroy(b); }
  } guard{t0, t1};
  CGB_CUDA(cudaEventRecord(t0, s));
  CgState init;
  memset(&init, 0, sizeof(init));
  init.tol = abs_tol;
  init.max_iter = max_iter;
  init.record_history = o.record_history ? 1 : 0;
  init.hist_cap = w.hist_cap;
  init.done = 1;
  w.h_st[0] = init;
  CGB_CUDA(cudaMemcpyAsync(w.st, &w.h_st[0], sizeof(CgState), cudaMemcpyHostToDevice, s));
  CGB_CUDA(cudaMemsetAsync(w.p0, 0, sizeof(double) * (size_t)std::max<int64_t>(n, 1), s));
  // K0: r = b - A x0 (halo of x0 first), rr = allreduce(r.r)
  CGB_TRY(halo_exchange(d, d_x, 0));
  if (n > 0) CGB_TRY(cgb_spmv(ctx, A, d->allgather ? d->pfull : d_x, w.Ap));
  CGB_TRY(halo_apply(d, w.Ap, nullptr, nullptr, nullptr, 0));
  if (d->peer) {
    peer_init_kernel<<<A->grid_vec, kBlock, 0, s>>>(n, d_b, w.Ap, w.r, w.st, w.partials, w.hist, d->pctx());
  } else {
    dist_init_kernel<<<A->grid_vec, kBlock, 0, s>>>(n, d_b, w.Ap, w.r, w.st, w.partials, d->scal + 2);
    CGB_CUDA(cudaGetLastError());
    CGB_NCCL(g_nccl.AllReduce(d->scal + 2, d->scal + 3, 1, ncclDouble, ncclSum, d->comm, s));
    dist_arm_kernel<<<1, 32, 0, s>>>(w.st, d->scal + 3, w.hist);
  }
  CGB_CUDA(cudaGetLastError());

  // kernels of ours per iteration (NCCL's own kernels not counted); all-gather mode: p-update, all-gather, flag wait,
  // one SpMV launch per column panel, all-reduce(p.Ap), update
  const int per_iter = d->allgather ? 5 + (int)std::max<size_t>(1, csr_panels(A)) : d->peer ? 5 : 6;
  int64_t launches = d->peer ? 4 : 5;
  const uint64_t gkey = ((uint64_t)batch << 1) | (d->peer ? 1u : 0u);
  if (persistent) { // the whole loop, the last x update included, in one cooperative kernel (dist_persistent.cuh)
    CGB_TRY(dist_persistent_launch(d, d_x, peer_timeout_cycles(ctx)));
    launches += 1;
    CGB_CUDA(cudaMemcpyAsync(&w.h_st[0], w.st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    const char *tm = getenv("CGB_DIST_TIMING"); // 1: rank 0 prints where a trip's time goes, as its CTA 0 saw it; 2: every rank
    if (tm && (d->rank == 0 || atoi(tm) >= 2) && w.h_st[0].iter > 0) {
      DistBar hb;
      CGB_CUDA(cudaMemcpy(&hb, d->bar, sizeof(hb), cudaMemcpyDeviceToHost));
      static const char *nm[9] = {"p-update", "barrier", "pack", "SpMV local", "halo flag wait", "halo rows", "all-reduce p.Ap", "r-update", "all-reduce r.r"};
      double tot = 0;
      for (int k = 0; k < 9; k++) tot += (double)hb.phase_ns[k];
      char line[640];
      int off = snprintf(line, sizeof(line), "[cgb dist timing] persistent loop, %lld trips, CTA 0 of rank %d, us per trip:", (long long)w.h_st[0].iter, d->rank);
      for (int k = 0; k < 9; k++) off += snprintf(line + off, sizeof(line) - (size_t)off, " %s %.1f |", nm[k], 1e-3 * (double)hb.phase_ns[k] / (double)w.h_st[0].iter);
      snprintf(line + off, sizeof(line) - (size_t)off, " total %.1f\n", 1e-3 * tot / (double)w.h_st[0].iter);
      fputs(line, stderr); // one write per rank: lines of different ranks do not interleave
    }
  }
  if (!persistent && use_graph && (!w.graph || w.graph_x != d_x || w.graph_key != gkey)) {
    if (w.graph) {
      cudaGraphExecDestroy(w.graph);
      w.graph = nullptr;
    }
    cudaGraph_t g = nullptr;
    CGB_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    int rc = CGB_OK;
    for (int i = 0; i < batch && rc == CGB_OK; i++) rc = dist_enqueue_iteration(d, d_x);
    cudaError_t e = cudaStreamEndCapture(s, &g);
    if (rc != CGB_OK) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    if (e != cudaSuccess) return fail(CGB_ERR_CUDA, "cudaStreamEndCapture (dist): %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&w.graph, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) {
      w.graph = nullptr;
      return fail(CGB_ERR_CUDA, "cudaGraphInstantiate (dist): %s", cudaGetErrorString(e));
    }
    w.graph_key = gkey;
    w.graph_x = d_x;
  }
  pt.mark("solve: workspace, K0, graph");
  // every rank takes the same decisions: the polled state is built from all-reduced scalars only
  while (!persistent) {
    if (use_graph) {
      CGB_CUDA(cudaGraphLaunch(w.graph, s));
    } else {
      for (int i = 0; i < batch; i++) CGB_TRY(dist_enqueue_iteration(d, d_x));
    }
    launches += (int64_t)batch * per_iter;
    CGB_CUDA(cudaMemcpyAsync(&w.h_st[0], w.st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
    CGB_CUDA(cudaStreamSynchronize(s));
    if (ctx->progress) ctx->progress((int64_t)w.h_st[0].iter, n_global, w.h_st[0].rr, ctx->progress_user);
    if (w.h_st[0].done || w.h_st[0].iter >= w.h_st[0].max_iter || w.h_st[0].comm_error) break;
  }
  if (w.h_st[0].comm_error) {
    d->dead = true;
    return fail(CGB_ERR_NCCL, "cgb_dist_cg_solve: a peer did not answer within the time limit (rank %d, trip %lld; CGB_PEER_TIMEOUT_MS)", d->rank,
                (long long)w.h_st[0].iter);
  }
  if (n > 0 && !persistent) cg_xflush_kernel<<<A->grid_pup, kBlock, 0, s>>>(n, d_x, w.p0, w.p1, 0, w.st, 0); // last trip's x update
  CGB_CUDA(cudaGetLastError());
  if (!persistent) launches += 1;
  CGB_CUDA(cudaEventRecord(t1, s));
  CGB_CUDA(cudaEventSynchronize(t1));
  pt.mark("solve: iterations");
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
    stats->kernel_launches = launches;
    stats->loop_mode = persistent ? CGB_LOOP_PERSISTENT : use_graph ? CGB_LOOP_GRAPH : CGB_LOOP_STEPWISE;
    { const int eng = csr_engine(A); stats->format = eng == kEngSell ? CGB_FMT_SELL : eng == kEngDict ? CGB_FMT_DICT : CGB_FMT_CSR; }
  }
  return CGB_OK;
}

int cgb_dist_residual(cgb_dist *d, const double *d_x, const double *d_b, double *rr, double *xx) {
  CGB_REQUIRE(d && d_x && d_b && rr && xx, "cgb_dist_residual: NULL argument");
  CGB_REQUIRE(((uintptr_t)d_x & 15) == 0, "cgb_dist_residual: x must be 16-byte aligned (128-bit vector accesses)");
  cgb_ctx *ctx = d->ctx;
  cgb_csr *A = d->Aloc;
  ScopedDevice sd(ctx->device);
  cudaStream_t s = ctx->stream;
  CGB_TRY(csr_prepare(A));
  if (d->Ahalo) CGB_TRY(csr_prepare(d->Ahalo));
  CGB_TRY(workspace_ensure(A, A->n, false, 0));
  Workspace &w = A->ws;
  if (d->dead) return fail(CGB_ERR_STATE, "cgb_dist_residual: this handle saw a peer time-out; destroy it on every rank and create a new one");
  if (d->peer) { // host-level barrier: rank skew stays out of the in-kernel waits
    int all = 0;
    CGB_TRY(all_agree(d, 1, &all));
  }
  CGB_TRY(halo_exchange(d, d_x, 0));
  if (A->n > 0) CGB_TRY(cgb_spmv(ctx, A, d->allgather ? d->pfull : d_x, w.Ap));
  CGB_TRY(halo_apply(d, w.Ap, nullptr, nullptr, nullptr, 0));
  if (d->peer) {
    peer_resid_kernel<<<A->grid_vec, kBlock, 0, s>>>(A->n, d_b, w.Ap, d_x, w.st, w.partials, d->scal + 8, d->pctx());
    CGB_CUDA(cudaGetLastError());
    CGB_CUDA(cudaMemcpyAsync(ctx->pinned, d->scal + 8, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  } else {
    dist_resid_kernel<<<A->grid_vec, kBlock, 0, s>>>(A->n, d_b, w.Ap, d_x, w.st, w.partials, d->scal + 4);
    CGB_CUDA(cudaGetLastError());
    CGB_NCCL(g_nccl.AllReduce(d->scal + 4, d->scal + 4, 2, ncclDouble, ncclSum, d->comm, s));
    CGB_CUDA(cudaMemsetAsync(d->scal + 6, 0, sizeof(double), s));
    CGB_CUDA(cudaMemcpyAsync(ctx->pinned, d->scal + 4, 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  }
  CGB_CUDA(cudaStreamSynchronize(s));
  *rr = ctx->pinned[0];
  *xx = ctx->pinned[1];
  if (ctx->pinned[2] != 0.0) {
    d->dead = true;
    return fail(CGB_ERR_NCCL, "cgb_dist_residual: a peer did not answer within the time limit");
  }
  return CGB_OK;
}

} // extern "C"
