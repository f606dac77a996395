// peer.cuh — NVLink peer-memory transport of the row-partitioned CG (one process per GPU).
//
// The two collectives of a CG iteration are tiny and latency-bound (two 8-byte all-reduces, a halo
// of a few MB at most), so they are done by the compute kernels themselves over peer-mapped memory
// (cudaIpc handles of cudaMalloc'ed windows, NVLink 5 through NVSwitch) instead of separate NCCL
// launches:
//   * halo: the pack kernel stores the boundary entries of p straight into the owner's halo buffer
//     on the peer GPU and raises a sequence flag there; the halo-correction kernel of the receiver
//     waits for the flags it needs, so the transfer overlaps the local SpMV that runs in between;
//   * all-reduce: the last CTA of a reducing kernel writes its partial into a slot on every peer,
//     raises a sequence flag, waits for all peers' flags in its own window and sums the slots in
//     RANK ORDER (bitwise reproducible, identical on every rank).
// Ordering: data stores, __threadfence_system(), then st.release.sys of the flag; the reader
// spins with ld.acquire.sys and reads the data after it.  Slots are double-buffered on the parity
// of the sequence number; a rank can only be one round ahead of the slowest one, because finishing
// round s needs every rank's flag of round s.  Every wait has a time limit (CGB_PEER_TIMEOUT_MS, default 30 s) that stops
// the loop and reports an error instead of hanging the GPU; the handle is unusable afterwards.
//
// There is nothing of this in the reference (single device); SURVEY.md §8(e).
#pragma once
#include "common.cuh"

namespace cgb {

constexpr int kMaxRanks = 16;
// Time limit of every peer wait, in SM clock cycles: set once per process from CGB_PEER_TIMEOUT_MS (default 30 s;
// dist.cu: peer_timeout_cycles) — long enough for any honest rank skew, short enough not to hang a GPU for good.
__device__ long long g_peer_timeout_cycles = 60000000000LL;

struct PeerWin { // one per rank, written by the peers
  double slot[2][kMaxRanks];               // all-reduce contributions, [parity][source rank]
  unsigned long long sflag[2][kMaxRanks];  // sequence number of the contribution
  unsigned long long hflag[kMaxRanks];     // sequence number of the last halo delivered by each source rank
};

struct PeerSeq { // local, per rank; identical on every rank because all ranks run the same sequence
  unsigned long long red;  // all-reduces completed
  unsigned long long halo; // halo exchanges started
};

struct PeerCtx { // kernel argument (by value)
  int rank, nranks;
  PeerWin *const *win; // device array: win[q] = rank q's window (peer-mapped), win[rank] = local
  PeerSeq *seq;
};

#ifdef __CUDACC__
__device__ __forceinline__ void st_release_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_f64(double *p, double v) {
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double *p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

// Spin until *flag >= want.  Returns false on timeout.
__device__ __forceinline__ bool peer_wait(const unsigned long long *flag, unsigned long long want) {
  if (ld_acquire_sys_u64(flag) >= want) return true;
  const long long t0 = clock64();
  while (ld_acquire_sys_u64(flag) < want) {
    if (clock64() - t0 > g_peer_timeout_cycles) return false;
    __nanosleep(20);
  }
  return true;
}

// All-reduce (sum) of one double over the ranks.  Called by ALL threads of ONE CTA per rank (the
// last CTA of a reducing kernel); `v` needs to be valid in thread 0 only.  Returns the sum (rank
// order) in every thread; *ok is cleared on timeout.
__device__ __forceinline__ double peer_allreduce(const PeerCtx &pc, double v, bool *ok) {
  __shared__ double sh_v[kMaxRanks];
  __shared__ double sh_in;
  __shared__ int sh_ok;
  if (threadIdx.x == 0) {
    sh_in = v;
    sh_ok = 1;
  }
  __syncthreads();
  const unsigned long long s = pc.seq->red + 1;
  const int par = (int)(s & 1ull);
  const int q = threadIdx.x;
  if (q < pc.nranks) {
    PeerWin *w = pc.win[q];
    st_relaxed_sys_f64(&w->slot[par][pc.rank], sh_in);
    __threadfence_system();
    st_release_sys_u64(&w->sflag[par][pc.rank], s);
    const PeerWin *me = pc.win[pc.rank];
    if (peer_wait(&me->sflag[par][q], s)) sh_v[q] = ld_relaxed_sys_f64(&me->slot[par][q]);
    else {
      sh_v[q] = 0.0;
      sh_ok = 0;
    }
  }
  __syncthreads();
  double sum = 0.0;
  for (int i = 0; i < pc.nranks; i++) sum += sh_v[i];
  if (threadIdx.x == 0) pc.seq->red = s;
  if (!sh_ok) *ok = false;
  __syncthreads(); // shared buffers may be reused by the next call
  return sum;
}
#endif // __CUDACC__

} // namespace cgb
