// core.cu — context, device memory, events, error reporting of the cgb200 C-ABI.
// Interfaces replaced: acpp::sycl::queue / malloc_device / free / copy / fill / event as used by
// the reference in src/LinearAlgebraTypes.hpp:43-49,101-121,160-183,213-224 and src/CG.hpp:561-578.
#define CGB_NO_ALLOC_MACROS
#include "common.cuh"

#include <stdarg.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

namespace cgb {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int ctx_scratch(cgb_ctx *ctx) {
  if (ctx->partials) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_CUDA(cached_malloc((void **)&ctx->partials, sizeof(double) * 2 * kMaxPartials));
  CGB_CUDA(cached_malloc((void **)&ctx->ticket, sizeof(unsigned int) * 4));
  CGB_CUDA(cudaMemsetAsync(ctx->ticket, 0, sizeof(unsigned int) * 4, ctx->stream));
  return CGB_OK;
}

// ---- device-memory cache (see common.cuh) ---------------------------------------------------------------------
namespace {
struct DevCache {
  std::multimap<size_t, void *> free_blocks;   // size -> block
  std::unordered_map<void *, size_t> live;     // blocks handed out
  size_t cached_bytes = 0;
  size_t cap = 0; // most bytes kept parked on this device (0: not yet determined)
};
struct AllocCache {
  std::mutex mu;
  std::map<int, DevCache> dev;
  std::multimap<size_t, void *> host_free;
  std::unordered_map<void *, size_t> host_live;
  bool on = true;
  size_t max_bytes = 0; // CGB_ALLOC_CACHE_MAX_GB; 0: per device, a quarter of its memory, at most 32 GB
  AllocCache() {
    if (const char *e = getenv("CGB_ALLOC_CACHE")) on = atoi(e) != 0;
    if (const char *e = getenv("CGB_ALLOC_CACHE_MAX_GB")) max_bytes = (size_t)(atof(e) * (double)((size_t)1 << 30));
  }
};
// Parked blocks are invisible to every other allocator of the process (NCCL, cub, torch, cupy): the default bound is
// a quarter of the device's memory and never more than 32 GB -- room for the largest end-to-end case measured
// (C4: 19.3 GB of matrix and work vectors per call) and no more.  Call with `dev` current.
size_t device_cap(const AllocCache &c, DevCache &d) {
  if (d.cap) return d.cap;
  if (c.max_bytes) return d.cap = c.max_bytes;
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) {
    cudaGetLastError();
    total_b = (size_t)64 << 30;
  }
  return d.cap = std::min((size_t)32 << 30, total_b / 4);
}
AllocCache &alloc_cache() {
  static AllocCache *c = new AllocCache(); // never destroyed: frees may come from static destructors
  return *c;
}
void trim_device_locked(DevCache &d, size_t keep) {
  while (d.cached_bytes > keep && !d.free_blocks.empty()) {
    auto it = std::prev(d.free_blocks.end()); // largest first
    cudaFree(it->second);
    d.cached_bytes -= it->first;
    d.free_blocks.erase(it);
  }
}
} // namespace

cudaError_t cached_malloc(void **p, size_t bytes) {
  AllocCache &c = alloc_cache();
  if (!c.on) return cudaMalloc(p, bytes);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const size_t want = ((bytes ? bytes : 1) + 511) & ~(size_t)511;
  std::lock_guard<std::mutex> lk(c.mu);
  DevCache &d = c.dev[dev];
  auto it = d.free_blocks.lower_bound(want);
  if (it != d.free_blocks.end() && it->first <= want + want / 4 + 4096) {
    *p = it->second;
    d.cached_bytes -= it->first;
    d.live[it->second] = it->first;
    d.free_blocks.erase(it);
    return cudaSuccess;
  }
  e = cudaMalloc(p, want);
  if (e != cudaSuccess) { // out of memory with blocks parked in the cache: give them back and try once more
    cudaGetLastError();
    cudaDeviceSynchronize();
    trim_device_locked(d, 0);
    e = cudaMalloc(p, want);
  }
  if (e == cudaSuccess) d.live[*p] = want;
  return e;
}

cudaError_t cached_free(void *p) {
  if (!p) return cudaSuccess;
  AllocCache &c = alloc_cache();
  if (!c.on) return cudaFree(p);
  cudaPointerAttributes at;
  int dev = 0;
  if (cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeDevice) dev = at.device;
  else {
    cudaGetLastError();
    cudaGetDevice(&dev);
  }
  std::unique_lock<std::mutex> lk(c.mu);
  DevCache &d = c.dev[dev];
  auto it = d.live.find(p);
  if (it == d.live.end()) { // not ours (e.g. adopted from the caller): the driver's own free
    lk.unlock();
    return cudaFree(p);
  }
  const size_t sz = it->second;
  d.live.erase(it);
  lk.unlock();
  ScopedDevice sd(dev); // the block's device, whatever the caller has current
  cudaError_t e = cudaDeviceSynchronize(); // cudaFree's semantics: nothing in flight touches the block once it changes hands
  lk.lock();
  d.free_blocks.emplace(sz, p);
  d.cached_bytes += sz;
  const size_t cap = device_cap(c, d);
  if (d.cached_bytes > cap) trim_device_locked(d, cap);
  return e;
}

cudaError_t cached_host_alloc(void **p, size_t bytes) {
  AllocCache &c = alloc_cache();
  const size_t want = ((bytes ? bytes : 1) + 4095) & ~(size_t)4095;
  if (c.on && want <= ((size_t)1 << 20)) {
    std::lock_guard<std::mutex> lk(c.mu);
    auto it = c.host_free.find(want);
    if (it != c.host_free.end()) {
      *p = it->second;
      c.host_live[*p] = want;
      c.host_free.erase(it);
      return cudaSuccess;
    }
    cudaError_t e = cudaMallocHost(p, want);
    if (e == cudaSuccess) c.host_live[*p] = want;
    return e;
  }
  return cudaMallocHost(p, want);
}

cudaError_t cached_host_free(void *p) {
  if (!p) return cudaSuccess;
  AllocCache &c = alloc_cache();
  {
    std::lock_guard<std::mutex> lk(c.mu);
    auto it = c.host_live.find(p);
    if (it != c.host_live.end()) { // small pinned blocks (state mirrors): kept for the next workspace
      c.host_free.emplace(it->second, p);
      c.host_live.erase(it);
      return cudaSuccess;
    }
  }
  return cudaFreeHost(p);
}

void cache_trim() {
  AllocCache &c = alloc_cache();
  std::lock_guard<std::mutex> lk(c.mu);
  int prev = -1;
  cudaGetDevice(&prev);
  for (auto &kv : c.dev) {
    cudaSetDevice(kv.first);
    cudaDeviceSynchronize();
    trim_device_locked(kv.second, 0);
  }
  if (prev >= 0) cudaSetDevice(prev);
  for (auto &kv : c.host_free) cudaFreeHost(kv.second);
  c.host_free.clear();
}

// ---- pageable host memory: staged copies --------------------------------------------------------------------
// The reference's upload / download sites hand the queue plain std::vector storage (src/LinearAlgebraTypes.hpp:
// 109-119, :177-183; src/CG.hpp:517-532).  A cudaMemcpyAsync from pageable memory goes through the driver's own
// single-threaded staging at a fraction of the PCIe rate, so copies of kStageMin bytes or more whose host side is
// not page-locked are pipelined here instead: kStageWorkers host threads move 4 MiB chunks between the caller's
// buffer and their own pair of pinned bounce buffers while the copy engine moves the previous chunks, each worker
// on its own stream, fenced against the context's stream with events on both sides.  When the call returns the
// caller's buffer has been read completely (upload) or written completely (download).
namespace {
constexpr size_t kStageMin = (size_t)8 << 20;
constexpr size_t kStageChunk = (size_t)4 << 20;
constexpr int kStageWorkersMax = 8;

struct Stager { // one per device: streams, events and bounce buffers live in that device's context
  std::mutex mu; // one staged copy at a time per device
  int device = -1;
  int workers = 0;
  void *buf[kStageWorkersMax][2] = {};
  cudaEvent_t ev[kStageWorkersMax][2] = {};
  cudaEvent_t done[kStageWorkersMax] = {};
  cudaStream_t st[kStageWorkersMax] = {};
  cudaEvent_t start = nullptr;
  bool ok = false;
  bool failed = false; // set-up failed once: this device keeps the driver's path, no retry
};
Stager &stager(int device) {
  static std::mutex mu;
  static std::map<int, Stager *> *all = new std::map<int, Stager *>(); // never destroyed (copies may run from static destructors)
  std::lock_guard<std::mutex> lk(mu);
  Stager *&s = (*all)[device];
  if (!s) s = new Stager();
  return *s;
}

// give back whatever a failed set-up had created
void stager_release(Stager &s) {
  for (int i = 0; i < kStageWorkersMax; i++) {
    for (int b = 0; b < 2; b++) {
      if (s.buf[i][b]) cudaFreeHost(s.buf[i][b]);
      if (s.ev[i][b]) cudaEventDestroy(s.ev[i][b]);
      s.buf[i][b] = nullptr;
      s.ev[i][b] = nullptr;
    }
    if (s.done[i]) cudaEventDestroy(s.done[i]);
    if (s.st[i]) cudaStreamDestroy(s.st[i]);
    s.done[i] = nullptr;
    s.st[i] = nullptr;
  }
  if (s.start) cudaEventDestroy(s.start);
  s.start = nullptr;
  cudaGetLastError();
}

// called with s.mu held and `device` current
bool stager_init(Stager &s, int device) {
  if (s.ok) return true;
  if (s.failed) return false;
  int hw = (int)std::thread::hardware_concurrency();
  int w = hw >= 16 ? 6 : hw >= 8 ? 4 : 2;
  if (const char *e = getenv("CGB_STAGE_THREADS")) w = std::max(1, std::min(kStageWorkersMax, atoi(e)));
  bool good = cudaEventCreateWithFlags(&s.start, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; i < w && good; i++) {
    good = cudaStreamCreateWithFlags(&s.st[i], cudaStreamNonBlocking) == cudaSuccess &&
           cudaEventCreateWithFlags(&s.done[i], cudaEventDisableTiming) == cudaSuccess;
    for (int b = 0; b < 2 && good; b++)
      good = cudaMallocHost(&s.buf[i][b], kStageChunk) == cudaSuccess &&
             cudaEventCreateWithFlags(&s.ev[i][b], cudaEventDisableTiming) == cudaSuccess;
  }
  if (!good) {
    cudaGetLastError();
    stager_release(s);
    s.failed = true;
    return false;
  }
  s.workers = w;
  s.device = device;
  s.ok = true;
  return true;
}

bool host_is_pageable(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return at.type == cudaMemoryTypeUnregistered;
}

// to_device: h -> d, else d -> h.  Returns false when staging is not available (caller falls back to a plain copy).
bool staged_copy(cgb_ctx *ctx, void *d, void *h, size_t bytes, bool to_device, cudaError_t *err) {
  if (const char *e = getenv("CGB_STAGE"))
    if (atoi(e) == 0) return false;
  Stager &s = stager(ctx->device);
  std::lock_guard<std::mutex> lk(s.mu);
  ScopedDevice sd(ctx->device);
  if (!stager_init(s, ctx->device)) return false;
  const size_t nchunks = (bytes + kStageChunk - 1) / kStageChunk;
  const int W = (int)std::min<size_t>((size_t)s.workers, nchunks);
  *err = cudaEventRecord(s.start, ctx->stream); // the device side of the copy is ordered after what the stream holds
  if (*err != cudaSuccess) return true;
  std::vector<cudaError_t> rc((size_t)W, cudaSuccess);
  auto work = [&](int w) {
    cudaSetDevice(s.device);
    cudaError_t e = cudaStreamWaitEvent(s.st[w], s.start, 0);
    size_t k = 0;
    char *hp = static_cast<char *>(h), *dp = static_cast<char *>(d);
    size_t prev_off = 0, prev_len = 0;
    int prev_b = 0;
    for (size_t c = (size_t)w; c < nchunks && e == cudaSuccess; c += (size_t)W, k++) {
      const int b = (int)(k & 1);
      const size_t off = c * kStageChunk, len = std::min(kStageChunk, bytes - off);
      e = cudaEventSynchronize(s.ev[w][b]); // the buffer's previous transfer (this call or the last one) is over
      if (e != cudaSuccess) break;
      if (to_device) {
        memcpy(s.buf[w][b], hp + off, len);
        e = cudaMemcpyAsync(dp + off, s.buf[w][b], len, cudaMemcpyHostToDevice, s.st[w]);
        if (e == cudaSuccess) e = cudaEventRecord(s.ev[w][b], s.st[w]);
      } else {
        e = cudaMemcpyAsync(s.buf[w][b], dp + off, len, cudaMemcpyDeviceToHost, s.st[w]);
        if (e == cudaSuccess) e = cudaEventRecord(s.ev[w][b], s.st[w]);
        if (k > 0 && e == cudaSuccess) { // the previous chunk has landed while this one is in flight
          e = cudaEventSynchronize(s.ev[w][prev_b]);
          memcpy(hp + prev_off, s.buf[w][prev_b], prev_len);
        }
        prev_off = off;
        prev_len = len;
        prev_b = b;
      }
    }
    if (!to_device && k > 0 && e == cudaSuccess) {
      e = cudaEventSynchronize(s.ev[w][prev_b]);
      memcpy(hp + prev_off, s.buf[w][prev_b], prev_len);
    }
    if (e == cudaSuccess) e = cudaEventRecord(s.done[w], s.st[w]);
    rc[(size_t)w] = e;
  };
  std::vector<std::thread> th;
  for (int w = 1; w < W; w++) th.emplace_back(work, w);
  work(0);
  for (auto &t : th) t.join();
  *err = cudaSuccess;
  for (int w = 0; w < W; w++) {
    if (rc[(size_t)w] != cudaSuccess) *err = rc[(size_t)w];
    else if (cudaStreamWaitEvent(ctx->stream, s.done[w], 0) != cudaSuccess) *err = cudaGetLastError();
  }
  return true;
}
} // namespace

// Host <-> device copy on the context's stream; pageable host memory of kStageMin bytes or more is staged.
cudaError_t copy_h2d(cgb_ctx *ctx, void *d_dst, const void *h_src, size_t bytes) {
  cudaError_t e = cudaSuccess;
  if (bytes >= kStageMin && host_is_pageable(h_src) && staged_copy(ctx, d_dst, const_cast<void *>(h_src), bytes, true, &e)) return e;
  return cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, ctx->stream);
}
cudaError_t copy_d2h(cgb_ctx *ctx, void *h_dst, const void *d_src, size_t bytes) {
  cudaError_t e = cudaSuccess;
  if (bytes >= kStageMin && host_is_pageable(h_dst) && staged_copy(ctx, const_cast<void *>(d_src), h_dst, bytes, false, &e)) return e;
  return cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream);
}

} // namespace cgb

using namespace cgb;

extern "C" {

const char *cgb_last_error(void) { return g_err; }
int cgb_version(void) { return CGB_VERSION; }

int cgb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int cgb_ctx_create(int device, cgb_ctx **out) {
  CGB_REQUIRE(out != nullptr, "cgb_ctx_create: out is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return fail(CGB_ERR_CUDA, "cgb_ctx_create: no CUDA device available (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  if (device < 0 || device >= count) return fail(CGB_ERR_INVALID, "cgb_ctx_create: device %d out of range [0,%d)", device, count);
  // cudaGetDeviceProperties costs 2-3 ms (and now and then 100 ms: it walks driver state); callers of the reference API
  // create a solver, and with it a context, per system, so the answer is kept per device for the life of the process
  static std::mutex prop_mu;
  static std::map<int, cudaDeviceProp> prop_of;
  cudaDeviceProp prop;
  {
    std::lock_guard<std::mutex> lk(prop_mu);
    auto it = prop_of.find(device);
    if (it == prop_of.end()) {
      CGB_CUDA(cudaGetDeviceProperties(&prop, device));
      prop_of[device] = prop;
    } else {
      prop = it->second;
    }
  }
  if (prop.major < 10)
    return fail(CGB_ERR_CUDA, "cgb_ctx_create: device %d is sm_%d%d; this library ships sm_100a code only", device,
                prop.major, prop.minor);
  ScopedDevice sd(device);
  cgb_ctx *ctx = new cgb_ctx();
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->coop_ok = prop.cooperativeLaunch;
  ctx->l2_bytes = (size_t)prop.l2CacheSize;
  ctx->l2_persist_max = (size_t)prop.persistingL2CacheMaxSize;
  ctx->l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
  e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete ctx;
    return fail(CGB_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e));
  }
  e = cached_host_alloc((void **)&ctx->pinned, 4096);
  if (e != cudaSuccess) {
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return fail(CGB_ERR_CUDA, "cudaMallocHost failed: %s", cudaGetErrorString(e));
  }
  *out = ctx;
  return CGB_OK;
}

int cgb_ctx_destroy(cgb_ctx *ctx) {
  if (!ctx) return CGB_OK;
  ScopedDevice sd(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->partials) cached_free(ctx->partials);
  if (ctx->ticket) cached_free(ctx->ticket);
  if (ctx->pinned) cached_host_free(ctx->pinned);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return CGB_OK;
}

int cgb_ctx_sync(cgb_ctx *ctx) {
  CGB_REQUIRE(ctx, "cgb_ctx_sync: ctx is NULL");
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int cgb_ctx_set_progress(cgb_ctx *ctx, cgb_progress_fn fn, void *user) {
  CGB_REQUIRE(ctx, "cgb_ctx_set_progress: ctx is NULL");
  ctx->progress = fn;
  ctx->progress_user = user;
  return CGB_OK;
}

void *cgb_ctx_stream(cgb_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
int cgb_ctx_device(cgb_ctx *ctx) { return ctx ? ctx->device : -1; }
int cgb_ctx_sm_count(cgb_ctx *ctx) { return ctx ? ctx->sm_count : 0; }

int cgb_malloc(cgb_ctx *ctx, size_t bytes, void **d_out) {
  CGB_REQUIRE(ctx && d_out, "cgb_malloc: NULL argument");
  ScopedDevice sd(ctx->device);
  *d_out = nullptr;
  if (bytes == 0) bytes = 16; // cudaMalloc aligns to >= 256 bytes
  CGB_CUDA(cached_malloc(d_out, bytes));
  return CGB_OK;
}

int cgb_free(cgb_ctx *ctx, void *d_ptr) {
  if (!d_ptr) return CGB_OK;
  CGB_REQUIRE(ctx, "cgb_free: ctx is NULL");
  ScopedDevice sd(ctx->device);
  // work queued on the stream may still use the buffer (reference: free after queue order): the cache
  // synchronises the device before the block can be handed out again
  CGB_CUDA(cached_free(d_ptr));
  return CGB_OK;
}

int cgb_memcpy_h2d(cgb_ctx *ctx, void *d_dst, const void *h_src, size_t bytes, int sync) {
  CGB_REQUIRE(ctx && (bytes == 0 || (d_dst && h_src)), "cgb_memcpy_h2d: NULL argument");
  if (bytes == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_CUDA(copy_h2d(ctx, d_dst, h_src, bytes));
  if (sync) CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int cgb_memcpy_d2h(cgb_ctx *ctx, void *h_dst, const void *d_src, size_t bytes, int sync) {
  CGB_REQUIRE(ctx && (bytes == 0 || (h_dst && d_src)), "cgb_memcpy_d2h: NULL argument");
  if (bytes == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_CUDA(copy_d2h(ctx, h_dst, d_src, bytes));
  if (sync) CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int cgb_memcpy_d2d(cgb_ctx *ctx, void *d_dst, const void *d_src, size_t bytes) {
  CGB_REQUIRE(ctx && (bytes == 0 || (d_dst && d_src)), "cgb_memcpy_d2d: NULL argument");
  if (bytes == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_CUDA(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
  return CGB_OK;
}

int cgb_memcpy(cgb_ctx *ctx, void *dst, const void *src, size_t bytes, int sync) {
  CGB_REQUIRE(ctx && (bytes == 0 || (dst && src)), "cgb_memcpy: NULL argument");
  if (bytes == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  // direction from the pointers; a pageable host side goes through the staged path (the reference's std::vector
  // uploads and downloads: src/LinearAlgebraTypes.hpp:109-119, src/CG.hpp:517-532)
  cudaPointerAttributes as, ad;
  const bool src_dev = cudaPointerGetAttributes(&as, src) == cudaSuccess && (as.type == cudaMemoryTypeDevice || as.type == cudaMemoryTypeManaged);
  const bool dst_dev = cudaPointerGetAttributes(&ad, dst) == cudaSuccess && (ad.type == cudaMemoryTypeDevice || ad.type == cudaMemoryTypeManaged);
  cudaGetLastError();
  if (dst_dev && !src_dev) CGB_CUDA(copy_h2d(ctx, dst, src, bytes));
  else if (src_dev && !dst_dev) CGB_CUDA(copy_d2h(ctx, dst, src, bytes));
  else CGB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ctx->stream));
  // pageable host memory may still be read after an async copy returns: callers that pass
  // borrowed std::vector storage rely on the wait the reference issues next (src/CG.hpp:169)
  if (sync) CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int cgb_memset(cgb_ctx *ctx, void *d_ptr, int byte, size_t bytes) {
  CGB_REQUIRE(ctx && (bytes == 0 || d_ptr), "cgb_memset: NULL argument");
  if (bytes == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_CUDA(cudaMemsetAsync(d_ptr, byte, bytes, ctx->stream));
  return CGB_OK;
}

int cgb_host_alloc(size_t bytes, void **h_out) {
  CGB_REQUIRE(h_out, "cgb_host_alloc: out is NULL");
  *h_out = nullptr;
  CGB_CUDA(cudaMallocHost(h_out, bytes ? bytes : 16));
  return CGB_OK;
}

int cgb_host_free(void *h_ptr) {
  if (h_ptr) CGB_CUDA(cudaFreeHost(h_ptr));
  return CGB_OK;
}

int cgb_event_record(cgb_ctx *ctx, cgb_event **out) {
  CGB_REQUIRE(ctx && out, "cgb_event_record: NULL argument");
  ScopedDevice sd(ctx->device);
  cgb_event *ev = new cgb_event();
  cudaError_t e = cudaEventCreate(&ev->ev);
  if (e == cudaSuccess) e = cudaEventRecord(ev->ev, ctx->stream);
  if (e != cudaSuccess) {
    if (ev->ev) cudaEventDestroy(ev->ev);
    delete ev;
    *out = nullptr;
    return fail(CGB_ERR_CUDA, "cgb_event_record: %s", cudaGetErrorString(e));
  }
  *out = ev;
  return CGB_OK;
}

int cgb_event_sync(cgb_event *ev) {
  CGB_REQUIRE(ev, "cgb_event_sync: event is NULL");
  CGB_CUDA(cudaEventSynchronize(ev->ev));
  return CGB_OK;
}

int cgb_event_elapsed_ms(cgb_event *start, cgb_event *stop, double *ms) {
  CGB_REQUIRE(start && stop && ms, "cgb_event_elapsed_ms: NULL argument");
  float f = 0.f;
  CGB_CUDA(cudaEventSynchronize(stop->ev));
  CGB_CUDA(cudaEventElapsedTime(&f, start->ev, stop->ev));
  *ms = (double)f;
  return CGB_OK;
}

int cgb_event_destroy(cgb_event *ev) {
  if (!ev) return CGB_OK;
  cudaEventDestroy(ev->ev);
  delete ev;
  return CGB_OK;
}

int cgb_trim(void) {
  cache_trim();
  return CGB_OK;
}

size_t cgb_memory_footprint(int64_t n, int64_t nnz) {
  // reference formula, src/CG.hpp:555-558
  return (size_t)(2 * nnz + 4 * n) * sizeof(double) + (size_t)(2 * n) * sizeof(int);
}

} // extern "C"
