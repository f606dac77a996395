// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Host-only stand-in for the subset of the SYCL 2020 / AdaptiveCpp API that the
// reference solver uses (reference: src/CG.hpp:15-17,44; src/LinearAlgebraTypes.hpp:11;
// src/VectorOperations.hpp:17-18). It lets g++ compile the reference's own, unmodified
// sources so they can serve as the parity oracle and as the timed CPU baseline.
//
// Execution model: every submission runs eagerly, in submission order, on the host.
// That is one valid schedule of the reference's event DAG (all depends_on edges point
// backwards in submission order). Two build flavours:
//   * default            : serial, left-to-right reductions  -> deterministic parity oracle
//   * -DSTANDIN_OPENMP   : parallel_for / reductions run under OpenMP -> timed CPU baseline
//                          (mirrors the reference's real CPU path, AdaptiveCpp's OpenMP host backend)
//
// The summation order of sycl::reduction is AdaptiveCpp-internal and not in the reference
// tree; the serial flavour fixes it as: identity + x0 + x1 + ... , then combined with the
// prior value of *ptr (no initialize_to_identity, so reductions ACCUMULATE, as in the reference).
#ifndef ORACLE_SYCL_STANDIN_HPP
#define ORACLE_SYCL_STANDIN_HPP

#include <cmath>
#include <condition_variable>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#ifdef STANDIN_OPENMP
#include <omp.h>
#endif

namespace acpp {
namespace sycl {

// ---- bookkeeping the harness reads (kernel-name -> number of submissions) ---------------
struct standin_stats_t {
  std::map<std::string, long> launches;
  long host_syncs = 0;
  // Bounded timing samples (bench.py only): after `trip_limit` BetaCalculation submissions — one per
  // CG loop trip, reference src/CG.hpp:413 — the stand-in raises the host-visible stop flag the
  // reference allocated with malloc_shared<bool> (src/CG.hpp:289), so the unmodified loop exits
  // after exactly that many trips.  0 = off (the parity oracle never sets it).
  long trip_limit = 0;
  long trips = 0;
  bool *shared_flag = nullptr;
  void clear() { launches.clear(); host_syncs = 0; trips = 0; }
  long count_containing(const std::string &needle) const {
    long c = 0;
    for (auto &kv : launches)
      if (kv.first.find(needle) != std::string::npos) c += kv.second;
    return c;
  }
};
inline standin_stats_t &standin_stats() {
  static standin_stats_t s;
  return s;
}
namespace detail {
struct unnamed_kernel;
template <class Name> inline void note_launch(const char *kind) {
  // Kernel names are incomplete local classes: typeid is unusable, the pretty name is not.
  std::string pf = __PRETTY_FUNCTION__;
  auto pos = pf.find("Name = ");
  std::string nm = pos == std::string::npos ? pf : pf.substr(pos + 7);
  auto end = nm.find_first_of(";]");
  if (end != std::string::npos) nm = nm.substr(0, end);
  standin_stats().launches[std::string(kind) + ":" + nm]++;
  if (nm.find("BetaCalculation") != std::string::npos) {
    auto &st = standin_stats();
    st.trips++;
    if (st.trip_limit > 0 && st.trips >= st.trip_limit && st.shared_flag) *st.shared_flag = true;
  }
}
} // namespace detail

// ---- basic vocabulary ---------------------------------------------------------------------
class exception : public std::exception {
  std::string msg;
public:
  explicit exception(std::string m = "sycl stand-in exception") : msg(std::move(m)) {}
  const char *what() const noexcept override { return msg.c_str(); }
};

template <int D = 1> struct range {
  std::size_t v;
  range(std::size_t n = 0) : v(n) {}
  std::size_t get(int) const { return v; }
  std::size_t operator[](int) const { return v; }
  std::size_t size() const { return v; }
};
template <int D = 1> struct id {
  std::size_t v;
  id(std::size_t n = 0) : v(n) {}
  operator std::size_t() const { return v; }
};
template <int D = 1> struct item {
  std::size_t v, r;
  item(std::size_t i, std::size_t n) : v(i), r(n) {}
  operator std::size_t() const { return v; }
  std::size_t get_id(int = 0) const { return v; }
  std::size_t get_linear_id() const { return v; }
  std::size_t get_range(int = 0) const { return r; }
};
template <int D = 1> struct nd_range {
  range<D> global, local;
  nd_range(range<D> g, range<D> l) : global(g), local(l) {}
};

namespace access {
enum class fence_space { local_space, global_space, global_and_local };
}

namespace detail {
// Reusable barrier for one emulated work-group (threads = work-items).
struct group_barrier {
  std::mutex m;
  std::condition_variable cv;
  std::size_t expected, waiting = 0, generation = 0;
  explicit group_barrier(std::size_t n) : expected(n) {}
  void arrive_and_wait() {
    std::unique_lock<std::mutex> lk(m);
    std::size_t gen = generation;
    if (++waiting == expected) {
      waiting = 0;
      ++generation;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return gen != generation; });
    }
  }
};
} // namespace detail

template <int D = 1> struct nd_item {
  std::size_t gid, lid, grp, lsz;
  detail::group_barrier *bar;
  std::size_t get_local_id(int = 0) const { return lid; }
  std::size_t get_global_id(int = 0) const { return gid; }
  std::size_t get_group(int = 0) const { return grp; }
  std::size_t get_local_range(int = 0) const { return lsz; }
  void barrier(access::fence_space = access::fence_space::global_and_local) const {
    bar->arrive_and_wait();
  }
};

template <class T = void> struct plus {
  T operator()(const T &a, const T &b) const { return a + b; }
};
template <> struct plus<void> {
  template <class A, class B> auto operator()(const A &a, const B &b) const { return a + b; }
};

// reduction descriptor + the reducer object handed to the kernel lambda
template <class T> struct reducer {
  T acc;
  reducer() : acc(T(0)) {}
  reducer &operator+=(const T &v) {
    acc += v;
    return *this;
  }
  reducer &combine(const T &v) {
    acc += v;
    return *this;
  }
};
template <class T> struct reduction_desc {
  T *target;
};
template <class T, class Op> reduction_desc<T> reduction(T *ptr, Op) { return reduction_desc<T>{ptr}; }

class event {
public:
  void wait() {}
  void wait_and_throw() {}
};

namespace info {
namespace device {
struct max_work_group_size {
  using return_type = std::size_t;
};
struct name {
  using return_type = std::string;
};
} // namespace device
} // namespace info

class device {
public:
  template <class Param> typename Param::return_type get_info() const {
    if constexpr (std::is_same_v<Param, info::device::max_work_group_size>)
      return 1024;
    else
      return typename Param::return_type{};
  }
};

class handler;

template <class T, int D = 1> class local_accessor {
  std::shared_ptr<std::vector<T>> store; // one buffer per work-group execution (groups run one at a time)
public:
  local_accessor(std::size_t n, handler &) : store(std::make_shared<std::vector<T>>(n)) {}
  local_accessor(range<1> n, handler &) : store(std::make_shared<std::vector<T>>(n.v)) {}
  T &operator[](std::size_t i) const { return (*store)[i]; }
};

class stream {
public:
  stream(std::size_t, std::size_t, handler &) {}
  template <class T> const stream &operator<<(const T &) const { return *this; }
};

struct interop_handle {};

class handler {
public:
  void depends_on(const event &) {}
  void depends_on(const std::vector<event> &) {}

  template <class Name = detail::unnamed_kernel, class F> void single_task(F f) {
    detail::note_launch<Name>("single_task");
    f();
  }

  // plain range
  template <class Name = detail::unnamed_kernel, class F> void parallel_for(range<1> r, F f) {
    detail::note_launch<Name>("parallel_for");
    const std::size_t n = r.v;
#ifdef STANDIN_OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (std::size_t i = 0; i < n; ++i) f(item<1>(i, n));
  }

  // one reduction
  template <class Name = detail::unnamed_kernel, class T, class F>
  void parallel_for(range<1> r, reduction_desc<T> red, F f) {
    detail::note_launch<Name>("parallel_for_red1");
    const std::size_t n = r.v;
    T total = T(0);
#ifdef STANDIN_OPENMP
#pragma omp parallel
    {
      reducer<T> loc;
#pragma omp for schedule(static) nowait
      for (std::size_t i = 0; i < n; ++i) f(item<1>(i, n), loc);
#pragma omp critical
      total += loc.acc;
    }
#else
    reducer<T> loc;
    for (std::size_t i = 0; i < n; ++i) f(item<1>(i, n), loc);
    total = loc.acc;
#endif
    *red.target = *red.target + total;
  }

  // two reductions
  template <class Name = detail::unnamed_kernel, class T, class U, class F>
  void parallel_for(range<1> r, reduction_desc<T> red1, reduction_desc<U> red2, F f) {
    detail::note_launch<Name>("parallel_for_red2");
    const std::size_t n = r.v;
    T total1 = T(0);
    U total2 = U(0);
#ifdef STANDIN_OPENMP
#pragma omp parallel
    {
      reducer<T> l1;
      reducer<U> l2;
#pragma omp for schedule(static) nowait
      for (std::size_t i = 0; i < n; ++i) f(item<1>(i, n), l1, l2);
#pragma omp critical
      {
        total1 += l1.acc;
        total2 += l2.acc;
      }
    }
#else
    reducer<T> l1;
    reducer<U> l2;
    for (std::size_t i = 0; i < n; ++i) f(item<1>(i, n), l1, l2);
    total1 = l1.acc;
    total2 = l2.acc;
#endif
    *red1.target = *red1.target + total1;
    *red2.target = *red2.target + total2;
  }

  // nd_range with barriers: each work-group is run by local-size real threads, groups one after
  // another. Only reached from dot_product_optimised / dot_product (never from CG::solve).
  template <class Name = detail::unnamed_kernel, class F> void parallel_for(nd_range<1> r, F f) {
    detail::note_launch<Name>("parallel_for_nd");
    const std::size_t lsz = r.local.v, groups = (r.global.v + lsz - 1) / lsz;
    for (std::size_t g = 0; g < groups; ++g) {
      detail::group_barrier bar(lsz);
      std::vector<std::thread> ts;
      ts.reserve(lsz);
      for (std::size_t l = 0; l < lsz; ++l)
        ts.emplace_back([&, l] { f(nd_item<1>{g * lsz + l, l, g, lsz, &bar}); });
      for (auto &t : ts) t.join();
    }
  }

  template <class T> void copy(const T *src, T *dst, std::size_t n) {
    detail::note_launch<detail::unnamed_kernel>("copy");
    std::memcpy(dst, src, n * sizeof(T));
  }
  template <class T> void fill(T *dst, const T &v, std::size_t n) {
    detail::note_launch<detail::unnamed_kernel>("fill");
    for (std::size_t i = 0; i < n; ++i) dst[i] = v;
  }
  template <class F> void AdaptiveCpp_enqueue_custom_operation(F f) {
    interop_handle h;
    f(h);
  }
};

class queue {
public:
  queue() {}
  template <class F> event submit(F f) {
    handler h;
    f(h);
    return event{};
  }
  template <class T> event fill(T *dst, const T &v, std::size_t n) {
    detail::note_launch<detail::unnamed_kernel>("fill");
#ifdef STANDIN_OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (std::size_t i = 0; i < n; ++i) dst[i] = v;
    return event{};
  }
  template <class T> event copy(const T *src, T *dst, std::size_t n) {
    detail::note_launch<detail::unnamed_kernel>("copy");
    std::memcpy(dst, src, n * sizeof(T));
    return event{};
  }
  template <class T> event memcpy(T *dst, const T *src, std::size_t bytes) {
    std::memcpy(dst, src, bytes);
    return event{};
  }
  template <class F> event single_task(F f) {
    detail::note_launch<detail::unnamed_kernel>("single_task");
    f();
    return event{};
  }
  void wait() { standin_stats().host_syncs++; }
  void wait_and_throw() { standin_stats().host_syncs++; }
  device get_device() const { return device{}; }
};

template <class T> T *malloc_device(std::size_t n, const queue &) {
  return static_cast<T *>(std::malloc(n * sizeof(T) > 0 ? n * sizeof(T) : 1));
}
template <class T> T *malloc_shared(std::size_t n, const queue &) {
  T *p = static_cast<T *>(std::malloc(n * sizeof(T) > 0 ? n * sizeof(T) : 1));
  if constexpr (std::is_same<T, bool>::value) standin_stats().shared_flag = p;
  return p;
}
template <class T> T *malloc_host(std::size_t n, const queue &) {
  return static_cast<T *>(std::malloc(n * sizeof(T) > 0 ? n * sizeof(T) : 1));
}
inline void free(void *p, const queue &) {
  if (p == static_cast<void *>(standin_stats().shared_flag)) standin_stats().shared_flag = nullptr;
  std::free(p);
}

using std::sqrt;
using std::fabs;
using std::isnan;

} // namespace sycl
} // namespace acpp

#endif // ORACLE_SYCL_STANDIN_HPP
