// AdaptiveCpp/sycl/sycl.hpp — compatibility shim, NOT a SYCL implementation.
//
// The reference's headers take an `acpp::sycl::queue` everywhere and return `acpp::sycl::event`
// (reference src/CG.hpp:44,61; src/LinearAlgebraTypes.hpp:60-63,146-149,213; src/VectorOperations.hpp:91).
// Code written against that API keeps compiling when it includes this file from the cgb200
// include directory: `queue` is a shared handle to a cgb200 context (one B200 + one in-order CUDA
// stream) and `event` wraps a CUDA event recorded on that stream.  Because the stream is in order,
// every depends_on() list of the reference is satisfied by construction.
//
// Only the host-side subset that the CGSolver classes and test/Tester.cpp need is provided
// (queue: copy / fill / wait / wait_and_throw; event: wait; malloc_device / free).  Device lambdas
// (parallel_for, single_task, reduction) are deliberately absent: every kernel of this build is a
// hand-written sm_100a CUDA kernel behind the C-ABI in cgb200.h.
#ifndef CGB200_SYCL_COMPAT_HPP
#define CGB200_SYCL_COMPAT_HPP

#include <cstddef>
#include <cstdlib>
#include <exception>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../../cgb200.h"

namespace acpp {
namespace sycl {

class exception : public std::runtime_error {
public:
  explicit exception(const std::string &what) : std::runtime_error(what) {}
};

namespace detail {
inline void check(int rc) {
  if (rc != CGB_OK) throw exception(cgb_last_error());
}
struct ctx_holder {
  cgb_ctx *ctx = nullptr;
  ctx_holder() {
    const char *dev = std::getenv("CGB_DEVICE");
    check(cgb_ctx_create(dev ? std::atoi(dev) : 0, &ctx));
  }
  ~ctx_holder() { cgb_ctx_destroy(ctx); }
  ctx_holder(const ctx_holder &) = delete;
  ctx_holder &operator=(const ctx_holder &) = delete;
};
struct event_holder {
  cgb_event *ev = nullptr;
  ~event_holder() { cgb_event_destroy(ev); }
};
} // namespace detail

class event {
public:
  event() = default;
  explicit event(cgb_ctx *ctx) : _h(std::make_shared<detail::event_holder>()) {
    detail::check(cgb_event_record(ctx, &_h->ev));
  }
  void wait() {
    if (_h && _h->ev) detail::check(cgb_event_sync(_h->ev));
  }
  cgb_event *native() const { return _h ? _h->ev : nullptr; }

private:
  std::shared_ptr<detail::event_holder> _h;
};

class queue {
public:
  queue() : _h(std::make_shared<detail::ctx_holder>()) {}
  cgb_ctx *native() const { return _h->ctx; }

  void wait() { detail::check(cgb_ctx_sync(native())); }
  void wait_and_throw() { wait(); }

  // USM copy: direction is inferred from the pointers (unified addressing)
  template <class T> event copy(const T *src, T *dst, std::size_t count) {
    detail::check(cgb_memcpy(native(), dst, src, count * sizeof(T), 0));
    return event(native());
  }
  template <class T> event fill(T *ptr, const T &value, std::size_t count) {
    if constexpr (std::is_same<T, double>::value) {
      detail::check(cgb_fill_f64(native(), ptr, value, count));
    } else {
      if (value != T()) throw exception("acpp::sycl compat: fill supports double, or zero for other types");
      detail::check(cgb_memset(native(), ptr, 0, count * sizeof(T)));
    }
    return event(native());
  }
  event memset(void *ptr, int value, std::size_t bytes) {
    detail::check(cgb_memset(native(), ptr, value, bytes));
    return event(native());
  }

private:
  std::shared_ptr<detail::ctx_holder> _h;
};

template <class T> T *malloc_device(std::size_t count, const queue &q) {
  void *p = nullptr;
  detail::check(cgb_malloc(q.native(), count * sizeof(T), &p));
  return static_cast<T *>(p);
}
inline void free(void *ptr, const queue &q) { cgb_free(q.native(), ptr); }

} // namespace sycl
} // namespace acpp

#endif // CGB200_SYCL_COMPAT_HPP
