// VectorOperations.hpp — the per-op kernels of the CGSolver API, executed on a B200.
//
// Drop-in for the reference header of the same name (reference src/VectorOperations.hpp:39-488):
// same class template, method names, argument order and default arguments.  Each method enqueues
// one hand-written sm_100a kernel through the C-ABI (cgb200.h) on the queue's in-order stream and
// returns an event recorded behind it; the `dependencies` lists are accepted and ignored because
// stream order already satisfies them.
//
// Semantics kept from the reference:
//   * scalar operands (a, b) are DEVICE pointers read when the kernel runs;
//   * every reduction ADDS INTO *result (sycl::reduction without initialize_to_identity, :300);
//   * operations work on `vector_size` elements set by setVectorSize(), the trailing size/count
//     argument is ignored exactly as in the reference;
//   * sapbx / sambx / saxpby are safe with Result aliasing X or Y (src/CG.hpp:390,392,418).
#ifndef CGB200_VECTOR_OPERATIONS_HPP
#define CGB200_VECTOR_OPERATIONS_HPP

#include <AdaptiveCpp/sycl/sycl.hpp>

#include <algorithm>
#include <cassert>
#include <cstddef>
#include <type_traits>
#include <vector>

#include "LinearAlgebraTypes.hpp"

namespace CGSolver {

template <class DT, Debuglevel debug = Debuglevel::None> class VectorOperations {
  static_assert(std::is_same<DT, double>::value, "the cgb200 kernels are fp64 (the reference instantiates double only)");

public:
  VectorOperations(asycl::queue &queue) : _queue(queue), workgroupsize(calculateWorkgroupSize()), vector_size(0) {}

  void setVectorSize(size_t size) { this->vector_size = size; }

  // *result += sum left[i] * right[i]     (reference :110-208; tree reduction there, one
  // deterministic two-stage kernel here)
  asycl::event dot_product_optimised(Vector<DT> &left, Vector<DT> &right, DT *result,
                                     std::vector<asycl::event> dependencies = {}, size_t count = 0) {
    (void)dependencies;
    (void)count;
    assert(vector_size != 0);
    return dot(left.ptr(), right.ptr(), result);
  }

  // deprecated raw-pointer flavour (reference :212-285)
  asycl::event dot_product(DT *left, DT *right, DT *result, std::vector<asycl::event> dependencies = {},
                           size_t vec_size = 0) {
    (void)dependencies;
    (void)vec_size;
    assert(vector_size != 0);
    return dot(left, right, result);
  }

  // reference :287-309
  asycl::event dot_product_trivial(Vector<DT> &left, Vector<DT> &right, Scalar<DT> &result,
                                   std::vector<asycl::event> dependencies = {}, std::size_t size = 0) {
    (void)dependencies;
    (void)size;
    return dot(left.ptr(), right.ptr(), result.ptr());
  }

  // *result += sum v[i]^2 (no square root)      reference :311-331
  asycl::event norm(Vector<DT> &vector, Scalar<DT> &result, std::vector<asycl::event> dependencies = {},
                    std::size_t size = 0) {
    (void)dependencies;
    (void)size;
    asycl::detail::check(cgb_norm2_acc(_queue.native(), (int64_t)vector_size, vector.ptr(), result.ptr()));
    return asycl::event(_queue.native());
  }

  // Result = (*a) X + (*b) Y                    reference :349-367
  inline asycl::event saxpby(Vector<DT> &X, Vector<DT> &Y, DT *a, DT *b, Vector<DT> &Result,
                             std::vector<asycl::event> events = {}, size_t vec_size = 0) {
    (void)events;
    (void)vec_size;
    asycl::detail::check(cgb_axpby(_queue.native(), (int64_t)vector_size, X.ptr(), Y.ptr(), a, b, Result.ptr()));
    return asycl::event(_queue.native());
  }

  // Result = X - (*b) Y                         reference :380-397
  inline asycl::event sambx(Vector<DT> &X, Vector<DT> &Y, DT *b, Vector<DT> &Result,
                            std::vector<asycl::event> events = {}, size_t count = 0) {
    (void)events;
    (void)count;
    asycl::detail::check(cgb_xpby(_queue.native(), (int64_t)vector_size, X.ptr(), Y.ptr(), b, -1, Result.ptr()));
    return asycl::event(_queue.native());
  }

  // Result = X + (*b) Y                         reference :410-428
  inline asycl::event sapbx(Vector<DT> &X, Vector<DT> &Y, DT *b, Vector<DT> &Result,
                            std::vector<asycl::event> events = {}, size_t count = 0) {
    (void)events;
    (void)count;
    asycl::detail::check(cgb_xpby(_queue.native(), (int64_t)vector_size, X.ptr(), Y.ptr(), b, +1, Result.ptr()));
    return asycl::event(_queue.native());
  }

  // Result = A vec (CSR)                        reference :438-466
  inline asycl::event spmv(Matrix<DT> &A, Vector<DT> &vec, Vector<DT> &Result, size_t NNZ,
                           std::vector<asycl::event> events = {}, size_t count = 0) {
    (void)NNZ;
    (void)events;
    (void)count;
    assert(vector_size != 0 && A.N() == vector_size);
    asycl::detail::check(cgb_spmv(_queue.native(), A.csr(), vec.ptr(), Result.ptr()));
    return asycl::event(_queue.native());
  }

private:
  asycl::event dot(const DT *l, const DT *r, DT *result) {
    asycl::detail::check(cgb_dot_acc(_queue.native(), (int64_t)vector_size, l, r, result));
    return asycl::event(_queue.native());
  }

  asycl::queue _queue;
  size_t workgroupsize;
  size_t vector_size;

  // The reference caps its work-group size at min(128, device maximum) (:478-487).  The CUDA
  // kernels choose their own launch shape; the value is kept for API completeness only.
  size_t calculateWorkgroupSize() { return std::min(static_cast<size_t>(128), static_cast<size_t>(1024)); }
};

} // namespace CGSolver

#endif // CGB200_VECTOR_OPERATIONS_HPP
