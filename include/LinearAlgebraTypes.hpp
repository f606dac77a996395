// LinearAlgebraTypes.hpp — device storage classes of the CGSolver API on a B200.
//
// Drop-in for the reference header of the same name (reference src/LinearAlgebraTypes.hpp):
// same namespace, class names, template parameters, constructors and accessors, so code written
// against the reference (its test/Tester.cpp included) compiles unchanged.  Nothing here computes:
// the classes own device memory obtained through the C-ABI (cgb200.h) and, for Matrix, the opaque
// cgb_csr handle that carries the SpMV launch plan and the cached CG workspace.
//
//   CGSolver::Debuglevel        reference :26-30
//   CGSolver::Asycl_deleter     reference :43-49
//   CGSolver::Matrix<DT>        reference :57-132   (CSR: DT values, int columns, int row offsets)
//   CGSolver::Vector<DT>        reference :143-203
//   CGSolver::Scalar<DT>        reference :210-250
#ifndef CGB200_LINEAR_ALGEBRA_TYPES_HPP
#define CGB200_LINEAR_ALGEBRA_TYPES_HPP

#include <AdaptiveCpp/sycl/sycl.hpp>

#include <cassert>
#include <cstddef>
#include <memory>
#include <type_traits>
#include <string>
#include <vector>

namespace CGSolver {

namespace asycl = acpp::sycl;

enum Debuglevel {
  None,   // silent
  Verbose // progress lines on std::clog
};

// shared_ptr deleter that returns device memory to the queue's context (keeps the context alive).
template <class DT> struct Asycl_deleter {
  asycl::queue _q;
  Asycl_deleter(asycl::queue &q) : _q(q) {}
  void operator()(DT *ptr) { asycl::free(ptr, _q); }
};

namespace detail {
template <class T> inline std::shared_ptr<T[]> device_array(asycl::queue &q, std::size_t count) {
  return std::shared_ptr<T[]>(asycl::malloc_device<T>(count, q), Asycl_deleter<T>(q));
}
// cgb_csr handle over arrays the Matrix owns; holds the queue so the context outlives the handle.
struct csr_handle {
  asycl::queue q;
  cgb_csr *h = nullptr;
  explicit csr_handle(const asycl::queue &queue) : q(queue) {}
  ~csr_handle() { cgb_csr_destroy(h); }
  csr_handle(const csr_handle &) = delete;
  csr_handle &operator=(const csr_handle &) = delete;
};
} // namespace detail

template <class DT> class Matrix {
public:
  explicit Matrix(asycl::queue q) : _size(0), _NNZ(0), _N(0), _queue(q) {}
  explicit Matrix(asycl::queue q, const std::size_t size) : _size(size), _NNZ(0), _N(0), _queue(q) {}
  explicit Matrix(asycl::queue q, std::vector<DT> &data, std::vector<int> &cols, std::vector<int> &rows)
      : _size(rows.size()), _NNZ(data.size()), _N(rows.size() - 1), _queue(q) {
    init(data, cols, rows);
    _queue.wait();
  }

  auto data() { return _data; }
  auto data_ptr() { return _data.get(); }
  auto columns() { return _columns; }
  auto columns_ptr() { return _columns.get(); }
  auto rows() { return _rows; }
  auto rows_ptr() { return _rows.get(); }
  auto N() const { return _N; }
  auto NNZ() const { return _NNZ; }

  // Three device allocations + three H2D copies ordered on the queue.  As in the reference the
  // host vectors are only borrowed: the copies have left them when this call returns.
  auto init(std::vector<DT> &data, std::vector<int> &cols, std::vector<int> &rows) {
    if (_size == 0) _size = rows.size();
    _N = _size - 1;
    _NNZ = data.size();
    _csr.reset();
    _data = detail::device_array<DT>(_queue, _NNZ);
    _columns = detail::device_array<int>(_queue, _NNZ);
    _rows = detail::device_array<int>(_queue, _size);
    _queue.copy(data.data(), _data.get(), _NNZ);
    _queue.copy(cols.data(), _columns.get(), _NNZ);
    _queue.copy(rows.data(), _rows.get(), _size);
  }

  // --- additive (not in the reference) ---------------------------------------------------------
  // The library handle of this matrix: created on first use over the arrays above (no copy, no
  // ownership transfer), shared by copies of the Matrix, destroyed before the arrays.
  cgb_csr *csr() {
    static_assert(std::is_same<DT, double>::value, "the cgb200 kernels are fp64");
    if (!_rows) return nullptr;
    if (!_csr) {
      auto h = std::make_shared<detail::csr_handle>(_queue);
      asycl::detail::check(cgb_csr_from_device(_queue.native(), (int64_t)_N, (int64_t)_NNZ, _data.get(),
                                               _columns.get(), _rows.get(), 0, &h->h));
      _csr = h;
    }
    return _csr->h;
  }
  // Adopt a matrix generated on the device (cgb_gen_stencil): the Matrix takes the handle.
  static Matrix from_handle(asycl::queue q, cgb_csr *h) {
    Matrix m(q);
    m._N = (std::size_t)cgb_csr_n(h);
    m._NNZ = (std::size_t)cgb_csr_nnz(h);
    m._size = m._N + 1;
    auto keep = std::make_shared<detail::csr_handle>(q);
    keep->h = h;
    m._csr = keep;
    // aliasing pointers: lifetime tied to the handle, which owns the arrays
    m._data = std::shared_ptr<DT[]>(keep, cgb_csr_vals(h));
    m._columns = std::shared_ptr<int[]>(keep, cgb_csr_cols(h));
    m._rows = std::shared_ptr<int[]>(keep, cgb_csr_rowptr(h));
    return m;
  }

  // Assemble on the GPU from triplets (cgb_csr_from_triplets; flags: CGB_COO_*).  This is the bulk replacement
  // for the reference's element-by-element HiFlow3 bridge (src/CG.hpp:104-151): hand over (i, j, v) lists once.
  static Matrix from_triplets(asycl::queue q, const std::vector<int> &rows, const std::vector<int> &cols,
                              const std::vector<DT> &vals, long long n = -1, int flags = 0) {
    static_assert(std::is_same<DT, double>::value, "the cgb200 kernels are fp64");
    cgb_csr *h = nullptr;
    asycl::detail::check(cgb_csr_from_triplets(q.native(), (int64_t)n, (int64_t)vals.size(), rows.data(), cols.data(),
                                               vals.data(), flags, &h));
    return from_handle(q, h);
  }
  // Matrix Market file straight to the device, in the dialect of the reference's read_file
  // (test/mm_reader.cpp:154-171): parallel parse on the host, assembly on the GPU, no host CSR.
  static Matrix from_file(asycl::queue q, const std::string &path) {
    cgb_csr *h = nullptr;
    asycl::detail::check(cgb_csr_from_mm_file(q.native(), path.c_str(), &h));
    return from_handle(q, h);
  }

private:
  std::size_t _size;
  std::size_t _NNZ;
  std::size_t _N;
  asycl::queue _queue;
  std::shared_ptr<DT[]> _data;
  std::shared_ptr<int[]> _columns;
  std::shared_ptr<int[]> _rows;
  std::shared_ptr<detail::csr_handle> _csr; // declared last: released first
};

template <class DT> class Vector {
public:
  explicit Vector(asycl::queue q) : _q(q), _N(0) {}
  explicit Vector(asycl::queue q, const std::size_t N) : _q(q), _N(N) { init_empty(_N); }
  explicit Vector(asycl::queue q, std::vector<DT> &data) : _q(q), _N(0) {
    init(data);
    _q.wait();
  }

  // Fresh zero-filled storage of the vector's size (the size argument only matters the first time).
  asycl::event init_empty(std::size_t size = 0) {
    if (size != 0 && _N == 0) _N = size;
    assert(_N != 0);
    _ptr = detail::device_array<DT>(_q, _N);
    return _q.fill(_ptr.get(), static_cast<DT>(0), _N);
  }

  asycl::event init(std::vector<DT> &data) {
    _N = data.size(); // the reference leaves _N untouched here; N() is more useful when it is right
    _ptr = detail::device_array<DT>(_q, data.size());
    return _q.copy(data.data(), _ptr.get(), data.size());
  }

  auto data() { return _ptr; }
  auto ptr() { return _ptr.get(); }
  auto N() { return _N; }

private:
  asycl::queue _q;
  std::size_t _N;
  std::shared_ptr<DT[]> _ptr;
};

template <class DT> class Scalar {
public:
  explicit Scalar(asycl::queue q, DT value_ = static_cast<DT>(0)) : _q(q) { init(value_); }

  // One device element, set with a blocking copy (reference :216-224).
  auto init(DT value_) {
    value = std::shared_ptr<DT>(asycl::malloc_device<DT>(1, _q), Asycl_deleter<DT>(_q));
    _q.copy(&value_, value.get(), 1);
    _q.wait();
  }

  auto ptr() { return value.get(); }
  constexpr operator DT *() const { return value.get(); }
  DT *operator*() { return value.get(); }

  // additive: blocking read-back
  DT get() {
    DT v;
    _q.copy(value.get(), &v, 1).wait();
    return v;
  }

private:
  asycl::queue _q;
  std::shared_ptr<DT> value;
};

} // namespace CGSolver

#endif // CGB200_LINEAR_ALGEBRA_TYPES_HPP
