// ops.cu — the per-op vector kernels behind VectorOperations (reference
// src/VectorOperations.hpp): dot (accumulate-into), norm, saxpby, sapbx, sambx, fill, scalars.
// The element-wise ops use separate multiply and add roundings (no FMA contraction) so they are
// bit-identical to the reference's expressions compiled without contraction.
#include "common.cuh"

using namespace cgb;

namespace {

__device__ __forceinline__ bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// MODE 0: a*x + b*y   (saxpby, :361-363)   MODE 1: x + b*y (sapbx, :422-424)   MODE 2: x - b*y (sambx, :391-393)
template <int MODE>
__device__ __forceinline__ double combine(double a, double b, double x, double y) {
  if (MODE == 0) return __dadd_rn(__dmul_rn(a, x), __dmul_rn(b, y));
  if (MODE == 1) return __dadd_rn(x, __dmul_rn(b, y));
  return __dsub_rn(x, __dmul_rn(b, y));
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) axpby_kernel(int64_t n, const double *x, const double *y, const double *pa,
                                                        const double *pb, double *out) {
  const double a = (MODE == 0) ? *pa : 1.0, b = *pb; // scalars are device pointers read at run time
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  if (aligned16(x) && aligned16(y) && aligned16(out)) {
    const int64_t n2 = n >> 1;
    const double2 *x2 = reinterpret_cast<const double2 *>(x);
    const double2 *y2 = reinterpret_cast<const double2 *>(y);
    double2 *o2 = reinterpret_cast<double2 *>(out);
    for (int64_t i = tid; i < n2; i += nthr) {
      const double2 xv = x2[i], yv = y2[i];
      double2 ov;
      ov.x = combine<MODE>(a, b, xv.x, yv.x);
      ov.y = combine<MODE>(a, b, xv.y, yv.y);
      o2[i] = ov;
    }
    if ((n & 1) && tid == 0) out[n - 1] = combine<MODE>(a, b, x[n - 1], y[n - 1]);
  } else {
    for (int64_t i = tid; i < n; i += nthr) out[i] = combine<MODE>(a, b, x[i], y[i]);
  }
}

// *result += sum x[i]*y[i]; per-CTA partials, last CTA sums them in a fixed order.
__global__ void __launch_bounds__(kBlock) dot_acc_kernel(int64_t n, const double *__restrict__ x,
                                                          const double *__restrict__ y, double *result,
                                                          double *partials, unsigned int *ticket) {
  __shared__ double red[32];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  if (aligned16(x) && aligned16(y)) {
    const int64_t n2 = n >> 1;
    const double2 *x2 = reinterpret_cast<const double2 *>(x);
    const double2 *y2 = reinterpret_cast<const double2 *>(y);
    for (int64_t i = tid; i < n2; i += nthr) {
      const double2 xv = x2[i], yv = y2[i];
      acc = fma(xv.x, yv.x, acc);
      acc = fma(xv.y, yv.y, acc);
    }
    if ((n & 1) && tid == 0) acc = fma(x[n - 1], y[n - 1], acc);
  } else {
    for (int64_t i = tid; i < n; i += nthr) acc = fma(x[i], y[i], acc);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (publish_and_elect(ticket, gridDim.x)) {
    const double s = reduce_partials(partials, gridDim.x, red);
    if (threadIdx.x == 0) *result = *result + s; // accumulate: no initialize_to_identity (:300)
  }
}

__global__ void fill_kernel(double *p, double v, int64_t n) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) p[i] = v;
}

int vec_grid(cgb_ctx *ctx, int64_t n) {
  int64_t need = (n / 2 + kBlock - 1) / kBlock;
  int64_t cap = (int64_t)ctx->sm_count * 8;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

} // namespace

extern "C" {

int cgb_dot_acc(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, double *d_result) {
  CGB_REQUIRE(ctx && d_result && (n == 0 || (d_x && d_y)) && n >= 0, "cgb_dot_acc: bad argument");
  if (n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  CGB_TRY(ctx_scratch(ctx));
  dot_acc_kernel<<<vec_grid(ctx, n), kBlock, 0, ctx->stream>>>(n, d_x, d_y, d_result, ctx->partials, ctx->ticket);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int cgb_norm2_acc(cgb_ctx *ctx, int64_t n, const double *d_x, double *d_result) {
  return cgb_dot_acc(ctx, n, d_x, d_x, d_result);
}

int cgb_axpby(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, const double *d_a, const double *d_b,
              double *d_out) {
  CGB_REQUIRE(ctx && d_a && d_b && (n == 0 || (d_x && d_y && d_out)) && n >= 0, "cgb_axpby: bad argument");
  if (n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  axpby_kernel<0><<<vec_grid(ctx, n), kBlock, 0, ctx->stream>>>(n, d_x, d_y, d_a, d_b, d_out);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int cgb_xpby(cgb_ctx *ctx, int64_t n, const double *d_x, const double *d_y, const double *d_b, int sign, double *d_out) {
  CGB_REQUIRE(ctx && d_b && (n == 0 || (d_x && d_y && d_out)) && n >= 0, "cgb_xpby: bad argument");
  CGB_REQUIRE(sign == 1 || sign == -1, "cgb_xpby: sign must be +1 or -1");
  if (n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  if (sign > 0)
    axpby_kernel<1><<<vec_grid(ctx, n), kBlock, 0, ctx->stream>>>(n, d_x, d_y, nullptr, d_b, d_out);
  else
    axpby_kernel<2><<<vec_grid(ctx, n), kBlock, 0, ctx->stream>>>(n, d_x, d_y, nullptr, d_b, d_out);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int cgb_fill_f64(cgb_ctx *ctx, double *d_ptr, double value, size_t n) {
  CGB_REQUIRE(ctx && (n == 0 || d_ptr), "cgb_fill_f64: bad argument");
  if (n == 0) return CGB_OK;
  ScopedDevice sd(ctx->device);
  if (value == 0.0) {
    CGB_CUDA(cudaMemsetAsync(d_ptr, 0, n * sizeof(double), ctx->stream));
    return CGB_OK;
  }
  fill_kernel<<<vec_grid(ctx, (int64_t)n), kBlock, 0, ctx->stream>>>(d_ptr, value, (int64_t)n);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}

int cgb_scalar_set(cgb_ctx *ctx, double *d_scalar, double value) {
  CGB_REQUIRE(ctx && d_scalar, "cgb_scalar_set: NULL argument");
  ScopedDevice sd(ctx->device);
  // blocking, like Scalar::init (src/LinearAlgebraTypes.hpp:216-224)
  CGB_CUDA(cudaMemcpyAsync(d_scalar, &value, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

int cgb_scalar_get(cgb_ctx *ctx, const double *d_scalar, double *h_value) {
  CGB_REQUIRE(ctx && d_scalar && h_value, "cgb_scalar_get: NULL argument");
  ScopedDevice sd(ctx->device);
  CGB_CUDA(cudaMemcpyAsync(h_value, d_scalar, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CGB_CUDA(cudaStreamSynchronize(ctx->stream));
  return CGB_OK;
}

} // extern "C"
