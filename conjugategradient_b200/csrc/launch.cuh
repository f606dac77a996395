// launch.cuh — host-side dispatch of the compile-time kernel parameters (lanes per row, row engine).
#pragma once
#include <type_traits>

#include "cg_kernels.cuh"
#include "matrix.cuh"

namespace cgb {

struct SpmvLaunch {
  int grid;
  int block;
  size_t smem;
  int lanes;
  bool tma;
};

inline SpmvLaunch spmv_launch_cfg(const cgb_csr *A) {
  SpmvLaunch c;
  c.tma = A->use_tma;
  c.lanes = c.tma ? A->lanes_tma : A->lanes;
  c.grid = c.tma ? A->grid_tma : A->grid_spmv;
  c.block = c.tma ? kTmaThreads : kBlock;
  c.smem = c.tma ? tma_smem_bytes(kTmaConsumers / c.lanes) : 0;
  return c;
}

inline CsrView csr_view(const cgb_csr *A) { return CsrView{A->n, A->nnz, A->rowptr, A->cols, A->vals}; }

// f(std::integral_constant<int, L>, std::bool_constant<TMA>)
template <class F> inline void dispatch_lanes(int lanes, bool tma, F &&f) {
#define CGB_CASE(LV)                                                  \
  if (tma) f(std::integral_constant<int, LV>{}, std::true_type{});    \
  else f(std::integral_constant<int, LV>{}, std::false_type{});       \
  break;
  switch (lanes) {
  case 1: CGB_CASE(1)
  case 2: CGB_CASE(2)
  case 4: CGB_CASE(4)
  case 8: CGB_CASE(8)
  case 16: CGB_CASE(16)
  default: CGB_CASE(32)
  }
#undef CGB_CASE
}

template <class K, class... Args>
inline void launch_kernel(K kernel, const SpmvLaunch &c, cudaStream_t s, Args... args) {
  if (c.smem > 48 * 1024)
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
  kernel<<<c.grid, c.block, c.smem, s>>>(args...);
}

} // namespace cgb
