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
  int eng; // row engine: kEngVec / kEngTma / kEngSell
};

inline bool csr_wants_sell(const cgb_csr *A) {
  return A->format == CGB_FMT_SELL || (A->format == CGB_FMT_AUTO && A->auto_sell);
}
// the column panels in use (0: the matrix is streamed as one piece)
inline size_t csr_panels(const cgb_csr *A) { return csr_wants_sell(A) ? A->panels.size() : 0; }

inline int csr_engine(const cgb_csr *A) {
  if (A->format == CGB_FMT_DICT && A->dict_ids) return kEngDict;
  if (csr_wants_sell(A) && (A->sell_vals || !A->panels.empty())) return kEngSell;
  return A->use_tma ? kEngTma : kEngVec;
}

inline SpmvLaunch spmv_launch_cfg(const cgb_csr *A) {
  SpmvLaunch c;
  c.eng = csr_engine(A);
  c.lanes = c.eng == kEngTma ? A->lanes_tma : ((c.eng == kEngSell || c.eng == kEngDict) ? 1 : A->lanes);
  c.grid = c.eng == kEngTma ? A->grid_tma : c.eng == kEngSell ? A->grid_sell : c.eng == kEngDict ? A->grid_dict : A->grid_spmv;
  c.block = c.eng == kEngTma ? kTmaThreads : kBlock;
  c.smem = c.eng == kEngTma ? tma_smem_bytes(kTmaConsumers / c.lanes) : c.eng == kEngDict ? dict_smem_bytes(A->dict_npat, A->dict_nent) : 0;
  return c;
}

inline CsrView csr_view(const cgb_csr *A) {
  return CsrView{A->n, A->nnz, A->rowptr, A->cols, A->vals, A->sell_slices, A->sell_off, A->sell_width, A->sell_cols, A->sell_vals, A->sell_perm,
                 A->sell_nlong, A->sell_long, A->sell_chunk, nullptr,
                 DictView{A->dict_ids, A->dict_ptr, A->dict_off, A->dict_val, A->dict_npat, A->dict_nent}};
}

// f(std::integral_constant<int, L>, std::integral_constant<int, ENG>)
template <class F> inline void dispatch_lanes(int lanes, int eng, F &&f) {
  if (eng == kEngSell) {
    f(std::integral_constant<int, 1>{}, std::integral_constant<int, kEngSell>{});
    return;
  }
  if (eng == kEngDict) {
    f(std::integral_constant<int, 1>{}, std::integral_constant<int, kEngDict>{});
    return;
  }
#define CGB_CASE(LV)                                                                                   \
  if (eng == kEngTma) f(std::integral_constant<int, LV>{}, std::integral_constant<int, kEngTma>{});    \
  else f(std::integral_constant<int, LV>{}, std::integral_constant<int, kEngVec>{});                   \
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

// ---- column panels ---------------------------------------------------------------------------
// A matrix with panels is A = [A_0 | ... | A_{P-1}] (sell.cu).  An SpMV-shaped pass becomes P launches:
// panels 0 .. P-2 accumulate their row sums into `acc` (plain spmv_kernel; acc = the pass's own output or
// a scratch vector), the last panel runs the pass's real kernel on csr_view_final(A, acc), whose engine adds
// acc[row] to the row sum before the epilogue sees it.  Without panels csr_view_final is csr_view.
inline CsrView csr_view_final(const cgb_csr *A, const double *acc) {
  const size_t P = csr_panels(A);
  if (P == 0) return csr_view(A);
  CsrView v = csr_view(A->panels[P - 1]);
  v.panel_acc = P > 1 ? acc : nullptr;
  return v;
}

inline void launch_panel_partials(const cgb_csr *A, const SpmvLaunch &c, cudaStream_t s, const double *x, double *acc,
                                  const CgState *gate) {
  for (size_t k = 0; k + 1 < csr_panels(A); k++) {
    CsrView v = csr_view(A->panels[k]);
    v.panel_acc = k > 0 ? acc : nullptr;
    launch_kernel(spmv_kernel<1, kEngSell>, c, s, v, x, acc, gate);
  }
}

} // namespace cgb
