// sell.cu — SELL-C-sigma shadow format (placeholder until the conversion lands).
#include "matrix.cuh"
extern "C" int cgb_csr_build_sell(cgb_csr *A, int sigma) {
  (void)A; (void)sigma;
  return cgb::fail(CGB_ERR_INVALID, "cgb_csr_build_sell: not built yet");
}
