// gemm.cu — exg_gemm: argument validation and dispatch between the tcgen05 and SIMT kernels.
#include "common.cuh"
#include <stdlib.h>

int exg_gemm_simt(const exg_gemm_params* p, void* stream);
int exg_gemm_tc(const exg_gemm_params* p, void* stream);
int exg_gemm_tc2(const exg_gemm_params* p, void* stream);
int exg_gemm_tc2_supported(const exg_gemm_params* p);

extern "C" int exg_gemm(const exg_gemm_params* p, void* stream) {
  EXG_CHECK_ARG(p && p->A && p->B && p->C, "gemm: null pointer");
  EXG_CHECK_ARG(p->M >= 0 && p->N >= 0 && p->K >= 0 && p->batch >= 1, "gemm: bad sizes M=%d N=%d K=%d batch=%d",
                p->M, p->N, p->K, p->batch);
  EXG_CHECK_ARG(exg_valid_dtype(p->in_dtype) && exg_valid_dtype(p->out_dtype), "gemm: bad dtype");
  EXG_CHECK_ARG(p->epilogue >= EXG_EPI_NONE && p->epilogue <= EXG_EPI_RELU_BWD, "gemm: bad epilogue %d", p->epilogue);
  if (p->epilogue == EXG_EPI_RESID)
    EXG_CHECK_ARG(p->aux && p->out_dtype == EXG_F32, "gemm: RESID needs fp32 aux and fp32 output");
  if (p->epilogue == EXG_EPI_QGELU_BWD || p->epilogue == EXG_EPI_RELU_BWD)
    EXG_CHECK_ARG(p->aux != nullptr, "gemm: *_BWD epilogue needs aux");
  if (p->M == 0 || p->N == 0) return 0;
  static const int force_simt = getenv("EXG_FORCE_SIMT") ? atoi(getenv("EXG_FORCE_SIMT")) : 0;
  int impl = p->impl;
  if (impl == EXG_GEMM_AUTO) {
    impl = (!force_simt && p->K > 0 && exg_gemm_tc_supported(p)) ? EXG_GEMM_TC : EXG_GEMM_SIMT;
  } else if (impl == EXG_GEMM_TC) {
    if (!exg_gemm_tc_supported(p) || p->K == 0) {
      exg_set_error("gemm: tcgen05 path does not support these operands (dtype/alignment)");
      return EXG_EUNSUPPORTED;
    }
  }
  static const int no_tc2 = getenv("EXG_NO_TC2") ? atoi(getenv("EXG_NO_TC2")) : 0;
  if (impl == EXG_GEMM_TC2) {
    if (!exg_gemm_tc2_supported(p)) {
      exg_set_error("gemm: persistent tcgen05 kernel does not support these operands / epilogue");
      return EXG_EUNSUPPORTED;
    }
  } else if (impl == EXG_GEMM_TC && !no_tc2 && p->impl == EXG_GEMM_AUTO && exg_gemm_tc2_supported(p)) {
    impl = EXG_GEMM_TC2;     // AUTO prefers the persistent TMA-epilogue kernel
  }
  ExgProfScope prof(impl == EXG_GEMM_SIMT ? EXG_PC_GEMM_SIMT : EXG_PC_GEMM_TC,
                    2.0 * p->M * p->N * p->K * p->batch, stream);
  if (impl == EXG_GEMM_TC2) return exg_gemm_tc2(p, stream);
  return impl == EXG_GEMM_TC ? exg_gemm_tc(p, stream) : exg_gemm_simt(p, stream);
}
