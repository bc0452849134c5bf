// tcgen05 / TMA / TMEM GEMM path (MDT_GEMM_TF32, MDT_GEMM_TF32X3) -- placeholder
// until the kernels land; every entry fails loudly.
#pragma once
#include "common.cuh"

namespace mdt {
struct Tc05Plan;
struct Tc05Buffers {
  float *params, *grads;
  Layout L;
  int D, H, Z, Bmax;
  float *x_stage, *h1, *ml, *eps, *sd, *z, *h3, *recon, *da4, *da3, *dz, *dml_kl, *dml, *da1;
  StepState* st;
  int sm_count, max_threads_sm;
  bool split3;
};
inline int tc05_unsupported() { return set_error(MDT_ERR_UNSUPPORTED, "tcgen05 GEMM path is not built yet"); }
inline int tc05_create(const Tc05Buffers*, cudaStream_t, Tc05Plan**) { return tc05_unsupported(); }
inline void tc05_destroy(Tc05Plan*) {}
inline int tc05_forward(Tc05Plan*, cudaStream_t, const float*, int, const float*, bool) { return tc05_unsupported(); }
inline int tc05_backward(Tc05Plan*, cudaStream_t, const float*, int) { return tc05_unsupported(); }
inline int tc05_params_changed(Tc05Plan*, cudaStream_t) { return tc05_unsupported(); }
inline int tc05_after_external_da4(Tc05Plan*, cudaStream_t, int) { return tc05_unsupported(); }
inline int tc05_linear(const float*, const float*, const float*, int, int, int, int, bool, float*, cudaStream_t) {
  return tc05_unsupported();
}
}  // namespace mdt
