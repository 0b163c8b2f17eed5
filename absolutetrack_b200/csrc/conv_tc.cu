// tcgen05 / TMEM / TMA bf16 implicit-GEMM encoder (AT_PREC_BF16).  Placeholder until the
// tensor-core path lands: fails loudly, never falls back.
#include "conv.cuh"

namespace at {
int64_t encoder_workspace_bytes_bf16(int c) { return (int64_t)c * 0 + 256; }
int encoder_forward_bf16(const float*, const void*, const void*, int, int, float*, void*, int64_t, cudaStream_t) {
  set_error("AT_PREC_BF16 encoder is not built into this library");
  return AT_ERR_UNSUPPORTED;
}
}  // namespace at

extern "C" int64_t at_model_bf16_blob_bytes(void) { return 256; }
extern "C" int at_model_prepack_bf16(const float*, void*, void*) {
  at::set_error("AT_PREC_BF16 encoder is not built into this library");
  return AT_ERR_UNSUPPORTED;
}
