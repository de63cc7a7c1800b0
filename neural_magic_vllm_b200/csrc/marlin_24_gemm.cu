// gptq_marlin_24_gemm placeholder (kernel lands in the next commit).
#include "common.cuh"
extern "C" int b200_gptq_marlin_24_gemm(const b200_gptq_marlin_24_gemm_params* p, void* stream) {
  (void)p;
  (void)stream;
  b200::set_error("gptq_marlin_24_gemm: not implemented yet");
  return B200_ENOTSUP;
}
