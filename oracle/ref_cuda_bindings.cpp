// TEST / BASELINE INFRASTRUCTURE ONLY -- never linked into the product.
//
// Registers the reference's OWN CUDA kernels (compiled where they lie under /root/reference/csrc
// by oracle/build_ref_cuda.py, for sm_100) as torch ops in the namespaces `_C_refcuda` and
// `_C_refcuda_cache_ops`, so they coexist in one process with the product's `_C` registration.
// Prototypes come from the reference's csrc/ops.h and csrc/cache.h (included by -I, not copied);
// the schema strings are the ones of csrc/torch_bindings.cpp:24-48,129-137,174-243.
//
// Not included: cutlass_scaled_mm (needs the CUTLASS v3.5.0 tree that the reference's cmake
// fetches at build time; not on disk) and everything SURVEY.md marks out of scope.
#include <torch/library.h>

#include "cache.h"
#include "ops.h"

TORCH_LIBRARY(_C_refcuda, ops) {
  ops.def(
      "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache,"
      " int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size,"
      " int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, float kv_scale, int tp_rank,"
      " int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,"
      " int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2(Tensor! out, Tensor exp_sums, Tensor max_logits, Tensor tmp_out,"
      " Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale,"
      " Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len,"
      " Tensor? alibi_slopes, str kv_cache_dtype, float kv_scale, int tp_rank,"
      " int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size,"
      " int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);

  ops.def("silu_and_mul(Tensor! out, Tensor input) -> ()");
  ops.impl("silu_and_mul", torch::kCUDA, &silu_and_mul);
  ops.def("rms_norm(Tensor! out, Tensor input, Tensor weight, float epsilon) -> ()");
  ops.impl("rms_norm", torch::kCUDA, &rms_norm);
  ops.def("fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, float epsilon) -> ()");
  ops.impl("fused_add_rms_norm", torch::kCUDA, &fused_add_rms_norm);
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query, Tensor! key, int head_size,"
      " Tensor cos_sin_cache, bool is_neox) -> ()");
  ops.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);

  ops.def("gptq_marlin_24_gemm", &gptq_marlin_24_gemm);
  ops.def("gptq_marlin_gemm", &gptq_marlin_gemm);
  ops.def("gptq_marlin_repack", &gptq_marlin_repack);
  ops.def("awq_gemm", &awq_gemm);            // csrc/torch_bindings.cpp:117-118
  ops.def("awq_dequantize", &awq_dequantize);  // :121-122

  ops.def("static_scaled_int8_quant(Tensor! out, Tensor input, Tensor scale) -> ()");
  ops.impl("static_scaled_int8_quant", torch::kCUDA, &static_scaled_int8_quant);
  ops.def("dynamic_scaled_int8_quant(Tensor! out, Tensor input, Tensor! scale) -> ()");
  ops.impl("dynamic_scaled_int8_quant", torch::kCUDA, &dynamic_scaled_int8_quant);
  ops.def("static_scaled_fp8_quant(Tensor! out, Tensor input, Tensor scale) -> ()");
  ops.impl("static_scaled_fp8_quant", torch::kCUDA, &static_scaled_fp8_quant);
  ops.def("dynamic_scaled_fp8_quant(Tensor! out, Tensor input, Tensor! scale) -> ()");
  ops.impl("dynamic_scaled_fp8_quant", torch::kCUDA, &dynamic_scaled_fp8_quant);
}

TORCH_LIBRARY(_C_refcuda_cache_ops, cache_ops) {
  cache_ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  cache_ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  cache_ops.def("copy_blocks(Tensor[]! key_caches, Tensor[]! value_caches, Tensor block_mapping) -> ()");
  cache_ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  cache_ops.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype, float kv_scale) -> ()");
  cache_ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  cache_ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache,"
      " Tensor slot_mapping, str kv_cache_dtype) -> ()");
  cache_ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  cache_ops.def("convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  cache_ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
}
