/*
 * b200_vllm.h -- C ABI of libb200vllm.so: the B200 (sm_100a) kernels behind nm-vllm's
 * custom-op surface (vllm/_custom_ops.py -> torch.ops._C / _C_cache_ops).
 *
 * This is the drop-in boundary.  Every entry point takes a POD parameter block holding raw
 * device pointers, sizes and element strides -- no torch types -- plus the CUDA stream to
 * launch on (a cudaStream_t passed as void*).  Entry points return 0 on success or a
 * negative B200_E* code; b200_last_error() returns the message for the calling thread.
 * No entry point synchronises, allocates device memory, or reads device data on the host,
 * so all of them are CUDA-graph capturable (reference: worker/model_runner.py:910-1060).
 *
 * Each function cites the reference interface it replaces (paths relative to the reference
 * checkout).  The torch-library registration that binds these to the reference's op schemas
 * is neural_magic_vllm_b200/csrc/torch_bindings.cpp (replaces csrc/torch_bindings.cpp).
 */
#ifndef B200_VLLM_H_
#define B200_VLLM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_VLLM_ABI_VERSION 2

/* every entry point is exported with default visibility (the library is built -fvisibility=hidden) */
#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

/* error codes */
#define B200_OK 0
#define B200_EINVAL (-1)   /* bad argument (shape / stride / alignment / enum) */
#define B200_ENOTSUP (-2)  /* combination the reference also rejects (head size, block size...) */
#define B200_ECUDA (-3)    /* CUDA runtime error at launch */

/* element types of activations / outputs */
typedef enum { B200_F16 = 0, B200_BF16 = 1, B200_F32 = 2 } b200_dtype;

/* kv_cache_dtype string of the reference (csrc/quantization/fp8/nvidia/quant_utils.cuh:516-565):
 * "auto" -> B200_KV_AUTO, "fp8" / "fp8_e4m3" -> B200_KV_FP8_E4M3, "fp8_e5m2" -> B200_KV_FP8_E5M2 */
typedef enum { B200_KV_AUTO = 0, B200_KV_FP8_E4M3 = 1, B200_KV_FP8_E5M2 = 2 } b200_kv_dtype;

/* quantized element types for the W8A8 GEMM */
typedef enum { B200_Q_INT8 = 0, B200_Q_FP8_E4M3 = 1 } b200_qtype;

B200_API int b200_abi_version(void);
B200_API const char* b200_last_error(void);
/* parses the reference's kv_cache_dtype string; returns a b200_kv_dtype or B200_EINVAL */
B200_API int b200_parse_kv_cache_dtype(const char* s);

/* ------------------------------------------------------------------------------------------
 * Paged attention (decode).  Replaces paged_attention_v1 / paged_attention_v2
 * (csrc/attention/attention_kernels.cu:805-826, 966-990; schemas csrc/torch_bindings.cpp:24-48).
 *
 * Layouts (elements of the cache type; x = 16 / sizeof(cache element)):
 *   query        [num_seqs, num_heads, head_size], row stride q_stride (may exceed H*D)
 *   key_cache    [num_blocks, num_kv_heads, head_size/x, block_size, x]
 *   value_cache  [num_blocks, num_kv_heads, head_size, block_size]
 *   block_tables [num_seqs, max_num_blocks_per_seq] int32,  seq_lens [num_seqs] int32
 *   out          [num_seqs, num_heads, head_size] contiguous
 *   v2 scratch (caller-owned): tmp_out [num_seqs, num_heads, max_num_partitions, head_size],
 *                exp_sums / max_logits [num_seqs, num_heads, max_num_partitions] fp32,
 *                max_num_partitions = ceil(max_seq_len / 512)
 * ------------------------------------------------------------------------------------------ */
#define B200_PARTITION_SIZE 512

typedef struct {
  void* out;
  const void* query;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes; /* [num_heads] or NULL */
  /* v2 only (NULL for v1) */
  float* exp_sums;
  float* max_logits;
  void* tmp_out;
  int32_t num_seqs, num_heads, num_kv_heads, head_size, block_size;
  int32_t max_num_blocks_per_seq, max_seq_len;
  int64_t q_stride, kv_block_stride, kv_head_stride; /* elements */
  float scale, kv_scale;
  int32_t dtype;    /* b200_dtype of query/out */
  int32_t kv_dtype; /* b200_kv_dtype */
  /* block-sparse extension (attention_kernels.cu:209-256); vert_stride <= 1 disables it */
  int32_t tp_rank, blocksparse_local_blocks, blocksparse_vert_stride;
  int32_t blocksparse_block_size, blocksparse_head_sliding_step;
} b200_paged_attention_params;

B200_API int b200_paged_attention_v1(const b200_paged_attention_params* p, void* stream);
B200_API int b200_paged_attention_v2(const b200_paged_attention_params* p, void* stream);

/* paged_attention_v2 that allocates nothing and needs no caller scratch.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs,
 * outputs and arithmetic as b200_paged_attention_v2; p->exp_sums / max_logits / tmp_out are
 * ignored and the partition scratch lives in a device array built into the library
 * (b200_paged_attention_v2_static_scratch_max_bytes() = 16 MiB, held by every GPU that loads the
 * library until exit).  A shape whose scratch (b200_paged_attention_v2_static_scratch_bytes)
 * exceeds that returns B200_EINVAL with a message.  All calls on a device share the array: calls
 * that overlap on the GPU (different streams without ordering) race; calls ordered on one stream
 * are safe.  Nothing in the array has to be zero on entry. */
B200_API int b200_paged_attention_v2_static(const b200_paged_attention_params* p, void* stream);
B200_API int64_t b200_paged_attention_v2_static_scratch_max_bytes(void);
B200_API int64_t b200_paged_attention_v2_static_scratch_bytes(int32_t num_seqs, int32_t num_heads,
                                                              int32_t head_size, int32_t max_seq_len,
                                                              int32_t dtype);

/* ------------------------------------------------------------------------------------------
 * Variable-length attention for prefill.  Stands in for the library call
 * of the reference's prefill path: flash_attn_varlen_func (vllm/attention/backends/flash_attn.py:
 * 320-346) / xformers (backends/xformers.py:340-424); semantics as backends/torch_sdpa.py:203-218.
 *   q [total_q, num_heads, head_size], k / v [total_k, num_kv_heads, head_size] (token strides in
 *   elements: q may be a slice of a fused qkv row), out [total_q, num_heads, head_size];
 *   cu_seqlens_q / cu_seqlens_k [num_seqs + 1] int32 prefix sums; causal mask aligned bottom-right.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  void* out;
  const void* q;
  const void* k;
  const void* v;
  const int32_t* cu_seqlens_q;
  const int32_t* cu_seqlens_k;
  int32_t num_seqs, num_heads, num_kv_heads, head_size;
  int32_t total_q, total_k, max_seqlen_q, max_seqlen_k;
  int64_t q_stride, k_stride, v_stride, out_stride; /* elements between consecutive tokens */
  float softmax_scale;
  int32_t causal;
  int32_t dtype; /* B200_F16 / B200_BF16 */
} b200_varlen_attention_params;

B200_API int b200_varlen_attention_fwd(const b200_varlen_attention_params* p, void* stream);


/* ------------------------------------------------------------------------------------------
 * Cache ops.  Replace csrc/cache_kernels.cu (schemas csrc/torch_bindings.cpp:207-244).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* key;   /* [num_tokens, num_heads, head_size], row stride key_stride */
  const void* value; /* same, row stride value_stride */
  void* key_cache;   /* [num_blocks, num_heads, head_size/x, block_size, x] */
  void* value_cache; /* [num_blocks, num_heads, head_size, block_size] */
  const int64_t* slot_mapping; /* [num_tokens]; < 0 = padding token, skipped */
  int32_t num_tokens, num_heads, head_size, block_size, x;
  int64_t key_stride, value_stride;
  float kv_scale;
  int32_t dtype;    /* b200_dtype of key/value */
  int32_t kv_dtype; /* b200_kv_dtype of the caches */
} b200_reshape_and_cache_params;

/* reshape_and_cache (cache_kernels.cu:253-278) */
B200_API int b200_reshape_and_cache(const b200_reshape_and_cache_params* p, void* stream);

typedef struct {
  const void* key;
  const void* value;
  void* k_cache; /* [num_blocks, block_size, num_heads, head_size] */
  void* v_cache;
  const int64_t* slot_mapping;
  int32_t num_tokens, num_heads, head_size, block_size;
  int64_t key_stride, value_stride, block_stride;
  int32_t dtype;
} b200_reshape_and_cache_flash_params;

/* reshape_and_cache_flash (cache_kernels.cu:280-314); kv dtype "auto" only, as the reference */
B200_API int b200_reshape_and_cache_flash(const b200_reshape_and_cache_flash_params* p, void* stream);

/* copy_blocks (cache_kernels.cu:101-148).  key_cache_ptrs / value_cache_ptrs are DEVICE arrays of
 * num_layers device pointers; block_mapping is a DEVICE [num_pairs, 2] int64 array (src, dst).
 * bytes_per_block = numel_per_block * element size; must be a multiple of 16. */
B200_API int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs,
                     const int64_t* block_mapping, int32_t num_layers, int32_t num_pairs,
                     int64_t bytes_per_block, void* stream);

/* swap_blocks (cache_kernels.cu:24-63).  block_mapping is a HOST [num_pairs, 2] int64 array.
 * kind: 0 = device->device, 1 = device->host, 2 = host->device. One async copy per pair. */
B200_API int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host,
                     int32_t num_pairs, int64_t bytes_per_block, int32_t kind, void* stream);

/* convert_fp8 (cache_kernels.cu:339-389): exactly one of src/dst is uint8 (fp8). to_fp8 != 0
 * quantises src (dtype) -> dst (fp8) as cvt_satfinite(x / scale); otherwise dequantises
 * src (fp8) -> dst (dtype) as float(half(fp8)) * scale (quant_utils.cuh:295-301, 455-463). */
B200_API int b200_convert_fp8(void* dst, const void* src, int64_t numel, float scale, int32_t dtype,
                     int32_t kv_dtype, int32_t to_fp8, void* stream);

/* _C_cuda_utils (csrc/cuda_utils_kernels.cu) */
B200_API int b200_get_device_attribute(int32_t attribute, int32_t device_id, int32_t* value);
B200_API int b200_get_max_shared_memory_per_block_device_attribute(int32_t device_id, int32_t* value);

/* ------------------------------------------------------------------------------------------
 * Activation quantisation feeding the W8A8 GEMM.
 * Replaces csrc/quantization/compressed_tensors/int8_quant_kernels.cu:70-115 and
 * csrc/quantization/fp8/common.cu:129-165.
 * ------------------------------------------------------------------------------------------ */
/* out int8 [num_tokens, hidden]; dynamic != 0: scales [num_tokens] written (absmax/127),
 * else *scales (device scalar) read. */
B200_API int b200_scaled_int8_quant(int8_t* out, const void* input, float* scales, int32_t num_tokens,
                           int32_t hidden, int32_t dtype, int32_t dynamic, void* stream);
/* out e4m3 [numel]; per-tensor scale (device scalar). dynamic != 0: *scale must be 0 on entry
 * and receives absmax/448 (as fp8/common.cu:78-105), two launches. */
B200_API int b200_scaled_fp8_quant(uint8_t* out, const void* input, float* scale, int64_t numel,
                          int32_t dtype, int32_t dynamic, void* stream);

/* ------------------------------------------------------------------------------------------
 * W8A8 GEMM with scale epilogue.  Replaces cutlass_scaled_mm
 * (csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:48-100, scaled_mm_c3x.cu:539-556).
 *   out[M,N] = a_scales[M or 1] * (b_scales[N or 1] * (a[M,K] . b[K,N])) (+ bias[N])
 *   a row-major (lda), b column-major (ldb = stride between columns), out row-major (ldc).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  void* out;
  const void* a;
  const void* b;
  const float* a_scales;
  const float* b_scales;
  const void* bias; /* NULL or [N] of out_dtype */
  int32_t m, n, k;
  int64_t lda, ldb, ldc;
  int32_t a_scales_numel, b_scales_numel; /* 1 = per-tensor, m / n = per-token / per-channel */
  int32_t qtype;     /* b200_qtype */
  int32_t out_dtype; /* B200_F16 / B200_BF16 */
  /* optional split-K scratch (device, 16-byte aligned) used only when m*n tiles cannot fill the
   * GPU (weight-streaming decode); NULL / too small => no split-K. Contents are don't-care. */
  void* workspace;
  int64_t workspace_bytes;
} b200_scaled_mm_params;

B200_API int b200_scaled_mm(const b200_scaled_mm_params* p, void* stream);
/* bytes of workspace that let b200_scaled_mm fill the GPU for this shape (0 = none needed) */
B200_API int64_t b200_scaled_mm_workspace_bytes(int32_t m, int32_t n, int32_t k);
/* cutlass_scaled_mm_supports_fp8 (scaled_mm_entry.cu:32-46): 1 for capability >= 89 */
B200_API int b200_scaled_mm_supports_fp8(int32_t cuda_device_capability);

/* ------------------------------------------------------------------------------------------
 * GPTQ-Marlin W4A16 / W8A16.  Replaces gptq_marlin_repack
 * (csrc/quantization/gptq_marlin/gptq_marlin_repack.cu:264-346) and gptq_marlin_gemm
 * (csrc/quantization/gptq_marlin/gptq_marlin.cu:1726-1866).
 * ------------------------------------------------------------------------------------------ */
/* b_q_weight [size_k/pack, size_n] int32 (GPTQ, packed along K) -> out [size_k/16, size_n*16/pack]
 * int32 (Marlin tiles); perm [size_k] int32 or NULL (act-order). pack = 32/num_bits. */
B200_API int b200_gptq_marlin_repack(const uint32_t* b_q_weight, const int32_t* perm, uint32_t* out,
                            int32_t size_k, int32_t size_n, int32_t num_bits, void* stream);

typedef struct {
  void* c;       /* [size_m, size_n] of dtype, contiguous */
  const void* a; /* [size_m, size_k] of dtype, row stride lda */
  const uint32_t* b_q_weight; /* Marlin-repacked [size_k/16, size_n*16/pack] */
  const void* b_scales;       /* [num_groups, size_n] of dtype, marlin_permute_scales order */
  const int32_t* g_idx;       /* [size_k] or NULL (no act-order) */
  const int32_t* perm;        /* [size_k] or NULL */
  void* a_tmp;                /* [size_m, size_k] scratch, required iff perm != NULL */
  int32_t* workspace;         /* caller's lock buffer; left zeroed (never written here) */
  int32_t size_m, size_n, size_k;
  int64_t lda;
  int32_t num_groups; /* b_scales rows; group_size = size_k / num_groups, or -1 semantics if 1 */
  int32_t num_bits;   /* 4 or 8 */
  int32_t is_k_full;
  int32_t dtype;      /* B200_F16 / B200_BF16 */
  /* optional split-K scratch (fp32 partials) for the weight-streaming regime; NULL => no split-K */
  void* splitk_ws;
  int64_t splitk_ws_bytes;
  int32_t weight_format; /* 0 = integers with zero point 2^(bits-1) (GPTQ); 1 = fp8 e4m3 (num_bits 8) */
} b200_gptq_marlin_gemm_params;

B200_API int b200_gptq_marlin_gemm(const b200_gptq_marlin_gemm_params* p, void* stream);
B200_API int64_t b200_gptq_marlin_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k);
/* extension (no single reference op): the row-parallel linear and the
 * fused_add_rms_norm behind it (csrc/layernorm_kernels.cu:199-249) as one call:
 *   residual [size_m, size_n] += gemm(a, w);  c = rms_norm(residual) * norm_weight
 * Where the GEMM splits K, the fp32 partial sums are reduced inside the norm kernel (one kernel
 * less per layer); otherwise the two operators run back to back.  Same bits as the unfused pair. */
B200_API int b200_gptq_marlin_gemm_add_rms_norm(const b200_gptq_marlin_gemm_params* p, void* residual,
                                                const void* norm_weight, float epsilon, void* stream);
/* test hook: which W4A16/W8A16 decode kernel gptq_marlin_gemm takes at M <= 64 (0 heuristic,
 * 1 first-generation split-K kernel, 2 stream-K kernel wherever eligible, -1 environment) */
B200_API void b200_debug_set_marlin_decode_kernel(int32_t choice);

/* debug only: cycle accounting of the stream-K decode kernel; returns the number of uint64 copied
 * (0 unless built with -DB200_STREAM_PROF) */
B200_API int b200_debug_stream_prof(unsigned long long* dst, int32_t n);
/* debug only: cycle totals of the prefill attention kernel's MMA issuer [0..7] and of one softmax
 * warp of query tile A [8..15] / B [16..23]; 0 unless built with -DB200_FA_PROF */
B200_API int b200_debug_fa_prof(unsigned long long* dst, int32_t n, int32_t reset);

/* ------------------------------------------------------------------------------------------
 * AWQ.  Replaces awq_gemm / awq_dequantize
 * (csrc/quantization/awq/gemm_kernels.cu:437-549; ops csrc/torch_bindings.cpp:117-122).  fp16.
 *   qweight [size_k, size_n / 8] int32, qzeros [size_k / group_size, size_n / 8] int32 (nibble j
 *   of a word = channel 8 c + {0,2,4,6,1,3,5,7}[j]), scales [size_k / group_size, size_n] fp16;
 *   w = (q - z) * s rounded to fp16 once.  size_n % 64 == 0, group_size % 32 == 0.
 * awq_gemm: c [size_m, size_n] = a [size_m, size_k] (row stride lda elements) x w.  splitk_ws
 * (optional, b200_awq_gemm_workspace_bytes) lets small size_m fill the GPU by splitting K; the
 * reference's split_k_iters argument only sizes ITS scratch and has no meaning here.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  void* c;
  const void* a;
  const void* qweight;
  const void* scales;
  const void* qzeros;
  int32_t size_m, size_n, size_k, group_size;
  int64_t lda;
  void* splitk_ws;
  int64_t splitk_ws_bytes;
} b200_awq_gemm_params;

B200_API int b200_awq_gemm(const b200_awq_gemm_params* p, void* stream);
B200_API int64_t b200_awq_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k);
B200_API int b200_awq_dequantize(void* out, const void* qweight, const void* scales, const void* qzeros,
                                 int32_t size_k, int32_t size_n, int32_t group_size, void* stream);

/* fp8_marlin_gemm (csrc/quantization/fp8/fp8_marlin.cu:1212-1308): fp8 e4m3 weights repacked with
 * gptq_marlin_repack(num_bits = 8), b_scales [1, size_n] in marlin_permute_scales order.  Takes the
 * same parameter block (num_bits = 8, num_groups = 1, no act-order). */
B200_API int b200_fp8_marlin_gemm(const b200_gptq_marlin_gemm_params* p, void* stream);

/* test hook: out [size_k, size_n] (dtype) = dequantised weights, produced by the same unpack /
 * scale device code the GEMM uses (bit-exact against the reference's w_ref,
 * quant_utils.py:73). No act-order. */
B200_API int b200_marlin_dequant(void* out, const uint32_t* b_q_weight, const void* b_scales,
                                 int32_t size_k, int32_t size_n, int32_t num_groups,
                                 int32_t num_bits, int32_t dtype, void* stream);

/* 2:4-sparse Marlin.  Replaces gptq_marlin_24_gemm
 * (csrc/quantization/marlin/sparse/marlin_24_cuda_kernel.cu:1015-1125). fp16 only. */
typedef struct {
  void* c;       /* [size_m, size_n] fp16 */
  const void* a; /* [size_m, size_k] fp16, contiguous */
  const uint32_t* b_q_weight; /* [size_k/2/16, size_n*16/pack] int32 */
  const uint16_t* b_meta;     /* [size_k/2/... ] int16 metadata, reference layout */
  const void* b_scales;       /* [num_groups, size_n] fp16 */
  int32_t* workspace;
  int32_t size_m, size_n, size_k;
  int32_t num_groups;
  int32_t num_bits;
  void* splitk_ws; /* optional fp32 split-K scratch, as in b200_gptq_marlin_gemm_params */
  int64_t splitk_ws_bytes;
} b200_gptq_marlin_24_gemm_params;

B200_API int b200_gptq_marlin_24_gemm(const b200_gptq_marlin_24_gemm_params* p, void* stream);
/* bytes of splitk_ws that let b200_gptq_marlin_24_gemm fill the GPU for this shape (0 = none needed) */
B200_API int64_t b200_gptq_marlin_24_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k);

/* ------------------------------------------------------------------------------------------
 * Decode-side glue between the linears and attention.
 * Rounding follows the reference kernels operator by operator (each scalar_t operator is an
 * fp32 operation rounded to the element type).
 * ------------------------------------------------------------------------------------------ */

/* rms_norm (csrc/layernorm_kernels.cu:21-45, 291-310): out = T(x * rsqrt(mean(x^2) + eps)) * weight.
 * input / out [num_tokens, hidden] contiguous, weight [hidden]. */
B200_API int b200_rms_norm(void* out, const void* input, const void* weight, float epsilon,
                           int32_t num_tokens, int32_t hidden, int32_t dtype, void* stream);

/* (fused_add_)rms_norm + activation quantisation in one pass (extension):
 * residual != NULL: residual += input first (rounded, stored back), as fused_add_rms_norm.
 * quant_mode 0: int8 per-token dynamic, scales [num_tokens] written; 1: int8 static, *scales read;
 * 2: fp8 e4m3 static, *scales read.  out [num_tokens, hidden] bytes.  Results are bit-identical to
 * (fused_add_)rms_norm followed by scaled_int8_quant / scaled_fp8_quant. */
B200_API int b200_rms_norm_quant(void* out, float* scales, const void* input, void* residual, const void* weight,
                        float epsilon, int32_t num_tokens, int32_t hidden, int32_t dtype, int32_t quant_mode,
                        void* stream);

/* fused_add_rms_norm (csrc/layernorm_kernels.cu:199-287, 326-352), in place:
 * residual = input + residual; input = rms_norm(residual) * weight. */
B200_API int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, float epsilon,
                                     int32_t num_tokens, int32_t hidden, int32_t dtype, void* stream);

/* rotary_embedding / batched_rotary_embedding (csrc/pos_encoding_kernels.cu:127-203), in place.
 * query [num_tokens, num_heads*head_size] with row stride query_stride (elements), key likewise;
 * cos_sin_cache [max_position, rot_dim] = (cos[rot_dim/2], sin[rot_dim/2]) per position;
 * cos_sin_cache_offsets: NULL (rotary_embedding) or [num_tokens] int64 added to the position
 * (batched_rotary_embedding). is_neox: rotate halves (NeoX) or interleaved pairs (GPT-J). */
B200_API int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                                   const void* cos_sin_cache, const int64_t* cos_sin_cache_offsets,
                                   int32_t num_tokens, int32_t rot_dim, int64_t query_stride,
                                   int64_t key_stride, int32_t num_heads, int32_t num_kv_heads,
                                   int32_t head_size, int32_t is_neox, int32_t dtype, void* stream);

/* activation ids of b200_act_and_mul / b200_activation */
typedef enum {
  B200_ACT_SILU = 0,      /* silu_and_mul        activation_kernels.cu:26-30  */
  B200_ACT_GELU = 1,      /* gelu_and_mul        :32-40  (erf)                */
  B200_ACT_GELU_TANH = 2, /* gelu_tanh_and_mul   :42-53                       */
  B200_ACT_GELU_NEW = 3,  /* gelu_new            :118-123                     */
  B200_ACT_GELU_FAST = 4, /* gelu_fast           :125-131                     */
  B200_ACT_GELU_QUICK = 5 /* gelu_quick          :133-137                     */
} b200_activation_id;

/* {silu,gelu,gelu_tanh}_and_mul (activation_kernels.cu:12-24, 57-89):
 * input [num_tokens, 2*d] contiguous, out [num_tokens, d] = act(input[:, :d]) * input[:, d:]. */
B200_API int b200_act_and_mul(void* out, const void* input, int32_t num_tokens, int32_t d, int32_t act,
                              int32_t dtype, void* stream);
/* gelu_new / gelu_fast / gelu_quick (activation_kernels.cu:94-157): element-wise, [num_tokens, d]. */
B200_API int b200_activation(void* out, const void* input, int32_t num_tokens, int32_t d, int32_t act,
                             int32_t dtype, void* stream);

/* B200 extension (no reference op): rotary_embedding followed by reshape_and_cache
 * (csrc/pos_encoding_kernels.cu:127-160 + csrc/cache_kernels.cu:152-203, 253-278) in one launch --
 * the pair the reference issues back to back on every decode step
 * (model_executor/layers/rotary_embedding.py:forward_cuda, attention/backends/*:forward).
 * Results are identical to calling the two ops in sequence.  Returns B200_ENOTSUP when rows are
 * not 16-byte aligned or x != 16/sizeof(cache element); callers then issue the two ops. */
typedef struct {
  const int64_t* positions; /* [num_tokens] */
  void* query;              /* rotated in place */
  void* key;                /* rotated in place, then cached */
  const void* value;
  const void* cos_sin_cache;
  void* key_cache;   /* [num_blocks, num_kv_heads, head_size/x, block_size, x] */
  void* value_cache; /* [num_blocks, num_kv_heads, head_size, block_size] */
  const int64_t* slot_mapping; /* [num_tokens]; < 0: token is rotated but not cached */
  int32_t num_tokens, num_heads, num_kv_heads, head_size, rot_dim, block_size, x;
  int64_t query_stride, key_stride, value_stride;
  float kv_scale;
  int32_t is_neox;
  int32_t dtype;    /* b200_dtype of query/key/value */
  int32_t kv_dtype; /* b200_kv_dtype of the caches */
} b200_rotary_cache_params;

B200_API int b200_rotary_embedding_and_cache(const b200_rotary_cache_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * One-shot all-reduce over NVLink peer memory for the latency-bound
 * decode all-reduces of the row-parallel linears.  Replaces the one-shot path of
 * csrc/custom_all_reduce.cuh:179-253 / custom_all_reduce.cu:50-148 (ops _C_custom_ar::*,
 * csrc/torch_bindings.cpp:262-294; Python: distributed/device_communicators/custom_all_reduce.py).
 *
 * Every rank allocates one buffer of b200_ar_state_bytes() + 2 * max_bytes with b200_ar_alloc
 * (zeroed; returns its 64-byte CUDA IPC handle), exchanges the handles out of band, opens the
 * peers' with b200_ar_open, and passes the world_size base pointers (its own at index rank) to
 * every b200_all_reduce call.  Sum in fp32 in rank order: all ranks produce identical bits.
 * Sizes: numel * element size must be a multiple of 16 and <= max_bytes (else B200_ENOTSUP: the
 * caller falls back to NCCL, as the reference does).  World sizes 2, 4, 6, 8.  Graph-capturable.
 * ------------------------------------------------------------------------------------------ */
B200_API int64_t b200_ar_state_bytes(void);
B200_API int b200_ar_alloc(int64_t max_bytes, void** ptr, unsigned char* ipc_handle /* [64] */);
B200_API int b200_ar_open(const unsigned char* ipc_handle /* [64] */, void** ptr);
B200_API int b200_ar_close(void* ptr);
B200_API int b200_ar_free(void* ptr);
B200_API int b200_all_reduce(void* const* bases, int32_t rank, int32_t world, const void* input,
                             void* output, int64_t numel, int32_t dtype, int64_t max_bytes,
                             void* stream);

/* B200 extension (no reference op): the all-reduce of the row-parallel partial sums fused with the
 * fused_add_rms_norm that follows it in every decoder layer (linear.py:792-793 +
 * layernorm.py forward_cuda -> csrc/layernorm_kernels.cu:199-287), in place:
 *   input = sum over ranks; residual += input; input = rms_norm(residual) * weight
 * with the rounding steps of the two separate ops.  B200_ENOTSUP when the message does not fit the
 * flag-in-data path (row > 32 KiB or total > 2 MiB): callers then issue the two ops. */
B200_API int b200_all_reduce_add_rms_norm(void* const* bases, int32_t rank, int32_t world, void* input,
                                          void* residual, const void* weight, float epsilon,
                                          int32_t num_tokens, int32_t hidden, int32_t dtype,
                                          int64_t max_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_VLLM_H_ */
