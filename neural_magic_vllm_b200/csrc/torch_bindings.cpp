// Torch-library registration of libb200vllm: the drop-in replacement of the reference's
// csrc/torch_bindings.cpp.  It registers the SAME op schemas under the same namespaces
// (torch.ops._C, torch.ops._C_cache_ops, torch.ops._C_cuda_utils) for the hot-path ops, so
// vllm/_custom_ops.py and everything above it call into this library unchanged.
//
// This file is the only place that sees torch types: each op validates its tensors with the
// reference's checks (cited inline), unpacks them into the POD parameter blocks of
// include/b200_vllm.h and calls the C ABI on the current CUDA stream.  No op synchronises or
// allocates outside the torch caching allocator, so all of them are CUDA-graph capturable.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <cstdlib>
#include <torch/library.h>

#include <optional>
#include <string>
#include <vector>

#include "b200_vllm.h"

#ifndef B200_TORCH_NS
#define B200_TORCH_NS _C
#endif
#define B200_CAT_(a, b) a##b
#define B200_CAT(a, b) B200_CAT_(a, b)
// TORCH_LIBRARY pastes its first argument: expand our macros one level earlier
#define B200_TORCH_LIBRARY(NS, M) TORCH_LIBRARY(NS, M)

namespace {

using at::Tensor;

// Message building without iostreams: every multi-part check message goes through cat().
inline void cat_one(std::string& s, const char* v) { s += v; }
inline void cat_one(std::string& s, const std::string& v) { s += v; }
inline void cat_one(std::string& s, bool v) { s += v ? "true" : "false"; }
inline void cat_one(std::string& s, int v) { s += std::to_string(v); }
inline void cat_one(std::string& s, long v) { s += std::to_string(v); }
inline void cat_one(std::string& s, long long v) { s += std::to_string(v); }
inline void cat_one(std::string& s, unsigned long v) { s += std::to_string(v); }
inline void cat_one(std::string& s, double v) { s += std::to_string(v); }
inline void cat_one(std::string& s, at::ScalarType v) { s += c10::toString(v); }
template <typename... A>
std::string cat(const A&... a) {
  std::string s;
  (cat_one(s, a), ...);
  return s;
}
template <typename... A>
std::string cat_or(const char* dflt, const A&... a) {
  if constexpr (sizeof...(A) == 0) return std::string(dflt);
  else return cat(a...);
}
#define B200_TCHECK(cond, ...)                                                              \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      TORCH_CHECK(false, cat_or("Expected " #cond " to be true, but got false.", ##__VA_ARGS__)); \
    }                                                                                       \
  } while (0)

void* cur_stream() { return static_cast<void*>(at::cuda::getCurrentCUDAStream().stream()); }

void check_rc(int rc) {
  if (rc == B200_OK) return;
  const std::string msg = b200_last_error();
  if (rc == B200_ENOTSUP) TORCH_CHECK_NOT_IMPLEMENTED(false, msg);
  TORCH_CHECK(false, msg);
}

int dtype_of(const Tensor& t, const char* what) {
  switch (t.scalar_type()) {
    case at::ScalarType::Half: return B200_F16;
    case at::ScalarType::BFloat16: return B200_BF16;
    case at::ScalarType::Float: return B200_F32;
    default: TORCH_CHECK(false, cat(what, ": unsupported data type ", t.scalar_type()));
  }
  return -1;
}

int kv_dtype_of(const std::string& s) {
  const int kv = b200_parse_kv_cache_dtype(s.c_str());
  B200_TCHECK(kv >= 0, "Unsupported data type of kv cache: ", s);
  return kv;
}

// ---------------------------------------------------------------------------------------------
// paged attention (csrc/attention/attention_kernels.cu:805-826, 966-990)
// ---------------------------------------------------------------------------------------------
void fill_attention(b200_paged_attention_params& p, Tensor& out, Tensor& query, Tensor& key_cache,
                    Tensor& value_cache, int64_t num_kv_heads, double scale, Tensor& block_tables,
                    Tensor& seq_lens, int64_t block_size, int64_t max_seq_len,
                    const c10::optional<Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                    double kv_scale, int64_t tp_rank, int64_t bs_local, int64_t bs_vert,
                    int64_t bs_block, int64_t bs_step) {
  B200_TCHECK(query.dim() == 3 && out.dim() == 3, "query/out must be [num_seqs, num_heads, head_size]");
  B200_TCHECK(query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query heads must be contiguous");
  B200_TCHECK(out.is_contiguous(), "out must be contiguous");
  B200_TCHECK(block_tables.scalar_type() == at::ScalarType::Int && block_tables.dim() == 2 &&
                  block_tables.stride(1) == 1,
              "block_tables must be a row-contiguous int32 [num_seqs, max_blocks] tensor");
  B200_TCHECK(seq_lens.scalar_type() == at::ScalarType::Int && seq_lens.is_contiguous(),
              "seq_lens must be a contiguous int32 tensor");
  B200_TCHECK(bs_vert <= 1 || bs_block > 0, "blocksparse_block_size must be positive");
  // attention_kernels.cu:984-986: blocksparse_block_size must divide into / by block_size
  if (bs_vert > 1) {
    B200_TCHECK(bs_block % block_size == 0, "Block sparse attention only supports "
                "blocksparse_block_size being a multiple of block_size");
  }
  p = b200_paged_attention_params{};
  p.out = out.data_ptr();
  p.query = query.data_ptr();
  p.key_cache = key_cache.data_ptr();
  p.value_cache = value_cache.data_ptr();
  p.block_tables = block_tables.data_ptr<int32_t>();
  p.seq_lens = seq_lens.data_ptr<int32_t>();
  p.alibi_slopes = nullptr;
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    B200_TCHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_contiguous(),
                "alibi_slopes must be a contiguous fp32 tensor");
    p.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  p.num_seqs = (int32_t)query.size(0);
  p.num_heads = (int32_t)query.size(1);
  p.head_size = (int32_t)query.size(2);
  p.num_kv_heads = (int32_t)num_kv_heads;
  p.block_size = (int32_t)block_size;
  p.max_num_blocks_per_seq = (int32_t)block_tables.size(1);
  p.max_seq_len = (int32_t)max_seq_len;
  p.q_stride = query.stride(0);
  p.kv_block_stride = key_cache.stride(0);
  p.kv_head_stride = key_cache.stride(1);
  p.scale = (float)scale;
  p.kv_scale = (float)kv_scale;
  p.dtype = dtype_of(query, "paged_attention");
  p.kv_dtype = kv_dtype_of(kv_cache_dtype);
  if (p.kv_dtype == B200_KV_AUTO) {
    B200_TCHECK(key_cache.scalar_type() == query.scalar_type() &&
                    value_cache.scalar_type() == query.scalar_type(),
                "kv_cache_dtype 'auto' requires caches of the query dtype");
  } else {
    B200_TCHECK(key_cache.scalar_type() == at::ScalarType::Byte &&
                    value_cache.scalar_type() == at::ScalarType::Byte,
                "fp8 kv caches must be uint8 tensors");
  }
  p.tp_rank = (int32_t)tp_rank;
  p.blocksparse_local_blocks = (int32_t)bs_local;
  p.blocksparse_vert_stride = (int32_t)bs_vert;
  p.blocksparse_block_size = (int32_t)bs_block;
  p.blocksparse_head_sliding_step = (int32_t)bs_step;
}

void paged_attention_v1(Tensor& out, Tensor& query, Tensor& key_cache, Tensor& value_cache,
                        int64_t num_kv_heads, double scale, Tensor& block_tables, Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len,
                        const c10::optional<Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                        double kv_scale, int64_t tp_rank, int64_t bs_local, int64_t bs_vert,
                        int64_t bs_block, int64_t bs_step) {
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  b200_paged_attention_params p;
  fill_attention(p, out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                 block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, bs_local,
                 bs_vert, bs_block, bs_step);
  // v1 has no scratch in its signature, so one CTA per (kv head, sequence) walks the whole
  // context.  B200_ATTN_V1_SPLIT = 1 serves the op with the partitioned kernel instead (scratch from
  // the torch caching allocator: stream-ordered, CUDA-graph safe; same arithmetic as
  // paged_attention_v2).  It was the default for grids between three and four CTAs per SM until
  // the kernel learned to run those as one wave of 3-warp CTAs (paged_attention.cu: batch 64 step
  // 4.17 -> 4.01 ms); measured slower than v1 everywhere else (batch 32: 3.23 -> 3.43 ms, batch
  // 256: 11.2 -> 11.8 ms; other partition sizes, 128 .. 384 tokens, no better).
  const int64_t parts = (max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  static const int split_knob = [] {
    const char* e = std::getenv("B200_ATTN_V1_SPLIT");
    return e ? std::atoi(e) : -1;
  }();
  const bool split = split_knob > 0 && parts > 1 && query.scalar_type() != at::ScalarType::Float;
  if (split) {
    Tensor tmp_out = at::empty({query.size(0), query.size(1), parts, query.size(2)}, query.options());
    Tensor exp_sums = at::empty({query.size(0), query.size(1), parts}, query.options().dtype(at::ScalarType::Float));
    Tensor max_logits = at::empty_like(exp_sums);
    p.exp_sums = exp_sums.data_ptr<float>();
    p.max_logits = max_logits.data_ptr<float>();
    p.tmp_out = tmp_out.data_ptr();
    check_rc(b200_paged_attention_v2(&p, cur_stream()));
    return;
  }
  check_rc(b200_paged_attention_v1(&p, cur_stream()));
}

// extension: variable-length causal attention for prefill (stands in for flash_attn_varlen_func /
// xformers of the reference's prefill path)
void varlen_attention_fwd(Tensor& out, const Tensor& q, const Tensor& k, const Tensor& v,
                          const Tensor& cu_seqlens_q, const Tensor& cu_seqlens_k, int64_t max_seqlen_q,
                          int64_t max_seqlen_k, double softmax_scale, bool causal) {
  B200_TCHECK(q.is_cuda() && k.is_cuda() && v.is_cuda() && out.is_cuda(), "varlen_attention_fwd: CUDA tensors");
  B200_TCHECK(q.dim() == 3 && k.dim() == 3 && v.dim() == 3 && out.dim() == 3,
              "varlen_attention_fwd: q / k / v / out must be [tokens, heads, head_size]");
  B200_TCHECK(q.stride(2) == 1 && k.stride(2) == 1 && v.stride(2) == 1 && out.stride(2) == 1 &&
                  q.stride(1) == q.size(2) && k.stride(1) == k.size(2) && v.stride(1) == v.size(2) &&
                  out.stride(1) == out.size(2),
              "varlen_attention_fwd: heads must be contiguous inside a token");
  B200_TCHECK(k.sizes() == v.sizes() && q.size(2) == k.size(2) && out.sizes() == q.sizes(), "varlen_attention_fwd: shapes");
  B200_TCHECK(q.scalar_type() == k.scalar_type() && q.scalar_type() == v.scalar_type() &&
                  q.scalar_type() == out.scalar_type(), "varlen_attention_fwd: dtypes must match");
  B200_TCHECK(cu_seqlens_q.scalar_type() == at::ScalarType::Int && cu_seqlens_k.scalar_type() == at::ScalarType::Int &&
                  cu_seqlens_q.is_cuda() && cu_seqlens_k.is_cuda() && cu_seqlens_q.is_contiguous() &&
                  cu_seqlens_k.is_contiguous() && cu_seqlens_q.numel() == cu_seqlens_k.numel() &&
                  cu_seqlens_q.numel() >= 1,
              "varlen_attention_fwd: cu_seqlens must be int32 CUDA vectors of num_seqs + 1 entries");
  const at::cuda::OptionalCUDAGuard guard(device_of(q));
  b200_varlen_attention_params p{};
  p.out = out.data_ptr();
  p.q = q.data_ptr();
  p.k = k.data_ptr();
  p.v = v.data_ptr();
  p.cu_seqlens_q = cu_seqlens_q.data_ptr<int32_t>();
  p.cu_seqlens_k = cu_seqlens_k.data_ptr<int32_t>();
  p.num_seqs = (int32_t)cu_seqlens_q.numel() - 1;
  p.num_heads = (int32_t)q.size(1);
  p.num_kv_heads = (int32_t)k.size(1);
  p.head_size = (int32_t)q.size(2);
  p.total_q = (int32_t)q.size(0);
  p.total_k = (int32_t)k.size(0);
  p.max_seqlen_q = (int32_t)max_seqlen_q;
  p.max_seqlen_k = (int32_t)max_seqlen_k;
  p.q_stride = q.stride(0);
  p.k_stride = k.stride(0);
  p.v_stride = v.stride(0);
  p.out_stride = out.stride(0);
  p.softmax_scale = (float)softmax_scale;
  p.causal = causal ? 1 : 0;
  p.dtype = dtype_of(q, "varlen_attention_fwd");
  check_rc(b200_varlen_attention_fwd(&p, cur_stream()));
}

void paged_attention_v2(Tensor& out, Tensor& exp_sums, Tensor& max_logits, Tensor& tmp_out,
                        Tensor& query, Tensor& key_cache, Tensor& value_cache, int64_t num_kv_heads,
                        double scale, Tensor& block_tables, Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const c10::optional<Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double kv_scale, int64_t tp_rank,
                        int64_t bs_local, int64_t bs_vert, int64_t bs_block, int64_t bs_step) {
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  b200_paged_attention_params p;
  fill_attention(p, out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                 block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, bs_local,
                 bs_vert, bs_block, bs_step);
  const int64_t parts = (max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  B200_TCHECK(exp_sums.scalar_type() == at::ScalarType::Float && exp_sums.is_contiguous() &&
                  max_logits.scalar_type() == at::ScalarType::Float && max_logits.is_contiguous(),
              "exp_sums / max_logits must be contiguous fp32 tensors");
  B200_TCHECK(exp_sums.numel() >= query.size(0) * query.size(1) * parts &&
                  max_logits.numel() >= query.size(0) * query.size(1) * parts &&
                  tmp_out.numel() >= query.size(0) * query.size(1) * parts * query.size(2),
              "paged_attention_v2 scratch tensors are too small for max_seq_len=", max_seq_len);
  B200_TCHECK(tmp_out.scalar_type() == query.scalar_type() && tmp_out.is_contiguous(),
              "tmp_out must be a contiguous tensor of the query dtype");
  p.exp_sums = exp_sums.data_ptr<float>();
  p.max_logits = max_logits.data_ptr<float>();
  p.tmp_out = tmp_out.data_ptr();
  check_rc(b200_paged_attention_v2(&p, cur_stream()));
}

// ---------------------------------------------------------------------------------------------
// cache ops (csrc/cache_kernels.cu)
// ---------------------------------------------------------------------------------------------
void reshape_and_cache(Tensor& key, Tensor& value, Tensor& key_cache, Tensor& value_cache,
                       Tensor& slot_mapping, const std::string& kv_cache_dtype, double kv_scale) {
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  B200_TCHECK(key.dim() == 3 && value.dim() == 3, "key/value must be [num_tokens, num_heads, head_size]");
  B200_TCHECK(key_cache.dim() == 5 && value_cache.dim() == 4, "bad kv cache rank");
  B200_TCHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be a contiguous int64 tensor");
  B200_TCHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value heads must be contiguous");
  B200_TCHECK(key_cache.is_contiguous() && value_cache.is_contiguous(), "kv caches must be contiguous");
  b200_reshape_and_cache_params p{};
  p.key = key.data_ptr();
  p.value = value.data_ptr();
  p.key_cache = key_cache.data_ptr();
  p.value_cache = value_cache.data_ptr();
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.num_tokens = (int32_t)key.size(0);
  p.num_heads = (int32_t)key.size(1);
  p.head_size = (int32_t)key.size(2);
  p.block_size = (int32_t)key_cache.size(3);
  p.x = (int32_t)key_cache.size(4);
  p.key_stride = key.stride(0);
  p.value_stride = value.stride(0);
  p.kv_scale = (float)kv_scale;
  p.dtype = dtype_of(key, "reshape_and_cache");
  p.kv_dtype = kv_dtype_of(kv_cache_dtype);
  if (p.kv_dtype == B200_KV_AUTO) {
    B200_TCHECK(key_cache.scalar_type() == key.scalar_type(), "cache dtype must match key dtype");
  } else {
    B200_TCHECK(key_cache.scalar_type() == at::ScalarType::Byte, "fp8 kv caches must be uint8");
  }
  check_rc(b200_reshape_and_cache(&p, cur_stream()));
}

void reshape_and_cache_flash(Tensor& key, Tensor& value, Tensor& k_cache, Tensor& v_cache,
                             Tensor& slot_mapping, const std::string& kv_cache_dtype) {
  // cache_kernels.cu:288-290
  B200_TCHECK(kv_cache_dtype == "auto", "unsupported kv_cache_dtype (str), supported kv_cache_dtype: auto");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  B200_TCHECK(k_cache.stride(0) == v_cache.stride(0), "k/v cache block strides differ");
  b200_reshape_and_cache_flash_params p{};
  p.key = key.data_ptr();
  p.value = value.data_ptr();
  p.k_cache = k_cache.data_ptr();
  p.v_cache = v_cache.data_ptr();
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.num_tokens = (int32_t)key.size(0);
  p.num_heads = (int32_t)key.size(1);
  p.head_size = (int32_t)key.size(2);
  p.block_size = (int32_t)k_cache.size(1);
  p.key_stride = key.stride(0);
  p.value_stride = value.stride(0);
  p.block_stride = k_cache.stride(0);
  p.dtype = dtype_of(key, "reshape_and_cache_flash");
  check_rc(b200_reshape_and_cache_flash(&p, cur_stream()));
}

void copy_blocks(std::vector<Tensor> const& key_caches, std::vector<Tensor> const& value_caches,
                 const Tensor& block_mapping) {
  const int64_t num_layers = (int64_t)key_caches.size();
  B200_TCHECK(num_layers == (int64_t)value_caches.size(), "key/value cache lists differ in length");
  if (num_layers == 0) return;
  const auto dev = key_caches[0].device();
  B200_TCHECK(dev.is_cuda(), "copy_blocks: caches must be CUDA tensors");
  const at::cuda::OptionalCUDAGuard guard(dev);
  const int64_t num_pairs = block_mapping.size(0);
  if (num_pairs == 0) return;
  // pointer tables: built on the host, shipped with one async copy from pinned memory each
  auto opts = at::TensorOptions().dtype(at::kLong).device(at::kCPU).pinned_memory(true);
  Tensor kp = at::empty({num_layers}, opts), vp = at::empty({num_layers}, opts);
  for (int64_t l = 0; l < num_layers; ++l) {
    kp.data_ptr<int64_t>()[l] = reinterpret_cast<int64_t>(key_caches[l].data_ptr());
    vp.data_ptr<int64_t>()[l] = reinterpret_cast<int64_t>(value_caches[l].data_ptr());
  }
  Tensor kpd = kp.to(dev, /*non_blocking=*/true), vpd = vp.to(dev, /*non_blocking=*/true);
  Tensor bm = block_mapping.to(dev, at::kLong, /*non_blocking=*/true).contiguous();
  const int64_t bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  check_rc(b200_copy_blocks(kpd.data_ptr<int64_t>(), vpd.data_ptr<int64_t>(), bm.data_ptr<int64_t>(),
                            (int32_t)num_layers, (int32_t)num_pairs, bytes, cur_stream()));
}

void swap_blocks(Tensor& src, Tensor& dst, const Tensor& block_mapping) {
  // cache_kernels.cu:24-63
  const auto sd = src.device(), dd = dst.device();
  int kind;
  if (sd.is_cuda() && dd.is_cuda()) {
    B200_TCHECK(sd.index() == dd.index(), "src and dst must be on the same GPU");
    kind = 0;
  } else if (sd.is_cuda() && dd.is_cpu()) {
    kind = 1;
  } else if (sd.is_cpu() && dd.is_cuda()) {
    kind = 2;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  B200_TCHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  Tensor bm = block_mapping.to(at::kLong).contiguous();
  const int64_t bytes = src.element_size() * src[0].numel();
  const at::cuda::OptionalCUDAGuard guard(sd.is_cuda() ? sd : dd);
  check_rc(b200_swap_blocks(src.data_ptr(), dst.data_ptr(), bm.data_ptr<int64_t>(),
                            (int32_t)bm.size(0), bytes, kind, cur_stream()));
}

void convert_fp8(Tensor& dst_cache, Tensor& src_cache, double scale, const std::string& kv_cache_dtype) {
  // cache_kernels.cu:339-389
  const auto sd = src_cache.device(), dd = dst_cache.device();
  B200_TCHECK(sd.is_cuda(), "src must be on a GPU");
  B200_TCHECK(dd.is_cuda(), "dst must be on a GPU");
  B200_TCHECK(sd.index() == dd.index(), "src and dst must be on the same GPU");
  const at::cuda::OptionalCUDAGuard guard(sd);
  B200_TCHECK(src_cache.is_contiguous() && dst_cache.is_contiguous(), "caches must be contiguous");
  B200_TCHECK(src_cache.numel() == dst_cache.numel(), "src and dst sizes differ");
  // cache_kernels.cu:372-388: only "fp8" / "fp8_e4m3" convert.  "fp8_e5m2" (and anything else) is a
  // host-side TORCH_CHECK(false) there; "auto" reaches `assert(false)` in device code
  // (quant_utils.cuh:513-523, kAuto) -- an error either way, raised here on the host.
  B200_TCHECK(kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3", "Unsupported data type: ", kv_cache_dtype);
  const int kv = B200_KV_FP8_E4M3;
  const bool to_fp8 = dst_cache.scalar_type() == at::ScalarType::Byte;
  const Tensor& wide = to_fp8 ? src_cache : dst_cache;
  B200_TCHECK((to_fp8 ? src_cache : dst_cache).scalar_type() != at::ScalarType::Byte &&
                  (to_fp8 || src_cache.scalar_type() == at::ScalarType::Byte),
              "convert_fp8: exactly one of src/dst must be uint8");
  check_rc(b200_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(),
                            (float)scale, dtype_of(wide, "convert_fp8"), kv, to_fp8 ? 1 : 0,
                            cur_stream()));
}

// ---------------------------------------------------------------------------------------------
// cuda utils (csrc/cuda_utils_kernels.cu)
// ---------------------------------------------------------------------------------------------
int64_t get_device_attribute(int64_t attribute, int64_t device_id) {
  int32_t v = 0;
  check_rc(b200_get_device_attribute((int32_t)attribute, (int32_t)device_id, &v));
  return v;
}

int64_t get_max_shared_memory_per_block_device_attribute(int64_t device_id) {
  int32_t v = 0;
  check_rc(b200_get_max_shared_memory_per_block_device_attribute((int32_t)device_id, &v));
  return v;
}

}  // namespace

#include "torch_bindings_quant.inc"
#include "torch_bindings_glue.inc"

B200_TORCH_LIBRARY(B200_TORCH_NS, ops) {
  ops.def(
      "paged_attention_v1("
      "    Tensor! out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float kv_scale, int tp_rank,"
      "    int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v1", c10::DispatchKey::CUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2("
      "    Tensor! out, Tensor exp_sums, Tensor max_logits,"
      "    Tensor tmp_out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float kv_scale, int tp_rank,"
      "    int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v2", c10::DispatchKey::CUDA, &paged_attention_v2);
  ops.def(
      "varlen_attention_fwd(Tensor! out, Tensor q, Tensor k, Tensor v, Tensor cu_seqlens_q, Tensor cu_seqlens_k, "
      "int max_seqlen_q, int max_seqlen_k, float softmax_scale, bool causal) -> ()");
  ops.impl("varlen_attention_fwd", c10::DispatchKey::CUDA, &varlen_attention_fwd);
  b200_register_quant_ops(ops);
  b200_register_glue_ops(ops);
}

B200_TORCH_LIBRARY(B200_CAT(B200_TORCH_NS, _cache_ops), cache_ops) {
  cache_ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  // src or dst may live on the host: dispatch on whichever argument is CUDA
  cache_ops.impl("swap_blocks", c10::DispatchKey::CUDA, &swap_blocks);
  cache_ops.impl("swap_blocks", c10::DispatchKey::CPU, &swap_blocks);
  cache_ops.def(
      "copy_blocks(Tensor[]! key_caches, Tensor[]! value_caches, Tensor block_mapping) -> ()");
  cache_ops.impl("copy_blocks", c10::DispatchKey::CUDA, &copy_blocks);
  cache_ops.def(
      "reshape_and_cache(Tensor key, Tensor value,"
      "                  Tensor! key_cache, Tensor! value_cache,"
      "                  Tensor slot_mapping,"
      "                  str kv_cache_dtype,"
      "                  float kv_scale) -> ()");
  cache_ops.impl("reshape_and_cache", c10::DispatchKey::CUDA, &reshape_and_cache);
  cache_ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value,"
      "                        Tensor! key_cache,"
      "                        Tensor! value_cache,"
      "                        Tensor slot_mapping,"
      "                        str kv_cache_dtype) -> ()");
  cache_ops.impl("reshape_and_cache_flash", c10::DispatchKey::CUDA, &reshape_and_cache_flash);
  cache_ops.def(
      "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  cache_ops.impl("convert_fp8", c10::DispatchKey::CUDA, &convert_fp8);
}

B200_TORCH_LIBRARY(B200_CAT(B200_TORCH_NS, _cuda_utils), cuda_utils) {
  cuda_utils.def("get_device_attribute(int attribute, int device_id) -> int");
  cuda_utils.impl("get_device_attribute", c10::DispatchKey::CompositeExplicitAutograd, &get_device_attribute);
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute(int device_id) -> int");
  cuda_utils.impl("get_max_shared_memory_per_block_device_attribute",
                  c10::DispatchKey::CompositeExplicitAutograd,
                  &get_max_shared_memory_per_block_device_attribute);
}
