// paged_attention_v2 for callers that cannot allocate: the partition scratch (tmp_out / exp_sums /
// max_logits) lives in a file-scope device array of fixed size instead of caller-owned buffers.
// No counterpart in the reference; added for code under CUDA
// graph capture, tight serving loops and harnesses that forbid cudaMalloc.  The arithmetic is
// b200_paged_attention_v2's (paged_attention.cu): this file only supplies the scratch pointers.
//
// Every GPU that loads the library holds the array (kStaticScratchBytes) until exit.  All calls on
// a device share it and it keeps the last call's contents: the partition kernel writes every
// (sequence, head, partition) entry the reduce kernel reads, so nothing has to start at zero --
// but two calls that overlap on the GPU (different streams, no ordering) race on it.
#include <mutex>

#include "common.cuh"

namespace b200 {
namespace {

constexpr size_t kStaticScratchBytes = 16u << 20;  // 16 MiB per device
static __device__ __align__(256) unsigned char g_attn_scratch[kStaticScratchBytes];

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// Device address of the array on the CURRENT device.  In host code `g_attn_scratch` is only a
// stand-in for the symbol: the address a kernel may use comes from cudaGetSymbolAddress, once per
// device, with that device current.
int scratch_base(unsigned char** out) {
  static unsigned char* base[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    set_error("paged_attention_v2_static: cannot identify the current device (%s)", cudaGetErrorString(e));
    return B200_ECUDA;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (base[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_attn_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      cudaGetLastError();
      set_error("paged_attention_v2_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return B200_ECUDA;
    }
    base[dev] = static_cast<unsigned char*>(ptr);
  }
  *out = base[dev];
  return B200_OK;
}

}  // namespace
}  // namespace b200

extern "C" {

int64_t b200_paged_attention_v2_static_scratch_max_bytes(void) { return (int64_t)b200::kStaticScratchBytes; }

int64_t b200_paged_attention_v2_static_scratch_bytes(int32_t num_seqs, int32_t num_heads, int32_t head_size,
                                                     int32_t max_seq_len, int32_t dtype) {
  if (num_seqs <= 0 || num_heads <= 0 || head_size <= 0 || max_seq_len < 0) return -1;
  int64_t parts = ((int64_t)max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  if (parts < 1) parts = 1;
  const int64_t entries = (int64_t)num_seqs * num_heads * parts;
  const int64_t elem = dtype == B200_F32 ? 4 : 2;
  return (int64_t)b200::align256((size_t)entries * 4) * 2 + (int64_t)b200::align256((size_t)(entries * head_size * elem));
}

int b200_paged_attention_v2_static(const b200_paged_attention_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "paged_attention_v2_static: null params");
  B200_REQUIRE(pp->num_seqs > 0 && pp->num_heads > 0 && pp->head_size > 0 && pp->max_seq_len >= 0, B200_EINVAL,
               "paged_attention_v2_static: bad sizes");
  const int64_t need = b200_paged_attention_v2_static_scratch_bytes(pp->num_seqs, pp->num_heads, pp->head_size,
                                                                    pp->max_seq_len, pp->dtype);
  B200_REQUIRE(need >= 0 && need <= (int64_t)kStaticScratchBytes, B200_EINVAL,
               "paged_attention_v2_static: scratch for num_seqs %d x num_heads %d x %d partitions x head_size %d needs "
               "%lld bytes, the built-in array holds %lld: use b200_paged_attention_v2 with caller-owned scratch",
               pp->num_seqs, pp->num_heads, (pp->max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE,
               pp->head_size, (long long)need, (long long)kStaticScratchBytes);
  unsigned char* base = nullptr;
  const int rc = scratch_base(&base);
  if (rc != B200_OK) return rc;
  int64_t parts = ((int64_t)pp->max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  if (parts < 1) parts = 1;
  const size_t stats = align256((size_t)pp->num_seqs * pp->num_heads * parts * 4);
  b200_paged_attention_params p = *pp;  // inputs / outputs stay the caller's
  p.exp_sums = reinterpret_cast<float*>(base);
  p.max_logits = reinterpret_cast<float*>(base + stats);
  p.tmp_out = base + 2 * stats;
  return b200_paged_attention_v2(&p, stream);
}

}  // extern "C"
