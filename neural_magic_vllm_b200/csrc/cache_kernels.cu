// Paged KV-cache maintenance kernels for B200.
// Replaces csrc/cache_kernels.cu of the reference (reshape_and_cache :152-278, _flash :206-314,
// copy_blocks :67-148, swap_blocks :24-63, convert_fp8 :316-389).
//
// All of these are HBM/latency-bound byte movers: the work is laid out so that every global access
// that CAN be 16 bytes wide is (K-cache rows are x = 16 B contiguous; whole blocks for copy_blocks),
// and the unavoidable token-strided V-cache stores are issued from consecutive lanes for consecutive
// head-dim rows so one warp store instruction touches 32 distinct 32-B sectors exactly once.
#include "common.cuh"

namespace b200 {

// One CTA per token. A work item is one x-wide group of the head dimension of one head:
// 16 B of K-cache, x strided elements of V-cache.
template <typename T, typename CacheT, int KV>
__global__ void reshape_and_cache_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                         CacheT* __restrict__ key_cache,
                                         CacheT* __restrict__ value_cache,
                                         const int64_t* __restrict__ slot_mapping,
                                         const int64_t key_stride, const int64_t value_stride,
                                         const int num_heads, const int head_size,
                                         const int block_size, const int x, const float kv_scale) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding token (cache_kernels.cu:166-169)
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int groups_per_head = head_size / x;
  const int n_groups = num_heads * groups_per_head;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  CacheT* kdst = key_cache + block_idx * (int64_t)num_heads * head_size * block_size;
  CacheT* vdst = value_cache + block_idx * (int64_t)num_heads * head_size * block_size;

  for (int gidx = threadIdx.x; gidx < n_groups; gidx += blockDim.x) {
    const int head = gidx / groups_per_head;
    const int xg = gidx % groups_per_head;
    const T* kp = ksrc + head * head_size + xg * x;
    const T* vp = vsrc + head * head_size + xg * x;
    CacheT* kd = kdst + ((int64_t)head * groups_per_head + xg) * block_size * x + block_off * x;
    CacheT* vd = vdst + ((int64_t)head * head_size + xg * x) * block_size + block_off;
    for (int j = 0; j < x; ++j) {
      T kv = kp[j];
      T vv = vp[j];
      if constexpr (KV == B200_KV_AUTO) {
        kd[j] = kv;
        vd[(int64_t)j * block_size] = vv;
      } else {
        kd[j] = fp8_quant<T, KV>(kv, kv_scale);
        vd[(int64_t)j * block_size] = fp8_quant<T, KV>(vv, kv_scale);
      }
    }
  }
}

// Vectorised variant: x elements of K move as one 16-byte store. Requires 16-B aligned key rows.
// Lanes are mapped so that consecutive lanes own consecutive x-groups (coalesced 16-B K loads and
// stores per lane; V stores of lane l, element j hit row (xg*x + j)).
template <typename T, typename CacheT, int KV, int X>
__global__ void reshape_and_cache_vec_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                             CacheT* __restrict__ key_cache,
                                             CacheT* __restrict__ value_cache,
                                             const int64_t* __restrict__ slot_mapping,
                                             const int64_t key_stride, const int64_t value_stride,
                                             const int num_heads, const int head_size,
                                             const int block_size, const float kv_scale) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int groups_per_head = head_size / X;
  const int n_groups = num_heads * groups_per_head;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  CacheT* kdst = key_cache + block_idx * (int64_t)num_heads * head_size * block_size;
  CacheT* vdst = value_cache + block_idx * (int64_t)num_heads * head_size * block_size;
  constexpr int SRC_VECS = (X * (int)sizeof(T)) / 16;  // 1 (same width) or 2/4 (fp8 caches)

  for (int gidx = threadIdx.x; gidx < n_groups; gidx += blockDim.x) {
    const int head = gidx / groups_per_head;
    const int xg = gidx % groups_per_head;
    union {
      uint4 v[SRC_VECS];
      T e[X];
    } kin, vin;
    const uint4* kp = reinterpret_cast<const uint4*>(ksrc + head * head_size + xg * X);
    const uint4* vp = reinterpret_cast<const uint4*>(vsrc + head * head_size + xg * X);
#pragma unroll
    for (int i = 0; i < SRC_VECS; ++i) {
      kin.v[i] = __ldg(kp + i);
      vin.v[i] = __ldg(vp + i);
    }
    CacheT* kd = kdst + ((int64_t)head * groups_per_head + xg) * block_size * X + block_off * X;
    CacheT* vd = vdst + ((int64_t)head * head_size + xg * X) * block_size + block_off;
    union {
      uint4 v;
      CacheT e[X];
    } kout;
    static_assert(sizeof(CacheT) * X == 16, "x-group must be 16 bytes of cache");
#pragma unroll
    for (int j = 0; j < X; ++j) {
      if constexpr (KV == B200_KV_AUTO) {
        kout.e[j] = kin.e[j];
        vd[(int64_t)j * block_size] = vin.e[j];
      } else {
        kout.e[j] = fp8_quant<T, KV>(kin.e[j], kv_scale);
        vd[(int64_t)j * block_size] = fp8_quant<T, KV>(vin.e[j], kv_scale);
      }
    }
    *reinterpret_cast<uint4*>(kd) = kout.v;
  }
}

template <typename T>
__global__ void reshape_and_cache_flash_kernel(const T* __restrict__ key,
                                               const T* __restrict__ value, T* __restrict__ k_cache,
                                               T* __restrict__ v_cache,
                                               const int64_t* __restrict__ slot_mapping,
                                               const int64_t block_stride, const int64_t key_stride,
                                               const int64_t value_stride, const int num_heads,
                                               const int head_size, const int block_size) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int n = num_heads * head_size;
  T* kd = k_cache + block_idx * block_stride + block_off * n;
  T* vd = v_cache + block_idx * block_stride + block_off * n;
  const T* ks = key + token * key_stride;
  const T* vs = value + token * value_stride;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    kd[i] = ks[i];
    vd[i] = vs[i];
  }
}

// grid (num_layers, num_pairs): whole-block 16-byte copies for K then V.
__global__ void copy_blocks_kernel(const int64_t* __restrict__ key_cache_ptrs,
                                   const int64_t* __restrict__ value_cache_ptrs,
                                   const int64_t* __restrict__ block_mapping,
                                   const int64_t vecs_per_block) {
  const int layer = blockIdx.x;
  const int pair = blockIdx.y;
  uint4* kc = reinterpret_cast<uint4*>(key_cache_ptrs[layer]);
  uint4* vc = reinterpret_cast<uint4*>(value_cache_ptrs[layer]);
  const int64_t src = block_mapping[2 * pair] * vecs_per_block;
  const int64_t dst = block_mapping[2 * pair + 1] * vecs_per_block;
  for (int64_t i = threadIdx.x; i < vecs_per_block; i += blockDim.x) kc[dst + i] = kc[src + i];
  for (int64_t i = threadIdx.x; i < vecs_per_block; i += blockDim.x) vc[dst + i] = vc[src + i];
}

template <typename T, int KV, bool TO_FP8>
__global__ void convert_fp8_kernel(void* __restrict__ dst, const void* __restrict__ src,
                                   const int64_t numel, const float scale) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += stride) {
    if constexpr (TO_FP8) {
      reinterpret_cast<uint8_t*>(dst)[i] = fp8_quant<T, KV>(reinterpret_cast<const T*>(src)[i], scale);
    } else {
      reinterpret_cast<T*>(dst)[i] = fp8_dequant<T, KV>(reinterpret_cast<const uint8_t*>(src)[i], scale);
    }
  }
}

template <typename T, typename CacheT, int KV>
static int launch_reshape_and_cache(const b200_reshape_and_cache_params* p, cudaStream_t stream) {
  constexpr int X = 16 / (int)sizeof(CacheT);
  const int n_groups = p->num_heads * (p->head_size / (p->x > 0 ? p->x : 1));
  const bool vec_ok = p->x == X && p->head_size % X == 0 &&
                      (reinterpret_cast<uintptr_t>(p->key) % 16 == 0) &&
                      (reinterpret_cast<uintptr_t>(p->value) % 16 == 0) &&
                      (reinterpret_cast<uintptr_t>(p->key_cache) % 16 == 0) &&
                      ((p->key_stride * (int64_t)sizeof(T)) % 16 == 0) &&
                      ((p->value_stride * (int64_t)sizeof(T)) % 16 == 0) &&
                      ((p->head_size * (int64_t)sizeof(T)) % 16 == 0);
  int threads = n_groups < 32 ? 32 : (n_groups > 256 ? 256 : ((n_groups + 31) / 32) * 32);
  if (vec_ok) {
    reshape_and_cache_vec_kernel<T, CacheT, KV, X><<<p->num_tokens, threads, 0, stream>>>(
        (const T*)p->key, (const T*)p->value, (CacheT*)p->key_cache, (CacheT*)p->value_cache,
        p->slot_mapping, p->key_stride, p->value_stride, p->num_heads, p->head_size, p->block_size,
        p->kv_scale);
  } else {
    reshape_and_cache_kernel<T, CacheT, KV><<<p->num_tokens, threads, 0, stream>>>(
        (const T*)p->key, (const T*)p->value, (CacheT*)p->key_cache, (CacheT*)p->value_cache,
        p->slot_mapping, p->key_stride, p->value_stride, p->num_heads, p->head_size, p->block_size,
        p->x, p->kv_scale);
  }
  return check_launch("reshape_and_cache");
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_reshape_and_cache(const b200_reshape_and_cache_params* p, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_REQUIRE(p != nullptr, B200_EINVAL, "reshape_and_cache: null params");
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->num_tokens > 0 && p->num_heads > 0 && p->head_size > 0 && p->block_size > 0 &&
                   p->x > 0 && p->head_size % p->x == 0,
               B200_EINVAL, "reshape_and_cache: bad shape (tokens=%d heads=%d head_size=%d block=%d x=%d)",
               p->num_tokens, p->num_heads, p->head_size, p->block_size, p->x);
#define RC_CASE(T)                                                                             \
  switch (p->kv_dtype) {                                                                       \
    case B200_KV_AUTO:                                                                         \
      return launch_reshape_and_cache<T, T, B200_KV_AUTO>(p, stream);                          \
    case B200_KV_FP8_E4M3:                                                                     \
      return launch_reshape_and_cache<T, uint8_t, B200_KV_FP8_E4M3>(p, stream);                \
    case B200_KV_FP8_E5M2:                                                                     \
      return launch_reshape_and_cache<T, uint8_t, B200_KV_FP8_E5M2>(p, stream);                \
    default:                                                                                   \
      break;                                                                                   \
  }                                                                                            \
  break;
  switch (p->dtype) {
    case B200_F16:
      RC_CASE(__half)
    case B200_BF16:
      RC_CASE(__nv_bfloat16)
    case B200_F32:
      RC_CASE(float)
    default:
      break;
  }
#undef RC_CASE
  set_error("reshape_and_cache: unsupported dtype %d / kv dtype %d", p->dtype, p->kv_dtype);
  return B200_EINVAL;
}

int b200_reshape_and_cache_flash(const b200_reshape_and_cache_flash_params* p, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_REQUIRE(p != nullptr, B200_EINVAL, "reshape_and_cache_flash: null params");
  if (p->num_tokens == 0) return B200_OK;
  const int n = p->num_heads * p->head_size;
  const int threads = n < 512 ? ((n + 31) / 32) * 32 : 512;
  switch (p->dtype) {
    case B200_F16:
      reshape_and_cache_flash_kernel<__half><<<p->num_tokens, threads, 0, stream>>>(
          (const __half*)p->key, (const __half*)p->value, (__half*)p->k_cache, (__half*)p->v_cache,
          p->slot_mapping, p->block_stride, p->key_stride, p->value_stride, p->num_heads,
          p->head_size, p->block_size);
      break;
    case B200_BF16:
      reshape_and_cache_flash_kernel<__nv_bfloat16><<<p->num_tokens, threads, 0, stream>>>(
          (const __nv_bfloat16*)p->key, (const __nv_bfloat16*)p->value, (__nv_bfloat16*)p->k_cache,
          (__nv_bfloat16*)p->v_cache, p->slot_mapping, p->block_stride, p->key_stride,
          p->value_stride, p->num_heads, p->head_size, p->block_size);
      break;
    case B200_F32:
      reshape_and_cache_flash_kernel<float><<<p->num_tokens, threads, 0, stream>>>(
          (const float*)p->key, (const float*)p->value, (float*)p->k_cache, (float*)p->v_cache,
          p->slot_mapping, p->block_stride, p->key_stride, p->value_stride, p->num_heads,
          p->head_size, p->block_size);
      break;
    default:
      set_error("reshape_and_cache_flash: unsupported dtype %d", p->dtype);
      return B200_EINVAL;
  }
  return check_launch("reshape_and_cache_flash");
}

int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs,
                     const int64_t* block_mapping, int32_t num_layers, int32_t num_pairs,
                     int64_t bytes_per_block, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (num_layers == 0 || num_pairs == 0) return B200_OK;
  B200_REQUIRE(bytes_per_block > 0 && bytes_per_block % 16 == 0, B200_EINVAL,
               "copy_blocks: bytes_per_block=%lld must be a positive multiple of 16",
               (long long)bytes_per_block);
  const int64_t vecs = bytes_per_block / 16;
  const int threads = vecs < 1024 ? (int)((vecs + 31) / 32) * 32 : 1024;
  dim3 grid(num_layers, num_pairs);
  copy_blocks_kernel<<<grid, threads, 0, stream>>>(key_cache_ptrs, value_cache_ptrs, block_mapping,
                                                   vecs);
  return check_launch("copy_blocks");
}

int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host,
                     int32_t num_pairs, int64_t bytes_per_block, int32_t kind, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  cudaMemcpyKind k;
  switch (kind) {
    case 0: k = cudaMemcpyDeviceToDevice; break;
    case 1: k = cudaMemcpyDeviceToHost; break;
    case 2: k = cudaMemcpyHostToDevice; break;
    default:
      set_error("swap_blocks: Invalid device combination");
      return B200_EINVAL;
  }
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  for (int i = 0; i < num_pairs; ++i) {
    const int64_t sb = block_mapping_host[2 * i], db = block_mapping_host[2 * i + 1];
    cudaError_t e = cudaMemcpyAsync(d + db * bytes_per_block, s + sb * bytes_per_block,
                                    (size_t)bytes_per_block, k, stream);
    if (e != cudaSuccess) {
      cudaGetLastError();
      set_error("swap_blocks: cudaMemcpyAsync: %s", cudaGetErrorString(e));
      return B200_ECUDA;
    }
  }
  return B200_OK;
}

int b200_convert_fp8(void* dst, const void* src, int64_t numel, float scale, int32_t dtype,
                     int32_t kv_dtype, int32_t to_fp8, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (numel == 0) return B200_OK;
  B200_REQUIRE(kv_dtype == B200_KV_FP8_E4M3 || kv_dtype == B200_KV_FP8_E5M2, B200_EINVAL,
               "convert_fp8: Unsupported data type: kv dtype enum %d", kv_dtype);
  const int threads = 256;
  int64_t blocks64 = (numel + threads - 1) / threads;
  const int blocks = (int)(blocks64 > num_sms() * 16 ? num_sms() * 16 : blocks64);
#define CV(T, KV)                                                                              \
  if (to_fp8)                                                                                  \
    convert_fp8_kernel<T, KV, true><<<blocks, threads, 0, stream>>>(dst, src, numel, scale);   \
  else                                                                                         \
    convert_fp8_kernel<T, KV, false><<<blocks, threads, 0, stream>>>(dst, src, numel, scale);
#define CV_KV(T)                         \
  if (kv_dtype == B200_KV_FP8_E4M3) {    \
    CV(T, B200_KV_FP8_E4M3)              \
  } else {                               \
    CV(T, B200_KV_FP8_E5M2)              \
  }
  switch (dtype) {
    case B200_F16: CV_KV(__half) break;
    case B200_BF16: CV_KV(__nv_bfloat16) break;
    case B200_F32: CV_KV(float) break;
    default:
      set_error("convert_fp8: unsupported dtype %d", dtype);
      return B200_EINVAL;
  }
#undef CV_KV
#undef CV
  return check_launch("convert_fp8");
}

}  // extern "C"
