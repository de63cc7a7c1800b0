// Decode-side glue between the quantized linears and paged attention (sm_100a) -- SURVEY.md §8(f) rank 1.
//   rms_norm / fused_add_rms_norm      replaces csrc/layernorm_kernels.cu:21-45, 199-287, 291-352
//   rotary_embedding (+ batched)       replaces csrc/pos_encoding_kernels.cu:10-125, 127-203
//   silu/gelu/gelu_tanh _and_mul,
//   gelu_new / gelu_fast / gelu_quick  replaces csrc/activation_kernels.cu:12-71, 77-157
//   rotary_embedding_and_cache         B200 extension: rotary + reshape_and_cache
//                                      (csrc/cache_kernels.cu:152-203) in ONE launch
//
// All of these are HBM/latency-bound element-wise work: at decode they move a few KiB per token,
// so what matters is one launch per op, 16-byte accesses and no second trip to HBM (the RMSNorm row
// is staged in shared memory between the sum-of-squares pass and the scale pass).
//
// Rounding follows the reference kernels step by step.  The reference computes in `scalar_t`
// (c10::Half / c10::BFloat16): every binary operator is "fp32 op, round to scalar_t".  Here a value
// of type T is carried as an fp32 holding a T-representable number and `rnd<T>` re-rounds after
// each operator, with __fmul_rn/__fadd_rn so that nvcc cannot contract across a rounding point.
#include "common.cuh"

#include <math.h>

namespace b200 {
namespace {

template <typename T>
__device__ __forceinline__ float rnd(float v) {
  return DT<T>::to_f(DT<T>::from_f(v));
}
template <>
__device__ __forceinline__ float rnd<float>(float v) {
  return v;
}
template <typename T>
__device__ __forceinline__ float tmul(float a, float b) {
  return rnd<T>(__fmul_rn(a, b));
}
template <typename T>
__device__ __forceinline__ float tadd(float a, float b) {
  return rnd<T>(__fadd_rn(a, b));
}
template <typename T>
__device__ __forceinline__ float tsub(float a, float b) {
  return rnd<T>(__fsub_rn(a, b));
}

template <typename T>
struct V16 {  // 16 bytes of T
  static constexpr int N = 16 / (int)sizeof(T);
  union {
    uint4 u;
    T e[N];
  };
  __device__ __forceinline__ void load(const T* p) { u = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(T* p) const { *reinterpret_cast<uint4*>(p) = u; }
  __device__ __forceinline__ float f(int i) const { return DT<T>::to_f(e[i]); }
  __device__ __forceinline__ void set(int i, float v) { e[i] = DT<T>::from_f(v); }
};

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = (l < nw) ? red[l] : 0.f;
  r = warp_sum(r);
  return r;  // every thread holds the total
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---------------------------------------------------------------------------------------------
// RMSNorm.  One CTA per token.  FUSED: z = input + residual (rounded to T), residual = z,
// input = T(z * rstd) * weight.  Plain: out = T(x * rstd) * weight.   (layernorm_kernels.cu:21-45,
// 199-249).  VEC: 16-byte accesses, the row parked in shared memory between the two passes.
// ---------------------------------------------------------------------------------------------
// SPLITK: `in` does not exist yet -- it is the sum of `splits` fp32 partial planes that a split-K
// GEMM left in its workspace (ws[s * plane + row * hidden + i], summed s = 0, 1, ... in that order
// and rounded to T exactly as the GEMM's own split-K sum kernel would); the rest of the kernel is
// unchanged, so GEMM -> fused_add_rms_norm and this fused form agree bit for bit.
template <typename T, bool FUSED, bool VEC, bool SPLITK = false>
__global__ void __launch_bounds__(1024) rms_norm_kernel(T* out, const T* in,
                                                        T* __restrict__ residual,
                                                        const T* __restrict__ weight, const float eps,
                                                        const int hidden, const float* __restrict__ ws,
                                                        const int splits, const int64_t plane) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ float red[32];
  pdl_wait();
  pdl_launch_dependents();
  T* row = reinterpret_cast<T*>(smem_raw);
  const int64_t base = (int64_t)blockIdx.x * hidden;
  float ss = 0.f;
  if constexpr (VEC) {
    constexpr int N = V16<T>::N;
    for (int i = threadIdx.x * N; i < hidden; i += blockDim.x * N) {
      V16<T> a;
      if constexpr (SPLITK) {
        static_assert(V16<T>::N == 8, "two float4 per vector");
        float4 lo = *reinterpret_cast<const float4*>(ws + base + i);
        float4 hi = *reinterpret_cast<const float4*>(ws + base + i + 4);
        for (int sp = 1; sp < splits; ++sp) {
          const float4 l2 = *reinterpret_cast<const float4*>(ws + sp * plane + base + i);
          const float4 h2 = *reinterpret_cast<const float4*>(ws + sp * plane + base + i + 4);
          lo.x += l2.x, lo.y += l2.y, lo.z += l2.z, lo.w += l2.w;
          hi.x += h2.x, hi.y += h2.y, hi.z += h2.z, hi.w += h2.w;
        }
        a.set(0, lo.x), a.set(1, lo.y), a.set(2, lo.z), a.set(3, lo.w);
        a.set(4, hi.x), a.set(5, hi.y), a.set(6, hi.z), a.set(7, hi.w);
      } else {
        a.load(in + base + i);
      }
      if constexpr (FUSED) {
        V16<T> r;
        r.load(residual + base + i);
#pragma unroll
        for (int j = 0; j < N; ++j) a.set(j, __fadd_rn(a.f(j), r.f(j)));
        a.store(residual + base + i);
      }
#pragma unroll
      for (int j = 0; j < N; ++j) ss = fmaf(a.f(j), a.f(j), ss);
      a.store(row + i);
    }
  } else {
    for (int i = threadIdx.x; i < hidden; i += blockDim.x) {
      float x = DT<T>::to_f(in[base + i]);
      if constexpr (FUSED) {
        x = tadd<T>(x, DT<T>::to_f(residual[base + i]));
        residual[base + i] = DT<T>::from_f(x);
      }
      ss = fmaf(x, x, ss);
    }
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / hidden + eps);
  if constexpr (VEC) {
    constexpr int N = V16<T>::N;
    // each thread re-reads exactly the vectors it staged: no barrier needed on `row`
    for (int i = threadIdx.x * N; i < hidden; i += blockDim.x * N) {
      V16<T> a, w;
      a.load(row + i);
      w.load(weight + i);
#pragma unroll
      for (int j = 0; j < N; ++j) a.set(j, __fmul_rn(rnd<T>(__fmul_rn(a.f(j), rstd)), w.f(j)));
      a.store(out + base + i);
    }
  } else {
    const T* src = FUSED ? residual : in;
    for (int i = threadIdx.x; i < hidden; i += blockDim.x) {
      const float x = DT<T>::to_f(src[base + i]);
      out[base + i] = DT<T>::from_f(__fmul_rn(rnd<T>(__fmul_rn(x, rstd)), DT<T>::to_f(weight[i])));
    }
  }
}

template <typename T, bool FUSED>
int launch_rms(void* out, const void* in, void* residual, const void* weight, float eps, int tokens,
               int hidden, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  const size_t row_bytes = (size_t)hidden * sizeof(T);
  const bool vec = hidden % N == 0 && aligned16(out) && aligned16(in) && aligned16(weight) &&
                   (!FUSED || aligned16(residual)) && row_bytes <= 48 * 1024;
  if (vec) {
    int threads = ((hidden / N + 31) / 32) * 32;
    threads = threads > 1024 ? 1024 : threads;
    launch_pdl(rms_norm_kernel<T, FUSED, true>, dim3(tokens), dim3(threads), row_bytes, st, (T*)out,
               (const T*)in, (T*)residual, (const T*)weight, eps, hidden, (const float*)nullptr, 0, (int64_t)0);
  } else {
    int threads = ((hidden + 31) / 32) * 32;
    threads = threads > 1024 ? 1024 : threads;
    launch_pdl(rms_norm_kernel<T, FUSED, false>, dim3(tokens), dim3(threads), 0, st, (T*)out,
               (const T*)in, (T*)residual, (const T*)weight, eps, hidden, (const float*)nullptr, 0, (int64_t)0);
  }
  return check_launch(FUSED ? "fused_add_rms_norm" : "rms_norm");
}

// split-K sum + residual add + RMSNorm in one kernel behind a split-K GEMM (marlin_gemm.cu):
// out = norm(round(sum of partials) + residual) * weight, residual updated in place
template <typename T>
int launch_splitk_add_rms(void* out, const float* ws, int splits, void* residual, const void* weight, float eps,
                          int tokens, int hidden, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  int threads = ((hidden / N + 31) / 32) * 32;
  threads = threads > 1024 ? 1024 : threads;
  launch_pdl(rms_norm_kernel<T, true, true, true>, dim3(tokens), dim3(threads), (size_t)hidden * sizeof(T), st, (T*)out,
             (const T*)nullptr, (T*)residual, (const T*)weight, eps, hidden, ws, splits, (int64_t)tokens * hidden);
  return check_launch("gptq_marlin_gemm_add_rms_norm (split-K sum + add + norm)");
}

}  // namespace

bool splitk_add_rms_eligible(int hidden, const void* out, const void* residual, const void* weight) {
  return hidden % 8 == 0 && (size_t)hidden * 2 <= 48 * 1024 && aligned16(out) && aligned16(residual) &&
         aligned16(weight);
}
int splitk_add_rms_norm(int dtype, void* out, const float* ws, int splits, void* residual, const void* weight,
                        float eps, int tokens, int hidden, cudaStream_t st) {
  return dtype == B200_F16
             ? launch_splitk_add_rms<__half>(out, ws, splits, residual, weight, eps, tokens, hidden, st)
             : launch_splitk_add_rms<__nv_bfloat16>(out, ws, splits, residual, weight, eps, tokens, hidden, st);
}

namespace {

// ---------------------------------------------------------------------------------------------
// (fused_add_)rms_norm + activation quantisation in one pass: the W8A8
// linears take int8 / fp8 activations, so the normalised row never goes to HBM in 16 bits.
// Arithmetic = the two reference ops back to back: y = T(T(z * rstd) * w) exactly as
// rms_norm_kernel stores it (layernorm_kernels.cu:21-45, 199-249), then
//   QMODE 0  int8, per-token dynamic: scale = absmax(y) / 127, q = cvt.rni.sat(y * (127 / absmax))
//   QMODE 1  int8, static:            q = cvt.rni.sat(y / *scale)   (int8_quant_kernels.cu:27-66)
//   QMODE 2  fp8 e4m3, static:        q = e4m3(clamp(y * (1 / *scale), +-448))  (fp8/common.cu:18-76)
// One CTA per token, the row parked in shared memory between the passes.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int8_t f2i8_rni_sat(float x) {
  int32_t d;
  asm("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(d) : "f"(x));
  return static_cast<int8_t>(d);
}
__device__ __forceinline__ float block_max_f(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();  // `red` may still be read by the previous reduction
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = (l < nw) ? red[l] : 0.f;
  return warp_max(r);
}

template <typename T, bool FUSED, int QMODE>
__global__ void __launch_bounds__(1024) rms_norm_quant_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                                              const T* __restrict__ in, T* __restrict__ residual,
                                                              const T* __restrict__ weight, const float eps,
                                                              const int hidden) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ float red[32];
  pdl_wait();
  pdl_launch_dependents();
  constexpr int N = V16<T>::N;
  T* row = reinterpret_cast<T*>(smem_raw);
  const int64_t base = (int64_t)blockIdx.x * hidden;
  float ss = 0.f;
  for (int i = threadIdx.x * N; i < hidden; i += blockDim.x * N) {
    V16<T> a;
    a.load(in + base + i);
    if constexpr (FUSED) {
      V16<T> r;
      r.load(residual + base + i);
#pragma unroll
      for (int j = 0; j < N; ++j) a.set(j, __fadd_rn(a.f(j), r.f(j)));
      a.store(residual + base + i);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) ss = fmaf(a.f(j), a.f(j), ss);
    a.store(row + i);
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / hidden + eps);
  float amax = 0.f;
  // each thread re-reads exactly the vectors it staged: no barrier needed on `row`
  for (int i = threadIdx.x * N; i < hidden; i += blockDim.x * N) {
    V16<T> a, w;
    a.load(row + i);
    w.load(weight + i);
#pragma unroll
    for (int j = 0; j < N; ++j) {
      a.set(j, __fmul_rn(rnd<T>(__fmul_rn(a.f(j), rstd)), w.f(j)));
      if constexpr (QMODE == 0) amax = fmaxf(amax, fabsf(a.f(j)));
    }
    a.store(row + i);
  }
  float mul = 1.f, div = 1.f;
  if constexpr (QMODE == 0) {
    amax = block_max_f(amax, red);
    if (threadIdx.x == 0) scales[blockIdx.x] = amax / 127.0f;
    mul = 127.0f / amax;
  } else if constexpr (QMODE == 1) {
    div = *scales;
  } else {
    mul = 1.0f / (*scales);
  }
  for (int i = threadIdx.x * N; i < hidden; i += blockDim.x * N) {
    V16<T> a;
    a.load(row + i);
    uint8_t q[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const float y = a.f(j);
      if constexpr (QMODE == 0) q[j] = (uint8_t)f2i8_rni_sat(y * mul);
      else if constexpr (QMODE == 1) q[j] = (uint8_t)f2i8_rni_sat(y / div);
      else q[j] = (uint8_t)__nv_cvt_float_to_fp8(fmaxf(-448.0f, fminf(y * mul, 448.0f)), __NV_SATFINITE, __NV_E4M3);
    }
    if constexpr (N == 8) *reinterpret_cast<uint2*>(out + base + i) = *reinterpret_cast<uint2*>(q);
    else *reinterpret_cast<uint32_t*>(out + base + i) = *reinterpret_cast<uint32_t*>(q);
  }
}

template <typename T, bool FUSED>
int launch_rms_quant(uint8_t* out, float* scales, const void* in, void* residual, const void* weight, float eps,
                     int tokens, int hidden, int qmode, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  const size_t row_bytes = (size_t)hidden * sizeof(T);
  B200_REQUIRE(hidden % N == 0 && aligned16(in) && aligned16(weight) && (!FUSED || aligned16(residual)) &&
                   (reinterpret_cast<uintptr_t>(out) & 7) == 0 && row_bytes <= 200 * 1024,
               B200_EINVAL, "rms_norm_quant: rows must be 16-byte aligned multiples of 16 bytes, <= 200 KiB");
  int threads = ((hidden / N + 31) / 32) * 32;
  threads = threads > 1024 ? 1024 : threads;
  auto go = [&](auto kern) -> int {
    if (row_bytes > 48 * 1024) {
      // (the three quant modes share one function-pointer type: one cache row per mode)
      static size_t configured[3][kMaxDevices] = {};
      if (ensure_dyn_smem(kern, row_bytes, configured[qmode]) != cudaSuccess) {
        cudaGetLastError();
        set_error("rms_norm_quant: cannot reserve %d B of shared memory", (int)row_bytes);
        return B200_ECUDA;
      }
    }
    launch_pdl(kern, dim3(tokens), dim3(threads), row_bytes, st, out, scales, (const T*)in, (T*)residual,
               (const T*)weight, eps, hidden);
    return check_launch("rms_norm_quant");
  };
  switch (qmode) {
    case 0: return go(rms_norm_quant_kernel<T, FUSED, 0>);
    case 1: return go(rms_norm_quant_kernel<T, FUSED, 1>);
    case 2: return go(rms_norm_quant_kernel<T, FUSED, 2>);
  }
  set_error("rms_norm_quant: quant mode %d (0 int8 dynamic, 1 int8 static, 2 fp8 static)", qmode);
  return B200_EINVAL;
}

// ---------------------------------------------------------------------------------------------
// Rotary embedding, in place on query / key (pos_encoding_kernels.cu:10-96).
//   x' = x*cos - y*sin ; y' = y*cos + x*sin, every operator rounded to T.
// One CTA per token; a work item is 16 bytes of x (NeoX: + the matching 16 bytes of y).
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void rot_pair(float x, float y, float c, float s, float& xo, float& yo) {
  xo = tsub<T>(tmul<T>(x, c), tmul<T>(y, s));
  yo = tadd<T>(tmul<T>(y, c), tmul<T>(x, s));
}

template <typename T, bool NEOX, bool VEC>
__device__ __forceinline__ void rotary_token(T* __restrict__ q, T* __restrict__ k,
                                             const T* __restrict__ cache, const int rot_dim,
                                             const int num_heads, const int num_kv_heads,
                                             const int head_size) {
  const int embed = rot_dim / 2;
  const T* cosp = cache;
  const T* sinp = cache + embed;
  const int heads = num_heads + num_kv_heads;
  if constexpr (VEC) {
    constexpr int N = V16<T>::N;
    if constexpr (NEOX) {
      const int per_head = embed / N;
      for (int it = threadIdx.x; it < heads * per_head; it += blockDim.x) {
        const int h = it / per_head, ro = (it % per_head) * N;
        T* arr = (h < num_heads ? q + (int64_t)h * head_size : k + (int64_t)(h - num_heads) * head_size);
        V16<T> x, y, c, s;
        x.load(arr + ro);
        y.load(arr + embed + ro);
        c.load(cosp + ro);
        s.load(sinp + ro);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          float xo, yo;
          rot_pair<T>(x.f(j), y.f(j), c.f(j), s.f(j), xo, yo);
          x.set(j, xo);
          y.set(j, yo);
        }
        x.store(arr + ro);
        y.store(arr + embed + ro);
      }
    } else {
      constexpr int P = N / 2;  // pairs per 16 bytes
      const int per_head = embed / P;
      for (int it = threadIdx.x; it < heads * per_head; it += blockDim.x) {
        const int h = it / per_head, ro = (it % per_head) * P;
        T* arr = (h < num_heads ? q + (int64_t)h * head_size : k + (int64_t)(h - num_heads) * head_size);
        V16<T> v;
        v.load(arr + 2 * ro);
#pragma unroll
        for (int j = 0; j < P; ++j) {
          float xo, yo;
          rot_pair<T>(v.f(2 * j), v.f(2 * j + 1), DT<T>::to_f(cosp[ro + j]), DT<T>::to_f(sinp[ro + j]),
                      xo, yo);
          v.set(2 * j, xo);
          v.set(2 * j + 1, yo);
        }
        v.store(arr + 2 * ro);
      }
    }
  } else {
    for (int it = threadIdx.x; it < heads * embed; it += blockDim.x) {
      const int h = it / embed, ro = it % embed;
      T* arr = (h < num_heads ? q + (int64_t)h * head_size : k + (int64_t)(h - num_heads) * head_size);
      const int xi = NEOX ? ro : 2 * ro, yi = NEOX ? embed + ro : 2 * ro + 1;
      float xo, yo;
      rot_pair<T>(DT<T>::to_f(arr[xi]), DT<T>::to_f(arr[yi]), DT<T>::to_f(cosp[ro]),
                  DT<T>::to_f(sinp[ro]), xo, yo);
      arr[xi] = DT<T>::from_f(xo);
      arr[yi] = DT<T>::from_f(yo);
    }
  }
}

template <typename T, bool NEOX, bool VEC>
__global__ void __launch_bounds__(512) rotary_kernel(const int64_t* __restrict__ positions,
                                                     T* __restrict__ query, T* __restrict__ key,
                                                     const T* __restrict__ cos_sin_cache,
                                                     const int64_t* __restrict__ offsets,
                                                     const int rot_dim, const int64_t q_stride,
                                                     const int64_t k_stride, const int num_heads,
                                                     const int num_kv_heads, const int head_size) {
  const int64_t t = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  int64_t pos = positions[t];
  if (offsets != nullptr) pos += offsets[t];  // batched_rotary_embedding (:98-125)
  rotary_token<T, NEOX, VEC>(query + t * q_stride, key + t * k_stride, cos_sin_cache + pos * rot_dim,
                             rot_dim, num_heads, num_kv_heads, head_size);
}

// B200 extension: rotary on q/k and the paged-cache write of the same token's K/V
// (cache_kernels.cu:152-203) in one pass: the rotated K chunk goes from registers both back to `key`
// and into its cache slot, so nothing is re-read.  A work item is 16 bytes of one head row:
// q rotary | k rotary + cache | k tail beyond rot_dim -> cache | v -> cache.
template <int BYTES>
__device__ __forceinline__ void store_small(void* dst, const void* src) {
  if constexpr (BYTES == 16) *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
  else if constexpr (BYTES == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src);
  else *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src);
}

template <typename T, typename CacheT, int KV, bool NEOX>
__global__ void __launch_bounds__(512)
    rotary_cache_kernel(const int64_t* __restrict__ positions, T* __restrict__ query,
                        T* __restrict__ key, const T* __restrict__ value,
                        const T* __restrict__ cos_sin_cache, const int rot_dim, const int64_t q_stride,
                        const int64_t k_stride, const int64_t v_stride, const int num_heads,
                        const int num_kv_heads, const int head_size, CacheT* __restrict__ key_cache,
                        CacheT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping,
                        const int block_size, const float kv_scale) {
  constexpr int N = V16<T>::N;
  constexpr int X = 16 / (int)sizeof(CacheT);  // elements per 16-byte K-cache group (X % N == 0)
  const int64_t t = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  const int embed = rot_dim / 2;
  const T* cosp = cos_sin_cache + positions[t] * rot_dim;
  const T* sinp = cosp + embed;
  T* qt = query + t * q_stride;
  T* kt = key + t * k_stride;
  const T* vt = value + t * v_stride;
  const int64_t slot = slot_mapping[t];
  const bool cached = slot >= 0;  // padding token: rotated, not cached (cache_kernels.cu:166-169)
  const int64_t block_idx = cached ? slot / block_size : 0;
  const int64_t block_off = cached ? slot % block_size : 0;
  CacheT* kdst = key_cache + block_idx * (int64_t)num_kv_heads * head_size * block_size;
  CacheT* vdst = value_cache + block_idx * (int64_t)num_kv_heads * head_size * block_size + block_off;
  const int groups = head_size / X;

  auto put_k = [&](const V16<T>& v, int head, int d0) {
    CacheT* kd = kdst + (((int64_t)head * groups + d0 / X) * block_size + block_off) * X + d0 % X;
    if constexpr (KV == B200_KV_AUTO) {
      store_small<16>(kd, &v.u);
    } else {
      __align__(8) uint8_t b[N];
#pragma unroll
      for (int j = 0; j < N; ++j) b[j] = fp8_quant<T, KV>(v.e[j], kv_scale);
      store_small<N>(kd, b);
    }
  };

  const int rot_per_head = NEOX ? embed / N : embed / (N / 2);
  const int tail_per_head = (head_size - rot_dim) / N;
  const int v_per_head = head_size / N;
  const int n_rot = (num_heads + num_kv_heads) * rot_per_head;
  const int n_q = num_heads * rot_per_head;
  const int n_tail = cached ? num_kv_heads * tail_per_head : 0;
  const int n_v = cached ? num_kv_heads * v_per_head : 0;
  for (int it = threadIdx.x; it < n_rot + n_tail + n_v; it += blockDim.x) {
    if (it < n_rot) {
      const bool is_k = it >= n_q;
      const int i = is_k ? it - n_q : it;
      const int h = i / rot_per_head, c = i % rot_per_head;
      T* arr = (is_k ? kt : qt) + (int64_t)h * head_size;
      if constexpr (NEOX) {
        const int ro = c * N;
        V16<T> x, y, cs, sn;
        x.load(arr + ro);
        y.load(arr + embed + ro);
        cs.load(cosp + ro);
        sn.load(sinp + ro);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          float xo, yo;
          rot_pair<T>(x.f(j), y.f(j), cs.f(j), sn.f(j), xo, yo);
          x.set(j, xo);
          y.set(j, yo);
        }
        x.store(arr + ro);
        y.store(arr + embed + ro);
        if (is_k && cached) {
          put_k(x, h, ro);
          put_k(y, h, embed + ro);
        }
      } else {
        constexpr int P = N / 2;
        const int ro = c * P;
        V16<T> v;
        v.load(arr + 2 * ro);
#pragma unroll
        for (int j = 0; j < P; ++j) {
          float xo, yo;
          rot_pair<T>(v.f(2 * j), v.f(2 * j + 1), DT<T>::to_f(cosp[ro + j]), DT<T>::to_f(sinp[ro + j]),
                      xo, yo);
          v.set(2 * j, xo);
          v.set(2 * j + 1, yo);
        }
        v.store(arr + 2 * ro);
        if (is_k && cached) put_k(v, h, 2 * ro);
      }
    } else if (it < n_rot + n_tail) {  // K beyond rot_dim: cached as is
      const int i = it - n_rot;
      const int h = i / tail_per_head, d0 = rot_dim + (i % tail_per_head) * N;
      V16<T> v;
      v.load(kt + (int64_t)h * head_size + d0);
      put_k(v, h, d0);
    } else {  // V: token-strided scatter, consecutive lanes -> consecutive 16-byte chunks of the row
      const int i = it - n_rot - n_tail;
      const int h = i / v_per_head, d0 = (i % v_per_head) * N;
      V16<T> v;
      v.u = __ldg(reinterpret_cast<const uint4*>(vt + (int64_t)h * head_size + d0));
      CacheT* vd = vdst + ((int64_t)h * head_size + d0) * block_size;
#pragma unroll
      for (int j = 0; j < N; ++j) {
        if constexpr (KV == B200_KV_AUTO) vd[(int64_t)j * block_size] = v.e[j];
        else vd[(int64_t)j * block_size] = fp8_quant<T, KV>(v.e[j], kv_scale);
      }
    }
  }
}

struct RotArgs {
  const int64_t* positions;
  void* query;
  void* key;
  const void* cache;
  const int64_t* offsets;
  int tokens, rot_dim, num_heads, num_kv_heads, head_size;
  int64_t q_stride, k_stride;
  bool neox;
};

template <typename T>
bool rotary_vec_ok(const RotArgs& a) {
  constexpr int N = V16<T>::N;
  return (a.rot_dim / 2) % N == 0 && a.head_size % N == 0 && a.q_stride % N == 0 &&
         a.k_stride % N == 0 && aligned16(a.query) && aligned16(a.key) && aligned16(a.cache);
}

template <typename T>
int launch_rotary(const RotArgs& a, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  const bool vec = rotary_vec_ok<T>(a);
  const int embed = a.rot_dim / 2;
  const int items = (a.num_heads + a.num_kv_heads) * (vec ? (a.neox ? embed / N : embed / (N / 2)) : embed);
  int threads = ((items + 31) / 32) * 32;
  threads = threads > 512 ? 512 : (threads < 32 ? 32 : threads);
#define ROT_LAUNCH(NEOX, VEC)                                                                    \
  launch_pdl(rotary_kernel<T, NEOX, VEC>, dim3(a.tokens), dim3(threads), 0, st, a.positions,      \
             (T*)a.query, (T*)a.key, (const T*)a.cache, a.offsets, a.rot_dim, a.q_stride,        \
             a.k_stride, a.num_heads, a.num_kv_heads, a.head_size)
  if (a.neox) {
    if (vec) ROT_LAUNCH(true, true); else ROT_LAUNCH(true, false);
  } else {
    if (vec) ROT_LAUNCH(false, true); else ROT_LAUNCH(false, false);
  }
#undef ROT_LAUNCH
  return check_launch("rotary_embedding");
}

// ---------------------------------------------------------------------------------------------
// Activations (activation_kernels.cu:26-53, 118-140).  Inputs/outputs are T-representable floats.
// ---------------------------------------------------------------------------------------------
template <typename T, int ACT>
__device__ __forceinline__ float act_fn(float x) {
  if constexpr (ACT == B200_ACT_SILU) {
    return rnd<T>(x / (1.0f + expf(-x)));
  } else if constexpr (ACT == B200_ACT_GELU) {
    return rnd<T>(x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f)));
  } else if constexpr (ACT == B200_ACT_GELU_TANH) {
    const float beta = 1.41421356237309504880f * 1.12837916709551257390f * 0.5f;
    const float cube = x * x * x;
    const float inner = beta * (x + 0.044715f * cube);
    return rnd<T>(0.5f * x * (1.0f + tanhf(inner)));
  } else if constexpr (ACT == B200_ACT_GELU_NEW) {
    const float x3 = tmul<T>(tmul<T>(x, x), x);
    const float a = tadd<T>(x, rnd<T>(0.044715f * x3));
    const float t = rnd<T>(tanhf(rnd<T>(0.79788456f * a)));
    return tmul<T>(tmul<T>(0.5f, x), tadd<T>(1.0f, t));
  } else if constexpr (ACT == B200_ACT_GELU_FAST) {
    const float a = rnd<T>(x * 0.79788456f);
    const float b = tmul<T>(rnd<T>(0.044715f * x), x);
    const float t = rnd<T>(tanhf(tmul<T>(a, tadd<T>(1.0f, b))));
    return tmul<T>(tmul<T>(0.5f, x), tadd<T>(1.0f, t));
  } else {  // B200_ACT_GELU_QUICK
    return rnd<T>(x / (1.0f + expf(-1.702f * x)));
  }
}

// grid (tokens, chunks).  GATED: input [tokens, 2, d] -> out = act(x) * y; else input [tokens, d].
template <typename T, int ACT, bool GATED, bool VEC>
__global__ void __launch_bounds__(256) act_kernel(T* __restrict__ out, const T* __restrict__ in,
                                                  const int d) {
  const int64_t t = blockIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  const T* x = in + t * (GATED ? 2 : 1) * (int64_t)d;
  const T* y = x + d;
  T* o = out + t * (int64_t)d;
  if constexpr (VEC) {
    constexpr int N = V16<T>::N;
    for (int i = (blockIdx.y * blockDim.x + threadIdx.x) * N; i < d; i += gridDim.y * blockDim.x * N) {
      V16<T> a, b;
      a.load(x + i);
      if constexpr (GATED) b.load(y + i);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float r = act_fn<T, ACT>(a.f(j));
        if constexpr (GATED) r = __fmul_rn(r, b.f(j));
        a.set(j, r);
      }
      a.store(o + i);
    }
  } else {
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < d; i += gridDim.y * blockDim.x) {
      float r = act_fn<T, ACT>(DT<T>::to_f(x[i]));
      if constexpr (GATED) r = __fmul_rn(r, DT<T>::to_f(y[i]));
      o[i] = DT<T>::from_f(r);
    }
  }
}

template <typename T, int ACT, bool GATED>
int launch_act(void* out, const void* in, int tokens, int d, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  const bool vec = d % N == 0 && aligned16(out) && aligned16(in);
  const int per_cta = 256 * (vec ? N : 1);
  int chunks = (d + per_cta - 1) / per_cta;
  // enough CTAs to cover the SMs at small batch; one pass per CTA at large batch
  const int cap = tokens >= 4 * num_sms() ? 1 : (4 * num_sms() + tokens - 1) / tokens;
  chunks = chunks > cap ? cap : chunks;
  dim3 grid(tokens, chunks < 1 ? 1 : chunks);
  if (vec)
    launch_pdl(act_kernel<T, ACT, GATED, true>, grid, dim3(256), 0, st, (T*)out, (const T*)in, d);
  else
    launch_pdl(act_kernel<T, ACT, GATED, false>, grid, dim3(256), 0, st, (T*)out, (const T*)in, d);
  return check_launch(GATED ? "act_and_mul" : "activation");
}

template <typename T, bool GATED>
int dispatch_act(void* out, const void* in, int tokens, int d, int act, cudaStream_t st) {
  switch (act) {
    case B200_ACT_SILU: return launch_act<T, B200_ACT_SILU, GATED>(out, in, tokens, d, st);
    case B200_ACT_GELU: return launch_act<T, B200_ACT_GELU, GATED>(out, in, tokens, d, st);
    case B200_ACT_GELU_TANH: return launch_act<T, B200_ACT_GELU_TANH, GATED>(out, in, tokens, d, st);
    case B200_ACT_GELU_NEW: return launch_act<T, B200_ACT_GELU_NEW, GATED>(out, in, tokens, d, st);
    case B200_ACT_GELU_FAST: return launch_act<T, B200_ACT_GELU_FAST, GATED>(out, in, tokens, d, st);
    case B200_ACT_GELU_QUICK: return launch_act<T, B200_ACT_GELU_QUICK, GATED>(out, in, tokens, d, st);
  }
  set_error("activation: unknown activation id %d", act);
  return B200_EINVAL;
}

template <bool GATED>
int act_entry(void* out, const void* in, int tokens, int d, int act, int dtype, void* stream) {
  B200_REQUIRE(tokens >= 0 && d >= 0, B200_EINVAL, "activation: negative size");
  if (tokens == 0 || d == 0) return B200_OK;
  B200_REQUIRE(out && in, B200_EINVAL, "activation: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return dispatch_act<__half, GATED>(out, in, tokens, d, act, st);
    case B200_BF16: return dispatch_act<__nv_bfloat16, GATED>(out, in, tokens, d, act, st);
    case B200_F32: return dispatch_act<float, GATED>(out, in, tokens, d, act, st);
  }
  set_error("activation: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

template <typename T, typename CacheT, int KV>
int launch_rotary_cache(const b200_rotary_cache_params* p, cudaStream_t st) {
  constexpr int N = V16<T>::N;
  constexpr int X = 16 / (int)sizeof(CacheT);
  const int embed = p->rot_dim / 2;
  const int items = (p->num_heads + p->num_kv_heads) * (p->is_neox ? embed / N : embed / (N / 2)) +
                    p->num_kv_heads * ((p->head_size - p->rot_dim) / N + p->head_size / N);
  int threads = ((items + 31) / 32) * 32;
  threads = threads > 512 ? 512 : (threads < 32 ? 32 : threads);
#define RC_LAUNCH(NEOX)                                                                           \
  launch_pdl(rotary_cache_kernel<T, CacheT, KV, NEOX>, dim3(p->num_tokens), dim3(threads), 0, st,  \
             p->positions, (T*)p->query, (T*)p->key, (const T*)p->value,                         \
             (const T*)p->cos_sin_cache, p->rot_dim, p->query_stride, p->key_stride,             \
             p->value_stride, p->num_heads, p->num_kv_heads, p->head_size, (CacheT*)p->key_cache, \
             (CacheT*)p->value_cache, p->slot_mapping, p->block_size, p->kv_scale)
  if (p->is_neox) RC_LAUNCH(true); else RC_LAUNCH(false);
#undef RC_LAUNCH
  return check_launch("rotary_embedding_and_cache");
}

template <typename T>
int rotary_cache_dtype(const b200_rotary_cache_params* p, cudaStream_t st) {
  RotArgs a{p->positions, p->query, p->key, p->cos_sin_cache, nullptr, p->num_tokens, p->rot_dim,
            p->num_heads, p->num_kv_heads, p->head_size, p->query_stride, p->key_stride,
            p->is_neox != 0};
  const int X = p->kv_dtype == B200_KV_AUTO ? 16 / (int)sizeof(T) : 16;
  B200_REQUIRE(rotary_vec_ok<T>(a) && p->head_size % X == 0 && p->x == X &&
                   (p->value_stride * sizeof(T)) % 16 == 0 && aligned16(p->value) &&
                   aligned16(p->key_cache),
               B200_ENOTSUP,
               "rotary_embedding_and_cache: needs 16-byte aligned rows and x == 16/sizeof(cache "
               "element); call rotary_embedding + reshape_and_cache instead");
  switch (p->kv_dtype) {
    case B200_KV_AUTO: return launch_rotary_cache<T, T, B200_KV_AUTO>(p, st);
    case B200_KV_FP8_E4M3: return launch_rotary_cache<T, uint8_t, B200_KV_FP8_E4M3>(p, st);
    case B200_KV_FP8_E5M2: return launch_rotary_cache<T, uint8_t, B200_KV_FP8_E5M2>(p, st);
  }
  set_error("rotary_embedding_and_cache: unsupported kv dtype %d", p->kv_dtype);
  return B200_EINVAL;
}

}  // namespace
}  // namespace b200

extern "C" {

int b200_rms_norm(void* out, const void* input, const void* weight, float epsilon,
                  int32_t num_tokens, int32_t hidden, int32_t dtype, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0 && hidden >= 0, B200_EINVAL, "rms_norm: negative size");
  if (num_tokens == 0 || hidden == 0) return B200_OK;  // empty batch: pointers may be null
  B200_REQUIRE(out && input && weight, B200_EINVAL, "rms_norm: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return launch_rms<__half, false>(out, input, nullptr, weight, epsilon, num_tokens, hidden, st);
    case B200_BF16: return launch_rms<__nv_bfloat16, false>(out, input, nullptr, weight, epsilon, num_tokens, hidden, st);
    case B200_F32: return launch_rms<float, false>(out, input, nullptr, weight, epsilon, num_tokens, hidden, st);
  }
  set_error("rms_norm: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, float epsilon,
                            int32_t num_tokens, int32_t hidden, int32_t dtype, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0 && hidden >= 0, B200_EINVAL, "fused_add_rms_norm: negative size");
  if (num_tokens == 0 || hidden == 0) return B200_OK;
  B200_REQUIRE(input && residual && weight, B200_EINVAL, "fused_add_rms_norm: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return launch_rms<__half, true>(input, input, residual, weight, epsilon, num_tokens, hidden, st);
    case B200_BF16: return launch_rms<__nv_bfloat16, true>(input, input, residual, weight, epsilon, num_tokens, hidden, st);
    case B200_F32: return launch_rms<float, true>(input, input, residual, weight, epsilon, num_tokens, hidden, st);
  }
  set_error("fused_add_rms_norm: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

int b200_rms_norm_quant(void* out, float* scales, const void* input, void* residual, const void* weight,
                        float epsilon, int32_t num_tokens, int32_t hidden, int32_t dtype, int32_t quant_mode,
                        void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0 && hidden >= 0, B200_EINVAL, "rms_norm_quant: negative size");
  if (num_tokens == 0 || hidden == 0) return B200_OK;
  B200_REQUIRE(out && scales && input && weight, B200_EINVAL, "rms_norm_quant: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* o = static_cast<uint8_t*>(out);
  if (residual != nullptr) {
    if (dtype == B200_F16)
      return launch_rms_quant<__half, true>(o, scales, input, residual, weight, epsilon, num_tokens, hidden, quant_mode, st);
    if (dtype == B200_BF16)
      return launch_rms_quant<__nv_bfloat16, true>(o, scales, input, residual, weight, epsilon, num_tokens, hidden, quant_mode, st);
  } else {
    if (dtype == B200_F16)
      return launch_rms_quant<__half, false>(o, scales, input, nullptr, weight, epsilon, num_tokens, hidden, quant_mode, st);
    if (dtype == B200_BF16)
      return launch_rms_quant<__nv_bfloat16, false>(o, scales, input, nullptr, weight, epsilon, num_tokens, hidden, quant_mode, st);
  }
  set_error("rms_norm_quant: unsupported dtype %d (fp16 / bf16)", dtype);
  return B200_EINVAL;
}

int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                          const void* cos_sin_cache, const int64_t* cos_sin_cache_offsets,
                          int32_t num_tokens, int32_t rot_dim, int64_t query_stride,
                          int64_t key_stride, int32_t num_heads, int32_t num_kv_heads,
                          int32_t head_size, int32_t is_neox, int32_t dtype, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0 && rot_dim >= 0 && rot_dim % 2 == 0 && rot_dim <= head_size &&
                   num_heads >= 0 && num_kv_heads >= 0 && head_size > 0,
               B200_EINVAL, "rotary_embedding: bad shape (tokens=%d rot_dim=%d heads=%d/%d head_size=%d)",
               num_tokens, rot_dim, num_heads, num_kv_heads, head_size);
  if (num_tokens == 0 || rot_dim == 0 || num_heads + num_kv_heads == 0) return B200_OK;
  B200_REQUIRE(positions && query && key && cos_sin_cache, B200_EINVAL, "rotary_embedding: null pointer");
  RotArgs a{positions, query, key, cos_sin_cache, cos_sin_cache_offsets, num_tokens, rot_dim,
            num_heads, num_kv_heads, head_size, query_stride, key_stride, is_neox != 0};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return launch_rotary<__half>(a, st);
    case B200_BF16: return launch_rotary<__nv_bfloat16>(a, st);
    case B200_F32: return launch_rotary<float>(a, st);
  }
  set_error("rotary_embedding: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

int b200_rotary_embedding_and_cache(const b200_rotary_cache_params* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_EINVAL, "rotary_embedding_and_cache: null params");
  B200_REQUIRE(p->num_tokens >= 0 && p->rot_dim > 0 && p->rot_dim % 2 == 0 &&
                   p->rot_dim <= p->head_size && p->num_heads > 0 && p->num_kv_heads > 0 &&
                   p->head_size > 0 && p->block_size > 0,
               B200_EINVAL, "rotary_embedding_and_cache: bad shape");
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->positions && p->query && p->key && p->value && p->cos_sin_cache && p->key_cache &&
                   p->value_cache && p->slot_mapping,
               B200_EINVAL, "rotary_embedding_and_cache: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_F16: return rotary_cache_dtype<__half>(p, st);
    case B200_BF16: return rotary_cache_dtype<__nv_bfloat16>(p, st);
    case B200_F32: return rotary_cache_dtype<float>(p, st);
  }
  set_error("rotary_embedding_and_cache: unsupported dtype %d", p->dtype);
  return B200_EINVAL;
}

int b200_act_and_mul(void* out, const void* input, int32_t num_tokens, int32_t d, int32_t act,
                     int32_t dtype, void* stream) {
  return b200::act_entry<true>(out, input, num_tokens, d, act, dtype, stream);
}

int b200_activation(void* out, const void* input, int32_t num_tokens, int32_t d, int32_t act,
                    int32_t dtype, void* stream) {
  return b200::act_entry<false>(out, input, num_tokens, d, act, dtype, stream);
}

}  // extern "C"
