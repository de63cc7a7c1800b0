// Activation quantisation feeding the W8A8 GEMM (sm_100a).
//   int8: per-token dynamic (absmax/127) or per-tensor static   -- replaces
//         csrc/quantization/compressed_tensors/int8_quant_kernels.cu:27-115
//   fp8 : per-tensor e4m3, dynamic (absmax/448, two launches) or static -- replaces
//         csrc/quantization/fp8/common.cu:27-165
// HBM-bound element-wise work: 16-byte loads, one row per CTA (int8) / grid-stride (fp8).
// Rounding follows the reference exactly: int8 = cvt.rni.sat(x / scale) (static) or
// cvt.rni.sat(x * (127 / absmax)) (dynamic); fp8 = e4m3(clamp(x * (1 / scale), +-448)).
#include "common.cuh"

namespace b200 {
namespace {

__device__ __forceinline__ int8_t f2i8(float x) {
  int32_t d;
  asm("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(d) : "f"(x));
  return static_cast<int8_t>(d);
}

template <typename T>
struct Vec8 {  // 8 elements of T
  T v[8];
};

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&f)[8]) {
  if constexpr (sizeof(T) == 2) {
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    const T* h = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = DT<T>::to_f(h[i]);
  } else {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
  }
}

__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = (l < nw) ? red[l] : 0.f;
  r = warp_max(r);
  __syncthreads();
  return r;
}

// one CTA per token
template <typename T, bool DYNAMIC>
__global__ void __launch_bounds__(256) int8_quant_kernel(int8_t* __restrict__ out,
                                                         const T* __restrict__ in,
                                                         float* __restrict__ scales, int hidden) {
  __shared__ float red[32];
  const int64_t row = blockIdx.x;
  const T* x = in + row * hidden;
  int8_t* o = out + row * hidden;
  const bool vec = (hidden % 8 == 0) && ((reinterpret_cast<uintptr_t>(in) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(out) & 7) == 0);
  float mul, div;
  if constexpr (DYNAMIC) {
    float amax = 0.f;
    if (vec) {
      for (int i = threadIdx.x * 8; i < hidden; i += blockDim.x * 8) {
        float f[8];
        load8(x + i, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(f[j]));
      }
    } else {
      for (int i = threadIdx.x; i < hidden; i += blockDim.x) amax = fmaxf(amax, fabsf(DT<T>::to_f(x[i])));
    }
    amax = block_max(amax, red);
    if (threadIdx.x == 0) scales[row] = amax / 127.0f;
    mul = 127.0f / amax;
    div = 1.f;
  } else {
    div = *scales;
    mul = 1.f;
  }
  if (vec) {
    for (int i = threadIdx.x * 8; i < hidden; i += blockDim.x * 8) {
      float f[8];
      load8(x + i, f);
      int8_t q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) q[j] = DYNAMIC ? f2i8(f[j] * mul) : f2i8(f[j] / div);
      *reinterpret_cast<uint2*>(o + i) = *reinterpret_cast<uint2*>(q);
    }
  } else {
    for (int i = threadIdx.x; i < hidden; i += blockDim.x) {
      const float f = DT<T>::to_f(x[i]);
      o[i] = DYNAMIC ? f2i8(f * mul) : f2i8(f / div);
    }
  }
}

// ---- fp8 per-tensor ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(512) fp8_absmax_kernel(float* __restrict__ scale,
                                                         const T* __restrict__ in, int64_t n) {
  __shared__ float red[32];
  float amax = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
  const int64_t nv = (reinterpret_cast<uintptr_t>(in) & 15) == 0 ? (n / 8) * 8 : 0;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < nv; i += stride) {
    float f[8];
    load8(in + i, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(f[j]));
  }
  for (int64_t i = nv + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    amax = fmaxf(amax, fabsf(DT<T>::to_f(in[i])));
  amax = block_max(amax, red);
  // non-negative floats order like ints (fp8/common.cu:8-14)
  if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(scale), __float_as_int(amax / 448.0f));
}

__device__ __forceinline__ uint8_t to_e4m3(float x, float inv) {
  float r = fmaxf(-448.0f, fminf(x * inv, 448.0f));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}

template <typename T>
__global__ void __launch_bounds__(512) fp8_quant_kernel(uint8_t* __restrict__ out,
                                                        const T* __restrict__ in,
                                                        const float* __restrict__ scale, int64_t n) {
  const float inv = 1.0f / (*scale);
  const bool al = ((reinterpret_cast<uintptr_t>(in) & 15) == 0) && ((reinterpret_cast<uintptr_t>(out) & 7) == 0);
  const int64_t nv = al ? (n / 8) * 8 : 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < nv; i += stride) {
    float f[8];
    load8(in + i, f);
    uint8_t q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] = to_e4m3(f[j], inv);
    *reinterpret_cast<uint2*>(out + i) = *reinterpret_cast<uint2*>(q);
  }
  for (int64_t i = nv + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = to_e4m3(DT<T>::to_f(in[i]), inv);
}

template <typename T>
int launch_int8(int8_t* out, const void* in, float* scales, int tokens, int hidden, bool dynamic,
                cudaStream_t st) {
  if (tokens == 0 || hidden == 0) return B200_OK;
  if (dynamic)
    int8_quant_kernel<T, true><<<tokens, 256, 0, st>>>(out, (const T*)in, scales, hidden);
  else
    int8_quant_kernel<T, false><<<tokens, 256, 0, st>>>(out, (const T*)in, scales, hidden);
  return check_launch("scaled_int8_quant");
}

template <typename T>
int launch_fp8(uint8_t* out, const void* in, float* scale, int64_t n, bool dynamic, cudaStream_t st) {
  if (n == 0) return B200_OK;
  int64_t want = (n + 512 * 8 - 1) / (512 * 8);
  int grid = (int)(want < 1 ? 1 : (want > num_sms() * 4 ? num_sms() * 4 : want));
  if (dynamic) fp8_absmax_kernel<T><<<grid, 512, 0, st>>>(scale, (const T*)in, n);
  fp8_quant_kernel<T><<<grid, 512, 0, st>>>(out, (const T*)in, scale, n);
  return check_launch("scaled_fp8_quant");
}

}  // namespace
}  // namespace b200

extern "C" {

int b200_scaled_int8_quant(int8_t* out, const void* input, float* scales, int32_t num_tokens,
                           int32_t hidden, int32_t dtype, int32_t dynamic, void* stream) {
  using namespace b200;
  B200_REQUIRE(out && input && scales, B200_EINVAL, "scaled_int8_quant: null pointer");
  B200_REQUIRE(num_tokens >= 0 && hidden >= 0, B200_EINVAL, "scaled_int8_quant: negative size");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return launch_int8<__half>(out, input, scales, num_tokens, hidden, dynamic != 0, st);
    case B200_BF16: return launch_int8<__nv_bfloat16>(out, input, scales, num_tokens, hidden, dynamic != 0, st);
    case B200_F32: return launch_int8<float>(out, input, scales, num_tokens, hidden, dynamic != 0, st);
  }
  set_error("scaled_int8_quant: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

int b200_scaled_fp8_quant(uint8_t* out, const void* input, float* scale, int64_t numel,
                          int32_t dtype, int32_t dynamic, void* stream) {
  using namespace b200;
  B200_REQUIRE(out && input && scale, B200_EINVAL, "scaled_fp8_quant: null pointer");
  B200_REQUIRE(numel >= 0, B200_EINVAL, "scaled_fp8_quant: negative size");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F16: return launch_fp8<__half>(out, input, scale, numel, dynamic != 0, st);
    case B200_BF16: return launch_fp8<__nv_bfloat16>(out, input, scale, numel, dynamic != 0, st);
    case B200_F32: return launch_fp8<float>(out, input, scale, numel, dynamic != 0, st);
  }
  set_error("scaled_fp8_quant: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

}  // extern "C"
