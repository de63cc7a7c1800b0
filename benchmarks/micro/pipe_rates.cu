// Issue-rate microbenchmark for the softmax inner loop's instruction mix on sm_100a: cycles per
// warp-instruction per SM sub-partition for MUFU.EX2, F2FP (fp32 pair -> bf16x2 / f16x2), FFMA2,
// FADD2, FMNMX3, and the MUFU + F2FP mix (do they share the XU pipe?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu && ./pipe_rates
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

constexpr int kIters = 256, kUnroll = 16;

template <int OP>
__global__ void rate_kernel(float* out, long long* cycles, float seed) {
  float x[kUnroll];
  uint32_t u[kUnroll];
#pragma unroll
  for (int i = 0; i < kUnroll; ++i) { x[i] = seed + threadIdx.x * 1e-3f + i; u[i] = i; }
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < kUnroll; ++i) {
      if constexpr (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if constexpr (OP == 1) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(x[(i + 1) % kUnroll]));
      if constexpr (OP == 2) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(x[(i + 1) % kUnroll]));
      if constexpr (OP == 3) asm volatile("{.reg .b64 a; mov.b64 a, {%0, %1}; fma.rn.ftz.f32x2 a, a, a, a; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[(i + 1) % kUnroll]));
      if constexpr (OP == 4) asm volatile("{.reg .b64 a; mov.b64 a, {%0, %1}; add.rn.ftz.f32x2 a, a, a; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[(i + 1) % kUnroll]));
      if constexpr (OP == 5) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(x[(i + 1) % kUnroll]), "f"(x[(i + 2) % kUnroll]));
      if constexpr (OP == 6) {  // MUFU + F2FP interleaved 2 : 1 (the softmax mix)
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        if (i & 1) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(x[i - 1]));
      }
      if constexpr (OP == 7) asm volatile("fma.rn.ftz.f32 %0, %0, %0, %0;" : "+f"(x[i]));
      if constexpr (OP == 8) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u[i]));
      if constexpr (OP == 9) {  // MUFU + FFMA2 + FADD2 + F2FP: the whole unmasked inner loop per pair
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        if (i & 1) {
          asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(x[i - 1]));
          asm volatile("{.reg .b64 a; mov.b64 a, {%0, %1}; fma.rn.ftz.f32x2 a, a, a, a; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[i - 1]));
          asm volatile("{.reg .b64 a; mov.b64 a, {%0, %1}; add.rn.ftz.f32x2 a, a, a; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[i - 1]));
        }
      }
    }
  }
  const long long t1 = clock64();
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < kUnroll; ++i) acc += x[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int insts_per_unroll_x16) {
  float* out; long long* cyc;
  cudaMalloc(&out, 1024 * sizeof(float)); cudaMalloc(&cyc, sizeof(long long));
  for (int warps_per_smsp = 1; warps_per_smsp <= 2; ++warps_per_smsp) {
    const int threads = 128 * warps_per_smsp;
    rate_kernel<OP><<<1, threads>>>(out, cyc, 0.5f);
    rate_kernel<OP><<<1, threads>>>(out, cyc, 0.5f);
    long long c; cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double n = (double)kIters * insts_per_unroll_x16;  // warp-instructions per warp
    printf("{\"op\": \"%s\", \"warps_per_smsp\": %d, \"cycles_per_warp_inst_per_smsp\": %.3f}\n", name, warps_per_smsp,
           (double)c / (n * warps_per_smsp));
  }
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("MUFU.EX2", 16);
  run<1>("F2FP.BF16.PACK_AB", 16);
  run<2>("F2FP.F16.PACK_AB", 16);
  run<3>("FFMA2", 16);
  run<4>("FADD2", 16);
  run<5>("FMNMX3", 16);
  run<7>("FFMA", 16);
  run<8>("MUFU.EX2.F16x2", 16);
  run<6>("mix 2 MUFU + 1 F2FP (per 3 inst)", 24);
  run<9>("mix 2 MUFU + F2FP + FFMA2 + FADD2 (per 5 inst)", 40);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
