// Probe of tcgen05.mma.sp (kind::f16, 2:4 structured sparsity) on sm_100a: which TMEM lane / bit of
// the metadata steers which (row, k) of the sparse A operand.  The PTX ISA text is not available on
// this box, so the layout is read off the hardware: B is the identity, so D[m][k] is the value the
// tensor core placed at logical k of row m.
//
//   A (compressed)  [128 rows][64 fp16] in shared memory, K-major, 128-byte swizzle; value = 1 + column
//   B               [128 n][128 k] identity, two 64-k boxes, K-major, 128-byte swizzle
//   metadata        4 TMEM columns (one per MMA of logical K = 32) x 128 lanes
//   D               [128][128] fp32 in TMEM
// Experiment 0 uses 0x4444'4444 everywhere (every group of four keeps positions 0, 1); experiment
// 1 + 2 * lane + half flips one 16-bit half of one lane's word in metadata column `col` to 0xEEEE
// (positions 2, 3) -- the rows / k range of D that change identify what that half-word covers.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../neural_magic_vllm_b200/csrc -I../../include \
//        -o sparse_mma_probe.bin sparse_mma_probe.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "umma.cuh"

using namespace b200;
using namespace b200::umma;

constexpr int kM = 128, kN = 128, kK = 128;  // logical K; compressed K = 64

__device__ __forceinline__ uint32_t sw128_off(int row, int byte_in_row) {  // K-major SWIZZLE_128B tile
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((byte_in_row >> 4) ^ (row & 7)) << 4) | (byte_in_row & 15)));
}

__device__ __forceinline__ void mma_sp_ss_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t tmem_e,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.sp.cta_group::1.kind::f16 [%0], %1, %2, [%3], %4, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(tmem_e), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void mma_sp_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t tmem_e,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.sp.cta_group::1.kind::f16 [%0], [%1], %2, [%3], %4, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(tmem_e), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ts = 1: the compressed A operand comes from TMEM (row m in lane m, two fp16 per 32-bit column,
// 8 columns per MMA) instead of shared memory
__global__ void __launch_bounds__(128, 1) probe_kernel(float* out, int n_exp, int col, uint32_t base_word, uint32_t flip16,
                                                       int ts) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_s = smem;              // 128 x 128 B
  uint8_t* b_s = smem + 16384;      // 2 boxes of 128 x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  // operands
  for (int i = tid; i < kM * 64; i += 128) {
    const int r = i / 64, c = i % 64;
    *reinterpret_cast<__half*>(a_s + sw128_off(r, c * 2)) = __float2half((float)(1 + c));
  }
  for (int i = tid; i < kN * kK; i += 128) {
    const int n = i / kK, k = i % kK;
    *reinterpret_cast<__half*>(b_s + (k / 64) * 16384 + sw128_off(n, (k % 64) * 2)) = __float2half(n == k ? 1.f : 0.f);
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 256);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> async proxy (MMA)
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  constexpr uint32_t kECol = 128, kACol = 160;
  if (ts) {  // compressed A into TMEM: column j of lane m = values 1 + 2j, 2 + 2j
    for (int j = 0; j < 32; j += 4) {
      uint32_t w[4];
      for (int i = 0; i < 4; ++i) {
        const __half2 h = __floats2half2_rn((float)(1 + 2 * (j + i)), (float)(2 + 2 * (j + i)));
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
      }
      asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(lane_addr + kACol + j), "r"(w[0]),
                   "r"(w[1]), "r"(w[2]), "r"(w[3])
                   : "memory");
    }
    tmem_st_wait();
  }
  // M = 128, N = 128, F16 x F16 -> F32, K-major both, sparse flag (bit 2)
  const uint32_t idesc = instr_desc<KIND_F16>(kM, kN) | (1u << 2);
  for (int e = 0; e < n_exp; ++e) {
    uint32_t w[4] = {base_word, base_word, base_word, base_word};
    if (e > 0) {
      const int lane = (e - 1) >> 1, half = (e - 1) & 1;
      if (tid == lane) w[col] = half ? ((w[col] & 0x0000ffffu) | (flip16 << 16)) : ((w[col] & 0xffff0000u) | flip16);
    }
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(lane_addr + kECol), "r"(w[0]),
                 "r"(w[1]), "r"(w[2]), "r"(w[3])
                 : "memory");
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    if (tid == 0) {
      for (int kk = 0; kk < 4; ++kk) {  // logical K = 32 per MMA: 32 B of compressed A, 64 B of B per row
        const uint64_t da = desc_advance(smem_desc_sw128(smem_u32(a_s)), kk * 32);
        const uint64_t db = desc_advance(smem_desc_sw128(smem_u32(b_s) + (kk >> 1) * 16384), (kk & 1) * 64);
        // the metadata address is an even column; the odd one is selected through sparse_id2
        // (instruction-descriptor bits [0, 2)), as cute's MMA_Traits<..._SS_SPARSE> does
        const uint32_t e_col = kECol + kk;
        if (ts)
          mma_sp_ts_f16(tmem_base, tmem_base + kACol + kk * 8, db, tmem_base + (e_col & ~1u), idesc | (e_col & 1u),
                        kk > 0 ? 1u : 0u);
        else
          mma_sp_ss_f16(tmem_base, da, db, tmem_base + (e_col & ~1u), idesc | (e_col & 1u), kk > 0 ? 1u : 0u);
      }
      mma_commit(bar);
    }
    mbar_wait_bounded(bar, e & 1);
    fence_after_sync();
    for (int c = 0; c < kN / 16; ++c) {
      uint32_t r[16];
      tmem_ld16(lane_addr + c * 16, r);
      tmem_ld_wait();
      for (int j = 0; j < 16; ++j) out[((size_t)e * kM + tid) * kN + c * 16 + j] = __uint_as_float(r[j]);
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
  }
  if (warp == 0) tmem_dealloc(tmem_base, 256);
}

int main(int argc, char** argv) {
  const int col = argc > 1 ? atoi(argv[1]) : 0;
  const int ts = argc > 2 ? atoi(argv[2]) : 0;
  const int n_exp = 1 + 2 * 128;
  float* d_out;
  cudaMalloc(&d_out, (size_t)n_exp * kM * kN * 4);
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 32768 + 2048);
  probe_kernel<<<1, 128, 16384 + 32768 + 2048>>>(d_out, n_exp, col, 0x44444444u, 0xEEEEu, ts);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) {
    printf("error: %s\n", cudaGetErrorString(err));
    return 1;
  }
  std::vector<float> h((size_t)n_exp * kM * kN);
  cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost);
  // baseline: print rows 0, 1, 8 of D for k < 32
  for (int r : {0, 1, 8, 127}) {
    printf("base row %3d:", r);
    for (int k = 0; k < 40; ++k) printf(" %g", h[(size_t)r * kN + k]);
    printf("\n");
  }
  // expected baseline: D[m][4 g + {0, 1}] = 1 + 2 g + {0, 1}
  int bad = 0;
  for (int r = 0; r < kM; ++r)
    for (int k = 0; k < kK; ++k) {
      const int g = k / 4, p = k % 4;
      const float want = p < 2 ? (float)(1 + 2 * g + p) : 0.f;
      if (h[(size_t)r * kN + k] != want) ++bad;
    }
  printf("baseline mismatches vs (positions 0,1 kept): %d of %d\n", bad, kM * kK);
  // flips
  for (int e = 1; e < n_exp; ++e) {
    const int lane = (e - 1) >> 1, half = (e - 1) & 1;
    int rmin = 999, rmax = -1, kmin = 999, kmax = -1, cnt = 0;
    for (int r = 0; r < kM; ++r)
      for (int k = 0; k < kK; ++k)
        if (h[((size_t)e * kM + r) * kN + k] != h[(size_t)r * kN + k]) {
          ++cnt;
          rmin = r < rmin ? r : rmin, rmax = r > rmax ? r : rmax, kmin = k < kmin ? k : kmin, kmax = k > kmax ? k : kmax;
        }
    if (lane < 2 || lane == 8 || lane == 127)
      printf("{\"col\": %d, \"lane\": %d, \"half\": %d, \"changed\": %d, \"rows\": [%d, %d], \"k\": [%d, %d]}\n", col, lane, half,
             cnt, rmin, rmax, kmin, kmax);
  }
  // machine-readable full map
  FILE* f = fopen("gpurun_out/sparse_probe_map.txt", argc > 1 ? "a" : "w");
  if (f) {
    for (int e = 1; e < n_exp; ++e) {
      const int lane = (e - 1) >> 1, half = (e - 1) & 1;
      int rmin = 999, rmax = -1, kmin = 999, kmax = -1;
      for (int r = 0; r < kM; ++r)
        for (int k = 0; k < kK; ++k)
          if (h[((size_t)e * kM + r) * kN + k] != h[(size_t)r * kN + k])
            rmin = r < rmin ? r : rmin, rmax = r > rmax ? r : rmax, kmin = k < kmin ? k : kmin, kmax = k > kmax ? k : kmax;
      fprintf(f, "%d %d %d %d %d %d %d\n", col, lane, half, rmin, rmax, kmin, kmax);
    }
    fclose(f);
  }
  return 0;
}
