// gptq_marlin_repack: GPTQ [K/pack, N] int32 (packed along K) -> Marlin tiles [K/16, N*16/pack]
// int32, optional act-order row permutation.  Integer, bit-exact against marlin_weights()
// (vllm/model_executor/layers/quantization/utils/marlin_utils.py:40-58, marlin_perms.py:16-41).
// Replaces csrc/quantization/gptq_marlin/gptq_marlin_repack.cu:264-346.
//
// One CTA per (k-tile of 16 rows, group of 64 columns): the 16x64 tile of quantised values is
// gathered (through perm) into shared memory as bytes with coalesced row reads, then each thread
// assembles output words in the fragment order.  One-time work at weight load: 0.5 B/weight read
// + 0.5 B/weight written.
#include "common.cuh"

namespace b200 {
namespace {

// position p of an output word takes entry IL[p] of the (block, row) octet / quartet
__device__ __constant__ int8_t kIL4[8] = {0, 2, 4, 6, 1, 3, 5, 7};
__device__ __constant__ int8_t kIL8[4] = {0, 2, 1, 3};

template <int BITS>
__global__ void __launch_bounds__(256) marlin_repack_kernel(const uint32_t* __restrict__ src,
                                                            const int32_t* __restrict__ perm,
                                                            uint32_t* __restrict__ dst, int size_k,
                                                            int size_n) {
  constexpr int PACK = 32 / BITS;
  constexpr uint32_t MASK = (1u << BITS) - 1u;
  constexpr int WORDS = 16 * 64 / PACK;  // output words per 16x64 tile
  __shared__ uint8_t tile[16][64 + 4];
  const int kt = blockIdx.y;
  const int n0 = blockIdx.x * 64;
  // gather: 16 rows x 64 columns, thread -> (row, col) with col fastest (coalesced 256-B rows)
  for (int e = threadIdx.x; e < 16 * 64; e += blockDim.x) {
    const int r = e >> 6, c = e & 63;
    int k = kt * 16 + r;
    if (perm != nullptr) k = perm[k];
    const uint32_t w = src[(int64_t)(k / PACK) * size_n + n0 + c];
    tile[r][c] = (uint8_t)((w >> ((k % PACK) * BITS)) & MASK);
  }
  __syncthreads();
  uint32_t* out = dst + (int64_t)kt * (size_n * 16 / PACK) + (int64_t)blockIdx.x * WORDS;
  for (int r = threadIdx.x; r < WORDS; r += blockDim.x) {
    uint32_t word = 0;
#pragma unroll
    for (int p = 0; p < PACK; ++p) {
      int old;  // index into the un-interleaved perm list of the tile
      if constexpr (BITS == 4) old = r * 8 + kIL4[p];
      else old = r * 4 + kIL8[p];
      const int i = old >> 5, j = (old >> 3) & 3, t = old & 7;
      const int block = t >> 2, q = t & 3;
      const int row = 2 * (i & 3) + (q & 1) + ((q >> 1) << 3);
      const int col = 16 * j + (i >> 2) + 8 * block;
      word |= (uint32_t)tile[row][col] << (p * BITS);
    }
    out[r] = word;
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_gptq_marlin_repack(const uint32_t* b_q_weight, const int32_t* perm, uint32_t* out,
                                       int32_t size_k, int32_t size_n, int32_t num_bits,
                                       void* stream) {
  using namespace b200;
  // gptq_marlin_repack.cu:268-291
  B200_REQUIRE(num_bits == 4 || num_bits == 8, B200_EINVAL, "num_bits must be 4 or 8. Got = %d",
               num_bits);
  B200_REQUIRE(size_k % 16 == 0, B200_EINVAL, "size_k = %d is not divisible by tile_k_size = 16",
               size_k);
  B200_REQUIRE(size_n % 64 == 0, B200_EINVAL, "size_n = %d is not divisible by tile_n_size = 64",
               size_n);
  B200_REQUIRE(b_q_weight && out, B200_EINVAL, "gptq_marlin_repack: null pointer");
  if (size_k == 0 || size_n == 0) return B200_OK;
  dim3 grid(size_n / 64, size_k / 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (num_bits == 4)
    marlin_repack_kernel<4><<<grid, 256, 0, st>>>(b_q_weight, perm, out, size_k, size_n);
  else
    marlin_repack_kernel<8><<<grid, 256, 0, st>>>(b_q_weight, perm, out, size_k, size_n);
  return check_launch("gptq_marlin_repack");
}
