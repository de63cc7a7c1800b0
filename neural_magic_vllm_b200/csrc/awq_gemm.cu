// awq_gemm / awq_dequantize on tcgen05 (sm_100a) -- SURVEY.md 8(f) rank 4.  Replaces
// csrc/quantization/awq/gemm_kernels.cu:18-549 (mma.sync m16n8k16 kernel + dequantize_weights) and
// consumes the AWQ checkpoint layout as it is:
//   qweight [K, N / 8] int32   one word = eight 4-bit values of ONE k row; nibble j holds output
//                              channel 8 c + {0, 2, 4, 6, 1, 3, 5, 7}[j] (dequantize.cuh:17-101)
//   qzeros  [K / G, N / 8] int32   same packing, one row per group of G input channels
//   scales  [K / G, N] fp16
//   w[k, n] = (q - z) * s, with (q - z) exact and ONE fp16 rounding (gemm_kernels.cu:402-431).
// fp16 only, like the reference.
//
// The AWQ word order makes eight CONSECUTIVE output channels of one k row fall out of one word:
// the dequantised weights are naturally n-contiguous.  So they go to shared memory as an MN-major
// A operand (rows = k, 128 bytes = 64 channels, 128-byte swizzle) with one 16-byte store per word,
// and tcgen05.mma reads them from there (D[n, token] = W^T[n, k] X^T[k, token]); no transpose, no
// repack into the Marlin layout.
//   warp 0     TMA producer: per 64-k stage the activation tile [BM x 64] (K-major, 128B swizzle),
//              the packed words [64 x 32] (256 channels: a 128-byte line per k row), and per 32-row
//              half the group's scale row and zero words
//   warp 1     MMA issuer (elected lane): per stage four K = 16 steps x two 128-channel sub-tiles
//              (kind::f16, fp32 D in TMEM)
//   warps 2-9  dequant: LDS.32 word -> lop3 magic numbers -> (q - z) exactly (hsub2 / hfma2 against
//              per-column constants built from the zero word) -> x scale -> STS.128; then the
//              epilogue (TMEM lane = output channel)
// Small M splits K over blockIdx.z with fp32 partials and a sum kernel.
#include "marlin_dequant.cuh"
#include "umma.cuh"

namespace b200 {
namespace {

using namespace umma;
using namespace marlin;

constexpr int kAwqTileN = 256;  // channels per CTA: one packed row segment = 128 bytes (a full line per k row)
constexpr int kAwqSubN = 128;   // MMA M: two sub-tiles per CTA
constexpr int kAwqSubs = kAwqTileN / kAwqSubN;
constexpr int kAwqStageK = 64;
constexpr int kAwqStages = 3;   // weight ring; with 2 A tiles and 6 activation tiles ~107 KiB per CTA at BM = 16: two CTAs per SM
constexpr int kAwqAStages = 2;
constexpr int kAwqDequantWarps = 8;
constexpr int kAwqThreads = 64 + kAwqDequantWarps * 32;
constexpr int kAwqATileBytes = kAwqStageK * kAwqTileN * 2;  // 32 KiB: four 64-channel column blocks of 8 KiB
constexpr int kAwqABlockBytes = kAwqStageK * 128;
constexpr int kAwqQBytes = kAwqStageK * (kAwqTileN / 8) * 4;  // 8 KiB of packed words
constexpr int kAwqSBytes = 2 * kAwqTileN * 2;                 // scale rows of the two 32-row halves
constexpr int kAwqZSlot = 128;                               // TMA destinations are 128-byte aligned
constexpr int kAwqZBytes = 2 * kAwqZSlot;                     // zero words (128 B) of the two halves
static_assert((kAwqTileN / 8) * 4 <= kAwqZSlot, "zero words of a tile fit one slot");

struct AwqArgs {
  void* c;
  float* partial;
  int m, n, k;
  int group_size;
  int kb_per_split, num_kb;
};

template <int BM>
struct PlanAwq {
  static constexpr int kXBytes = BM * 128;
  // two rings: the packed words / scales / zeros are free again as soon as the dequant warps hold
  // them in registers, the activation tile only once the MMAs that read it have completed
  static constexpr int kXStages = BM <= 16 ? 6 : 4;
  static constexpr int kQStageBytes = (kAwqQBytes + kAwqSBytes + kAwqZBytes + 1023) / 1024 * 1024;
  static constexpr int kBytes = kXStages * kXBytes + kAwqStages * kQStageBytes + kAwqAStages * kAwqATileBytes + 1024 + 512;
  static constexpr uint32_t kTmemCols = kAwqSubs * BM < 32 ? 32 : kAwqSubs * BM;
  static constexpr int kCtasPerSm = BM <= 32 ? 2 : 1;
};

// MN-major operand tile: rows = k, 128 bytes = 64 elements of the M / N dimension per row, 8-row
// swizzle atoms 1 KiB apart (SBO), 64-element column blocks `lbo_bytes` apart, SWIZZLE_128B
// (cute/atom/mma_traits_sm100.hpp canonical ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units)
__device__ __forceinline__ uint64_t awq_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((1024u >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Per-column constants from a zero word: zc[0], zc[2] = 1024 + z (subtracted from 1024 + q),
// zc[1], zc[3] = -(64 + z) (added to (1024 + 16 q) / 16): both give q - z exactly.
__device__ __forceinline__ void awq_zero_consts(uint32_t zq, uint32_t (&zc)[4]) {
  constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400;
  constexpr uint32_t NEG16TH = 0xac00ac00;  // -1/16
  const uint32_t z8 = zq >> 8;
  zc[0] = lop3<kAndOr>(zq, LO, EX);
  zc[1] = h2_fma(lop3<kAndOr>(zq, HI, EX), NEG16TH, 0u);
  zc[2] = lop3<kAndOr>(z8, LO, EX);
  zc[3] = h2_fma(lop3<kAndOr>(z8, HI, EX), NEG16TH, 0u);
}
// one packed word -> eight fp16 weights of channels 8 c .. 8 c + 7 (in order), (q - z) * s
__device__ __forceinline__ uint4 awq_dequant_word(uint32_t q, const uint32_t (&zc)[4], const uint4& s) {
  constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400, MUL = 0x2c002c00;  // 1/16
  const uint32_t q8 = q >> 8;
  uint4 r;
  r.x = mul2<__half>(h2_sub(lop3<kAndOr>(q, LO, EX), zc[0]), s.x);
  r.y = mul2<__half>(h2_fma(lop3<kAndOr>(q, HI, EX), MUL, zc[1]), s.y);
  r.z = mul2<__half>(h2_sub(lop3<kAndOr>(q8, LO, EX), zc[2]), s.z);
  r.w = mul2<__half>(h2_fma(lop3<kAndOr>(q8, HI, EX), MUL, zc[3]), s.w);
  return r;
}

// awq_dequantize: one thread per packed word (gemm_kernels.cu:369-435 dequantize_weights)
__global__ void __launch_bounds__(256) awq_dequantize_kernel(__half* __restrict__ out, const uint32_t* __restrict__ qweight,
                                                             const __half* __restrict__ scales,
                                                             const uint32_t* __restrict__ qzeros, int k, int n8,
                                                             int group_size) {
  const int64_t total = (int64_t)k * n8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / n8), c = (int)(i % n8);
    const int g = row / group_size;
    uint32_t zc[4];
    awq_zero_consts(qzeros[(int64_t)g * n8 + c], zc);
    const uint4 s = *reinterpret_cast<const uint4*>(scales + ((int64_t)g * n8 + c) * 8);
    *reinterpret_cast<uint4*>(out + i * 8) = awq_dequant_word(qweight[i], zc, s);
  }
}

template <int BM>
__global__ void __launch_bounds__(kAwqThreads, PlanAwq<BM>::kCtasPerSm)
awq_gemm_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_q,
                const __grid_constant__ CUtensorMap tm_s, const __grid_constant__ CUtensorMap tm_z,
                const AwqArgs args) {
  using P = PlanAwq<BM>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* x_ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stages = x_ring + P::kXStages * P::kXBytes;     // packed words + scales + zeros
  uint8_t* a_ring = stages + kAwqStages * P::kQStageBytes;  // dequantised weights, MN-major
  uint64_t* bars = reinterpret_cast<uint64_t*>(a_ring + kAwqAStages * kAwqATileBytes);
  uint64_t* full_p = bars;                       // [kAwqStages] weights landed
  uint64_t* empty_p = full_p + kAwqStages;       // [kAwqStages] weights in registers
  uint64_t* full_x = empty_p + kAwqStages;       // [kXStages] activations landed
  uint64_t* empty_x = full_x + P::kXStages;      // [kXStages] MMAs that read them completed
  uint64_t* full_a = empty_x + P::kXStages;
  uint64_t* empty_a = full_a + kAwqAStages;
  uint64_t* tmem_full = empty_a + kAwqAStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * kAwqTileN;
  const int split = blockIdx.z;
  const int kb_begin = split * args.kb_per_split;
  const int kb_end = min(kb_begin + args.kb_per_split, args.num_kb);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_s);
    tma_prefetch_desc(&tm_z);
    for (int s = 0; s < kAwqStages; ++s) {
      mbar_init(&full_p[s], 1);
      mbar_init(&empty_p[s], kAwqDequantWarps);
    }
    for (int s = 0; s < P::kXStages; ++s) {
      mbar_init(&full_x[s], 1);
      mbar_init(&empty_x[s], 1);
    }
    for (int s = 0; s < kAwqAStages; ++s) {
      mbar_init(&full_a[s], kAwqDequantWarps);
      mbar_init(&empty_a[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, P::kTmemCols);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ======================= TMA producer =======================
    const bool leader = elect_one();
    const uint64_t pol_w = l2_policy_evict_first();  // streamed once (per m tile)
    const uint64_t pol_x = l2_policy_evict_last();
    constexpr uint32_t tx_bytes = kAwqQBytes + kAwqSBytes + 2 * (kAwqTileN / 8) * 4;
    auto load_weights = [&](int stage, int kb) {
      uint8_t* q_dst = stages + stage * P::kQStageBytes;
      mbar_arrive_expect_tx(&full_p[stage], tx_bytes);
      tma_load_2d(q_dst, &tm_q, n0 / 8, kb * kAwqStageK, &full_p[stage], pol_w);
#pragma unroll
      for (int h = 0; h < 2; ++h) {  // G % 32 == 0: a 32-row half never straddles two groups
        const int g = (kb * kAwqStageK + 32 * h) / args.group_size;
        tma_load_2d(q_dst + kAwqQBytes + h * (kAwqTileN * 2), &tm_s, n0, g, &full_p[stage], pol_x);
        tma_load_2d(q_dst + kAwqQBytes + kAwqSBytes + h * kAwqZSlot, &tm_z, n0 / 8, g, &full_p[stage], pol_x);
      }
    };
    auto load_x = [&](int xs, int kb) {
      mbar_arrive_expect_tx(&full_x[xs], P::kXBytes);
      tma_load_2d(x_ring + xs * P::kXBytes, &tm_x, kb * kAwqStageK, m0, &full_x[xs], pol_x);
    };
    // programmatic dependent launch: weights, scales and zeros of the first ring of stages stream
    // in before the wait; only the activation tiles depend on the predecessor
    const int n_kb = kb_end - kb_begin;
    const int pre = min(n_kb, kAwqStages), pre_x = min(n_kb, P::kXStages);
    if (leader)
      for (int i = 0; i < pre; ++i) load_weights(i, kb_begin + i);
    pdl_wait();
    if (leader)
      for (int i = 0; i < pre_x; ++i) load_x(i, kb_begin + i);
    __syncwarp();
    // weights run `pre` steps ahead, activations `pre_x`: one loop, each ring with its own cursor
    int stage = pre == kAwqStages ? 0 : pre, xs = pre_x == P::kXStages ? 0 : pre_x;
    uint32_t phase = pre == kAwqStages ? 1 : 0, xphase = pre_x == P::kXStages ? 1 : 0;
    for (int i = pre; i < n_kb + (pre_x - pre); ++i) {
      if (i < n_kb) {
        mbar_wait_bounded(&empty_p[stage], phase ^ 1);
        if (leader) load_weights(stage, kb_begin + i);
        if (++stage == kAwqStages) stage = 0, phase ^= 1;
      }
      const int ix = i + (pre_x - pre);  // the activation step this iteration may issue
      if (ix >= pre_x && ix < n_kb) {
        mbar_wait_bounded(&empty_x[xs], xphase ^ 1);
        if (leader) load_x(xs, kb_begin + ix);
        if (++xs == P::kXStages) xs = 0, xphase ^= 1;
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    const bool leader = elect_one();
    // D_sub[128 channels, BM tokens]; A = weights, MN-major (bit 15); B = activations, K-major
    constexpr uint32_t idesc = instr_desc<KIND_F16>(kAwqSubN, BM) | (1u << 15);
    int xs = 0, as = 0;
    uint32_t aphase = 0, xphase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait_bounded(&full_a[as], aphase);
      mbar_wait_bounded(&full_x[xs], xphase);
      fence_after_sync();
      const uint64_t dx = smem_desc_sw128(smem_u32(x_ring + xs * P::kXBytes));
      const uint32_t aa = smem_u32(a_ring + as * kAwqATileBytes);
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < kAwqStageK / 16; ++kk)  // 16 k rows = two 8-row atoms of 1 KiB
#pragma unroll
          for (int sub = 0; sub < kAwqSubs; ++sub)
            mma_ss<KIND_F16>(tmem_base + sub * BM,
                             awq_desc_mn_sw128(aa + sub * 2 * kAwqABlockBytes + kk * 2048, kAwqABlockBytes),
                             desc_advance(dx, kk * 32), idesc, (kb > kb_begin || kk > 0) ? 1u : 0u);
        mma_commit(&empty_a[as]);
        mma_commit(&empty_x[xs]);
      }
      if (++xs == P::kXStages) xs = 0, xphase ^= 1;
      if (++as == kAwqAStages) as = 0, aphase ^= 1;
    }
    if (leader) mma_commit(tmem_full);
    pdl_launch_dependents();
  } else {
    // ======================= dequant, then epilogue =======================
    const int qd = warp & 3;                 // TMEM quadrant of the epilogue
    const int sub = (warp - 2) >> 2;         // epilogue: 128-channel sub-tile
    const int t = (warp - 2) * 32 + lane;    // 0 .. 255
    const int c = t & 31;                    // word column: channels n0 + 8 c .. + 7
    const int r0 = t >> 5;                   // rows r0 + 8 i
    const int blk = c >> 3, chunk = c & 7;   // 64-channel column block, 16-byte chunk inside its 128-byte row
    const uint32_t stages_u32 = smem_u32(stages), a_ring_u32 = smem_u32(a_ring);
    int stage = 0, as = 0;
    uint32_t phase = 0, aphase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait_bounded(&full_p[stage], phase);
      const uint32_t q_src = stages_u32 + stage * P::kQStageBytes;  // shared-space address
      uint32_t words[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) words[i] = lds32(q_src + (r0 + 8 * i) * (kAwqTileN / 2) + c * 4);
      uint4 sc[2];
      uint32_t zq[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        sc[h] = lds128(q_src + kAwqQBytes + h * (kAwqTileN * 2) + c * 16);
        zq[h] = lds32(q_src + kAwqQBytes + kAwqSBytes + h * kAwqZSlot + c * 4);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_p[stage]);
      mbar_wait_bounded(&empty_a[as], aphase ^ 1);
      const uint32_t a_dst = a_ring_u32 + as * kAwqATileBytes + blk * kAwqABlockBytes + r0 * 128 + ((chunk ^ r0) << 4);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t zc[4];
        awq_zero_consts(zq[h], zc);
#pragma unroll
        for (int i = 4 * h; i < 4 * h + 4; ++i)  // row r0 + 8 i: 8-row atom i, row r0 inside it
          sts128(a_dst + i * 1024, awq_dequant_word(words[i], zc, sc[h]));
      }
      // generic-proxy stores -> visible to the tensor core (async proxy) before the arrival
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[as]);
      if (++stage == kAwqStages) stage = 0, phase ^= 1;
      if (++as == kAwqAStages) as = 0, aphase ^= 1;
    }
    // epilogue: TMEM lane = output channel, column = token
    const int n = n0 + sub * kAwqSubN + qd * 32 + lane;
    mbar_wait_bounded(tmem_full, 0);
    fence_after_sync();
    const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + sub * BM;
#pragma unroll 1
    for (int cc = 0; cc < BM / 16; ++cc) {
      if (m0 + cc * 16 >= args.m) break;
      uint32_t rr[16];
      tmem_ld16(taddr + cc * 16, rr);
      tmem_ld_wait();
      if (n < args.n) {
        if (args.partial != nullptr) {
          float* ws = args.partial + ((int64_t)split * args.m) * args.n;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int mm = m0 + cc * 16 + j;
            if (mm < args.m) ws[(int64_t)mm * args.n + n] = __uint_as_float(rr[j]);
          }
        } else {
          __half* out = reinterpret_cast<__half*>(args.c);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int mm = m0 + cc * 16 + j;
            if (mm < args.m) out[(int64_t)mm * args.n + n] = __float2half_rn(__uint_as_float(rr[j]));
          }
        }
      }
    }
    fence_before_sync();
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, P::kTmemCols);
  }
}

__global__ void __launch_bounds__(256) awq_splitk_sum_kernel(__half* __restrict__ out, const float* __restrict__ ws,
                                                             int64_t total, int splits) {
  pdl_wait();
  pdl_launch_dependents();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += ws[(int64_t)s * total + i];  // fixed order: deterministic
    out[i] = __float2half_rn(acc);
  }
}

template <int BM>
int launch_awq(const b200_awq_gemm_params& p, cudaStream_t st) {
  using P = PlanAwq<BM>;
  const int groups = p.size_k / p.group_size;
  CUtensorMap tm_x, tm_q, tm_s, tm_z;
  int rc = make_tmap_2d(&tm_x, p.a, (uint64_t)p.size_m, (uint64_t)p.size_k, (uint64_t)p.lda * 2, 2, BM);
  if (rc != B200_OK) return rc;
  const uint64_t n8 = (uint64_t)p.size_n / 8;
  rc = make_tmap_2d_ex(&tm_q, p.qweight, (uint64_t)p.size_k, n8, n8 * 4, CU_TENSOR_MAP_DATA_TYPE_INT32, kAwqTileN / 8,
                       kAwqStageK, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_s, p.scales, (uint64_t)groups, (uint64_t)p.size_n, (uint64_t)p.size_n * 2,
                       CU_TENSOR_MAP_DATA_TYPE_UINT16, kAwqTileN, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_z, p.qzeros, (uint64_t)groups, n8, n8 * 4, CU_TENSOR_MAP_DATA_TYPE_INT32, kAwqTileN / 8, 1,
                       CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  AwqArgs a{};
  a.c = p.c;
  a.m = p.size_m, a.n = p.size_n, a.k = p.size_k;
  a.group_size = p.group_size;
  a.num_kb = (p.size_k + kAwqStageK - 1) / kAwqStageK;
  const int m_tiles = (p.size_m + BM - 1) / BM, n_tiles = (p.size_n + kAwqTileN - 1) / kAwqTileN;
  // weight streaming at small M: split K until every resident-CTA slot of the GPU is taken
  const int tiles = m_tiles * n_tiles;
  const int slots = P::kCtasPerSm * num_sms();
  int splits = tiles >= slots ? 1 : slots / tiles;
  if (splits > a.num_kb / 4) splits = a.num_kb / 4;
  const int64_t per = (int64_t)p.size_m * p.size_n * 4;
  if (p.splitk_ws == nullptr || per <= 0) splits = 1;
  else if (splits > p.splitk_ws_bytes / per) splits = (int)(p.splitk_ws_bytes / per);
  if (splits < 1) splits = 1;
  a.kb_per_split = (a.num_kb + splits - 1) / splits;
  splits = (a.num_kb + a.kb_per_split - 1) / a.kb_per_split;
  a.partial = splits > 1 ? reinterpret_cast<float*>(p.splitk_ws) : nullptr;
  auto kern = awq_gemm_kernel<BM>;
  static size_t smem_configured[kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, smem_configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("awq_gemm: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  cudaError_t le = launch_pdl(kern, dim3(m_tiles, n_tiles, splits), dim3(kAwqThreads), (size_t)P::kBytes, st, tm_x, tm_q,
                              tm_s, tm_z, a);
  if (le != cudaSuccess) {
    cudaGetLastError();
    set_error("awq_gemm: launch failed: %s", cudaGetErrorString(le));
    return B200_ECUDA;
  }
  if (splits > 1) {
    const int64_t total = (int64_t)p.size_m * p.size_n;
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    le = launch_pdl(awq_splitk_sum_kernel, dim3(blocks), dim3(256), (size_t)0, st, (__half*)p.c,
                    (const float*)a.partial, total, splits);
    if (le != cudaSuccess) {
      cudaGetLastError();
      set_error("awq_gemm split-K sum: launch failed: %s", cudaGetErrorString(le));
      return B200_ECUDA;
    }
  }
  return B200_OK;
}

int awq_check_weights(const char* op, const void* qweight, const void* scales, const void* qzeros, int k, int n,
                      int group_size) {
  B200_REQUIRE(k > 0 && n > 0 && group_size > 0, B200_EINVAL, "%s: bad sizes", op);
  B200_REQUIRE(qweight && scales && qzeros, B200_EINVAL, "%s: null pointer", op);
  // gemm_kernels.cu:506-513
  B200_REQUIRE(n % 64 == 0, B200_EINVAL, "%s: OC is not multiple of cta_N = 64", op);
  B200_REQUIRE(group_size % 32 == 0, B200_EINVAL, "%s: Group size should be a multiple of 32", op);
  B200_REQUIRE(k % group_size == 0, B200_EINVAL, "%s: IC = %d is not a multiple of the group size %d", op, k, group_size);
  B200_REQUIRE(((uintptr_t)qweight & 15) == 0 && ((uintptr_t)scales & 15) == 0 && ((uintptr_t)qzeros & 15) == 0,
               B200_EINVAL, "%s: operands must be 16-byte aligned", op);
  return B200_OK;
}

}  // namespace
}  // namespace b200

extern "C" {

int b200_awq_dequantize(void* out, const void* qweight, const void* scales, const void* qzeros, int32_t size_k,
                        int32_t size_n, int32_t group_size, void* stream) {
  using namespace b200;
  int rc = awq_check_weights("awq_dequantize", qweight, scales, qzeros, size_k, size_n, group_size);
  if (rc != B200_OK) return rc;
  B200_REQUIRE(out != nullptr && ((uintptr_t)out & 15) == 0, B200_EINVAL, "awq_dequantize: out must be 16-byte aligned");
  const int64_t total = (int64_t)size_k * (size_n / 8);
  int blocks = (int)((total + 255) / 256);
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  awq_dequantize_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__half*)out, (const uint32_t*)qweight,
                                                                  (const __half*)scales, (const uint32_t*)qzeros, size_k,
                                                                  size_n / 8, group_size);
  return check_launch("awq_dequantize");
}

int64_t b200_awq_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k) {
  if (m <= 0 || n <= 0 || k <= 0 || m > 1024) return 0;  // (awq.py:166-170 sends M >= 256 to dequantize + matmul)
  const int bm = m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : 128;
  const int tiles = ((m + bm - 1) / bm) * ((n + b200::kAwqTileN - 1) / b200::kAwqTileN);
  const int slots = (bm <= 32 ? 2 : 1) * b200::num_sms();
  int splits = tiles >= slots ? 1 : slots / tiles;
  const int num_kb = (k + b200::kAwqStageK - 1) / b200::kAwqStageK;
  if (splits > num_kb / 4) splits = num_kb / 4;
  return splits > 1 ? (int64_t)splits * m * n * 4 : 0;
}

int b200_awq_gemm(const b200_awq_gemm_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "awq_gemm: null params");
  const b200_awq_gemm_params& p = *pp;
  B200_REQUIRE(p.size_m >= 0, B200_EINVAL, "awq_gemm: bad sizes");
  int rc = awq_check_weights("awq_gemm", p.qweight, p.scales, p.qzeros, p.size_k, p.size_n, p.group_size);
  if (rc != B200_OK) return rc;
  if (p.size_m == 0) return B200_OK;
  B200_REQUIRE(p.c && p.a, B200_EINVAL, "awq_gemm: null pointer");
  B200_REQUIRE(p.size_k % 8 == 0 && p.lda >= p.size_k && p.lda % 8 == 0 && ((uintptr_t)p.a & 15) == 0, B200_EINVAL,
               "awq_gemm: activations must be 16-byte aligned with a 16-byte aligned row stride");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.size_m <= 16) return launch_awq<16>(p, st);
  if (p.size_m <= 32) return launch_awq<32>(p, st);
  if (p.size_m <= 64) return launch_awq<64>(p, st);
  return launch_awq<128>(p, st);
}

}  // extern "C"
