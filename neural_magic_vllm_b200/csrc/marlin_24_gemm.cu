// gptq_marlin_24_gemm on tcgen05 / TMEM (sm_100a): C[M,N] = A[M,K] * W, W 4-bit or 8-bit, 2:4 sparse
// along K, in the reference's Marlin-24 layout (values [K/32, 2N | 4N] int32, metadata [K/32, 2N] int16, scales
// [K/128 | 1, N] fp16).  Replaces csrc/quantization/marlin/sparse/marlin_24_cuda_kernel.cu:1015-1125
// (mma.sp.m16n8k32).  fp16 only, like the reference (:1100-1103: group 128 or channelwise).
//
// HBM traffic is the compressed stream (0.375 B per dense weight) read once, through three 2-D TMA
// boxes per 64-k stage (values, metadata, scales) plus the activation tile.  The tensor core takes
// the weights as they are stored -- 2:4 compressed -- through tcgen05.mma.sp (kind::f16):
//   A  = the kept values: four dequant warps (one per TMEM quadrant) unpack q - 8 with the
//        lop3 / hfma2 trick, apply the group scale (one rounding, bit-identical to the reference's
//        w_24_ref) and write the compressed fragment straight into TMEM (tcgen05.st 16x128b.x2: the
//        Marlin-24 value order IS that shape's register order); row = output column n
//   E  = the 2-bit position metadata, one 32-bit TMEM column per MMA of 32 dense k: lane
//        m%8 + 8 khalf + 16 (m/16) holds the 16 metadata bits of rows m (low half) and m + 8 (high
//        half) for dense k 16 khalf .. 16 khalf + 15 (read off the hardware with
//        benchmarks/micro/sparse_mma_probe.cu; cute's tmem_e_frg says the same).  One LDS.128 and
//        one PRMT per thread turn the Marlin-24 metadata words into that column.
//   B  = activations in shared memory (K-major, 128-byte swizzle), N = the token tile
//   D  = fp32 in TMEM, issued by one thread: two sparse MMAs per 64-k stage
// No expansion to dense anywhere: half the dequant stores and half the tensor work of a dense MMA.
//
// Layout facts used below (re-derived from marlin_24_perms.py:16-48 and format_24.py:22-49, and
// checked bit-exactly against tests/golden/marlin_24.npz):
//   values   tile = 16 compressed k rows (32 dense k) x 64 n = 128 int32; lane i (c = i/4, r = i%4)
//            owns words 4i..4i+3; word j holds n = 8c + 4b + j for b = 0,1 and the kept pairs of the
//            4-groups g = r and g = r + 4:  nibbles (0,4) = (b0, g=r), (1,5) = (b0, g=r+4),
//            (2,6) = (b1, g=r), (3,7) = (b1, g=r+4)
//            8-bit (interleave {0,2,1,3} instead of {0,2,4,6,1,3,5,7}): the tile is 256 int32, lane i
//            owns words 8i..8i+7, word 2j + b holds n = 8c + 4b + j with bytes = (g=r first value,
//            g=r+4 first, g=r second, g=r+4 second) -- the byte order unpack8 of the dense kernel takes
//   metadata per (32-k row, 64-n chunk) 256 bytes; the 8 bytes at unit u = (j%2) + 16 (j/2) + 2c are
//            four int16 [b0 groups 0-3, b1 groups 0-3, b0 groups 4-7, b1 groups 4-7] for n = 8c+4b+j;
//            group g's nibble = idx0 | idx1 << 2 (positions of the two kept values)
//   scales   grouped rows: position 8c + 2j + b holds column 8c + 4b + j; single row: natural order
// A TMEM quadrant q = 2 nt + b takes the 32 columns {64 nt + 8c + 4b + j}: TMEM lane p of the
// quadrant holds column 64 nt + 8 (p%8) + 4b + 2 (p/16) + (p%16)/8 (undone by the epilogue).
#include "marlin_dequant.cuh"
#include "umma.cuh"

namespace b200 {
namespace {

using namespace umma;
using namespace marlin;

constexpr int kTileN = 128;
constexpr int kStageK = 64;
#ifndef B200_M24_STAGES
#define B200_M24_STAGES 6
#endif
#ifndef B200_M24_ASTAGES
#define B200_M24_ASTAGES 3
#endif
constexpr int kStages = B200_M24_STAGES;
constexpr int kAStages = B200_M24_ASTAGES;
constexpr int kACols = kStageK / 4;          // 16 TMEM columns per stage: 32 kept fp16 per row
constexpr int kECol = kAStages * kACols;     // 48: metadata ring, two columns (even, odd) per stage
constexpr int kDCol = (kECol + 2 * kAStages + 15) / 16 * 16;  // 64: accumulator behind the A and metadata rings
static_assert(kECol % 2 == 0 && kECol + 2 * kAStages <= kDCol, "TMEM layout");
// dequant warps: one per TMEM quadrant; at BM = 64 (two CTAs per SM instead of four) two per quadrant,
// one for each 32-k half of a stage, so that an SM keeps 16 of them (as in marlin_gemm_tmem_kernel)
// 1 < M <= 16: the split-K sum kernel's traffic grows with the split count while the k ranges get
// short; measured on the Llama-2-7B layers (o: 16 -> 9 splits 8.7 -> 7.6 us, down 11.8 -> 11.2)
constexpr int kMaxSplitsSmallM = 9;

template <int BM>
__host__ __device__ constexpr int dequant_warps24() { return BM >= 64 ? 8 : 4; }
template <int BM>
__host__ __device__ constexpr int threads24() { return 64 + dequant_warps24<BM>() * 32; }
constexpr int kMetaBytes = 2 * 512;
constexpr int kScaleBytes = 1024;            // one row (256 B) padded: stages stay 1 KiB aligned

struct Args24 {
  void* c;
  float* partial;
  const uint16_t* scales;
  int m, n, k;
  int num_groups, group_shift;
  int kb_per_split, num_kb;
};

template <int BM, int BITS>
struct Plan24 {
  static constexpr int kValBytes = 2 * 1024 * BITS / 4;  // 2 rows of 32 dense k x 128 n
  static constexpr int kXBytes = BM * 128;
  static constexpr int kStageBytes = kXBytes + kValBytes + kMetaBytes + kScaleBytes;
  static constexpr int kBytes = kStages * kStageBytes + 1024 + 512;
  static constexpr uint32_t kTmemCols = kDCol + BM <= 128 ? 128 : 256;
};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
// D[tmem] (+)= A_sparse[tmem] * B[smem]: A holds the kept half of a 2:4 sparse M x 32 slab (8 columns
// of packed fp16 pairs), `tmem_e` (an even column; sparse_id2 = idesc bit 0 selects the odd one) the
// position metadata
__device__ __forceinline__ void mma_sp_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t tmem_e,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.sp.cta_group::1.kind::f16 [%0], [%1], %2, [%3], %4, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(tmem_e), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x2(uint32_t taddr, uint32_t r0, uint32_t r1) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(taddr), "r"(r0), "r"(r1) : "memory");
}

template <int BM, int BITS>
__global__ void __launch_bounds__(threads24<BM>(), (BM <= 32 ? 4 : 2))
marlin24_gemm_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                     const __grid_constant__ CUtensorMap tm_m, const __grid_constant__ CUtensorMap tm_s,
                     const Args24 args) {
  using P = Plan24<BM, BITS>;
  constexpr int kValBytes = P::kValBytes;
  constexpr int kWordsPerN = BITS / 2;  // int32 per output column in one 32-k row: 2 | 4
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stages = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kStages * P::kStageBytes);
  uint64_t* full_p = bars;
  uint64_t* empty_p = full_p + kStages;
  uint64_t* full_a = empty_p + kStages;
  uint64_t* empty_a = full_a + kAStages;
  uint64_t* tmem_full = empty_a + kAStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * kTileN;
  const int split = blockIdx.z;
  const int kb_begin = split * args.kb_per_split;
  const int kb_end = min(kb_begin + args.kb_per_split, args.num_kb);
  const bool grouped = args.num_groups > 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_m);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_p[s], 1);
      mbar_init(&empty_p[s], dequant_warps24<BM>() + 1);
    }
    for (int s = 0; s < kAStages; ++s) {
      mbar_init(&full_a[s], dequant_warps24<BM>());
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
    const bool leader = elect_one();  // issue under elect.sync: uniform-register operands (see scaled_mm.cu)
    const uint64_t pol_w = l2_policy_evict_first();
    const uint64_t pol_x = l2_policy_evict_last();
    const uint32_t tx_bytes = P::kXBytes + kValBytes + kMetaBytes + (grouped ? kTileN * 2 : 0);
    // programmatic dependent launch: values / metadata / scales of the first ring of stages
    // stream in before the wait; only the activation tiles depend on the predecessor
    const int pre = min(kb_end - kb_begin, kStages);
    if (leader) {
      for (int i = 0; i < pre; ++i) {
        const int kb = kb_begin + i;
        uint8_t* w_dst = stages + i * P::kStageBytes + P::kXBytes;
        mbar_arrive_expect_tx(&full_p[i], tx_bytes);
#pragma unroll
        for (int l = 0; l < BITS / 4; ++l)
          tma_load_2d(w_dst + l * 2048, &tm_w, n0 * kWordsPerN + l * 256, kb * 2, &full_p[i], pol_w);
        tma_load_2d(w_dst + kValBytes, &tm_m, n0 * 2, kb * 2, &full_p[i], pol_w);
        if (grouped)
          tma_load_2d(w_dst + kValBytes + kMetaBytes, &tm_s, n0, (kb * kStageK) >> args.group_shift,
                      &full_p[i], pol_x);
      }
    }
    pdl_wait();
    if (leader) {
      for (int i = 0; i < pre; ++i)
        tma_load_2d(stages + i * P::kStageBytes, &tm_x, (kb_begin + i) * kStageK, m0, &full_p[i], pol_x);
    }
    __syncwarp();
    int stage = pre == kStages ? 0 : pre;
    uint32_t phase = pre == kStages ? 1 : 0;
    for (int kb = kb_begin + pre; kb < kb_end; ++kb) {
      mbar_wait_bounded(&empty_p[stage], phase ^ 1);
      if (leader) {
        uint8_t* x_dst = stages + stage * P::kStageBytes;
        uint8_t* w_dst = x_dst + P::kXBytes;
        mbar_arrive_expect_tx(&full_p[stage], tx_bytes);
        tma_load_2d(x_dst, &tm_x, kb * kStageK, m0, &full_p[stage], pol_x);
#pragma unroll
        for (int l = 0; l < BITS / 4; ++l)
          tma_load_2d(w_dst + l * 2048, &tm_w, n0 * kWordsPerN + l * 256, kb * 2, &full_p[stage], pol_w);
        tma_load_2d(w_dst + kValBytes, &tm_m, n0 * 2, kb * 2, &full_p[stage], pol_w);
        if (grouped)
          tma_load_2d(w_dst + kValBytes + kMetaBytes, &tm_s, n0, (kb * kStageK) >> args.group_shift,
                      &full_p[stage], pol_x);
      }
      __syncwarp();
      if (++stage == kStages) stage = 0, phase ^= 1;
    }
  } else if (warp == 1) {
    // whole warp runs the loop (uniform operands); one elected lane issues (see scaled_mm.cu)
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = instr_desc<KIND_F16>(kTileN, BM) | (1u << 2);  // sparse A
      int stage = 0, as = 0;
      uint32_t aphase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        // a full_a arrival implies the stage's full_p: the dequant warps waited on it first
        mbar_wait_bounded(&full_a[as], aphase);
        fence_after_sync();
        const uint64_t dx = smem_desc_sw128(smem_u32(stages + stage * P::kStageBytes));
        if (leader) {
#pragma unroll
          for (int kt = 0; kt < 2; ++kt)  // 32 dense k per MMA: 8 columns of kept pairs, 64 bytes of B per row
            mma_sp_ts_f16(tmem_base + kDCol, tmem_base + as * kACols + kt * 8, desc_advance(dx, kt * 64),
                          tmem_base + kECol + as * 2, idesc | (uint32_t)kt, (kb > kb_begin || kt > 0) ? 1u : 0u);
          mma_commit(&empty_a[as]);
          mma_commit(&empty_p[stage]);
        }
        if (++stage == kStages) stage = 0;
        if (++as == kAStages) as = 0, aphase ^= 1;
      }
      if (leader) mma_commit(tmem_full);
      pdl_launch_dependents();  // main loop issued: the next kernel may become resident and wait
    }
  } else {
    const int qd = warp & 3;  // TMEM quadrant
    constexpr int kKtPerWarp = 2 / (dequant_warps24<BM>() / 4);  // 32-k halves of a stage per warp
    const int kt0 = ((warp - 2) >> 2) * kKtPerWarp;
    const int nt = qd >> 1, hb = qd & 1;
    const int c = lane >> 2, r = lane & 3;
    (void)r;
    uint32_t sc[4];                        // sc[j]: scale of n = 64 nt + 8c + 4 hb + j, both halves
#pragma unroll
    for (int j = 0; j < 4; ++j) sc[j] = 0x3c003c00u;
    if (!grouped) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + nt * 64 + 8 * c + 4 * hb + j;
        const uint32_t v = n < args.n ? args.scales[n] : 0;
        sc[j] = v | (v << 16);
      }
    }
    const uint32_t stage0 = smem_u32(stages) + P::kXBytes;
    // 4-bit: one 128-column box, rows 1 KiB, chunk nt at +512; 8-bit: one 64-column box per chunk
    const uint32_t val_off = BITS == 4 ? nt * 512 + lane * 16 : nt * 2048 + lane * 32;
    // metadata column: this thread writes TMEM lane `lane` of its quadrant = rows (c' , j = 2 bb'),
    // (c', j = 2 bb' + 1) with c' = lane % 8, bb' = lane / 16, k half (lane / 8) % 2; the two 8-byte
    // units of those rows are adjacent: 16 bytes at unit 16 bb' + 2 c'
    const uint32_t meta_off = kValBytes + nt * 256 + (lane >> 4) * 128 + (lane & 7) * 16;
    const int khalf = (lane >> 3) & 1;
    const uint32_t meta_sel = hb ? 0x7632u : 0x5410u;  // the int16 of this quadrant's b from both rows
    const uint32_t scale_off = kValBytes + kMetaBytes + (nt * 64 + 8 * c) * 2;
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
    int cur_group = -1;
    // unrolled over one full trip of the stage ring: stage, TMEM A-ring slot and its phase are
    // compile-time constants (kStages is a multiple of kAStages), as in marlin_gemm_tmem_kernel
    static_assert(kStages % kAStages == 0, "the unrolled dequant loop needs whole A-ring wraps");
    uint32_t phase = 0;
    for (int kb0 = kb_begin; kb0 < kb_end; kb0 += kStages, phase ^= 1) {
#pragma unroll
     for (int us = 0; us < kStages; ++us) {
      const int kb = kb0 + us;
      if (kb >= kb_end) break;
      const int stage = us, as = us % kAStages;
      const uint32_t aphase = (us / kAStages) & 1;
      mbar_wait_bounded(&full_p[stage], phase);
      const uint32_t base = stage0 + stage * P::kStageBytes;
      uint32_t words[2][4], mword[2];
#pragma unroll
      for (int kk = 0; kk < kKtPerWarp; ++kk) {  // registers indexed by kk (compile time), addresses by kt
        const int kt = kt0 + kk;
        if constexpr (BITS == 4) {
          const uint4 w = lds128(base + val_off + kt * 1024);
          words[kk][0] = w.x >> (8 * hb), words[kk][1] = w.y >> (8 * hb);
          words[kk][2] = w.z >> (8 * hb), words[kk][3] = w.w >> (8 * hb);
        } else {  // word 2 j + b of the lane's eight
          const uint4 w0 = lds128(base + val_off + kt * 1024), w1 = lds128(base + val_off + kt * 1024 + 16);
          words[kk][0] = hb ? w0.y : w0.x, words[kk][1] = hb ? w0.w : w0.z;
          words[kk][2] = hb ? w1.y : w1.x, words[kk][3] = hb ? w1.w : w1.z;
        }
        // [row j = 2 bb': k half 0, k half 1][row j = 2 bb' + 1: k half 0, k half 1], each word = b0 | b1 << 16
        const uint4 mw = lds128(base + meta_off + kt * 512);
        mword[kk] = __byte_perm(khalf ? mw.y : mw.x, khalf ? mw.w : mw.z, meta_sel);
      }
      if (grouped) {
        const int g = (kb * kStageK) >> args.group_shift;
        if (g != cur_group) {  // warp-uniform
          cur_group = g;
          const uint4 v = lds128(base + scale_off);
          sc[0] = dup16(v.x, hb), sc[1] = dup16(v.y, hb), sc[2] = dup16(v.z, hb), sc[3] = dup16(v.w, hb);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_p[stage]);
      mbar_wait_bounded(&empty_a[as], aphase ^ 1);
      fence_after_sync();
      const uint32_t t_dst = t_lane + as * kACols;
#pragma unroll
      for (int kk = 0; kk < kKtPerWarp; ++kk) {
        const int kt = kt0 + kk;
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
          uint32_t lo[2], hi[2];  // [row c | row c + 8]: kept pairs of groups r (lo) and r + 4 (hi)
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int j = 2 * bb + u;
            constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400;
            constexpr uint32_t SUB = 0x64086408, MUL = 0x2c002c00, ADD = 0xd480d480;
            const uint32_t q = words[kk][j];
            // exact q - zp, x scale (one rounding)
            if constexpr (BITS == 4) {
              lo[u] = mul2<__half>(h2_sub(lop3<kAndOr>(q, LO, EX), SUB), sc[j]);
              hi[u] = mul2<__half>(h2_fma(lop3<kAndOr>(q, HI, EX), MUL, ADD), sc[j]);
            } else {
              unpack8<__half>(q, lo[u], hi[u]);
              lo[u] = mul2<__half>(lo[u], sc[j]);
              hi[u] = mul2<__half>(hi[u], sc[j]);
            }
          }
          // kept-pair column g of row kt: groups r and r + 4 = columns r and 4 + r of the 8
          tmem_st_16x128b_x2(t_dst + ((uint32_t)(16 * bb) << 16) + kt * 8, lo[0], lo[1], hi[0], hi[1]);
        }
      }
      if constexpr (kKtPerWarp == 2) {
        tmem_st_32x32b_x2(t_lane + kECol + as * 2, mword[0], mword[1]);
      } else {  // this warp's k half only: one metadata column
        asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(t_lane + kECol + as * 2 + kt0),
                     "r"(mword[0])
                     : "memory");
      }
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[as]);
     }
    }
    // epilogue: TMEM lane p of quadrant (nt, hb) is column 8 (p%8) + 4 hb + 2 (p/16) + (p%16)/8
    {
      const int n = n0 + nt * 64 + 8 * (lane & 7) + 4 * hb + 2 * (lane >> 4) + ((lane & 15) >> 3);
      mbar_wait_bounded(tmem_full, 0);
      fence_after_sync();
      const uint32_t taddr = t_lane + kDCol;
#pragma unroll 1
      for (int cc = (warp - 2) >> 2; cc < BM / 16; cc += dequant_warps24<BM>() / 4) {  // token chunks interleaved over the quadrant's warps
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
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, P::kTmemCols);
  }
}

__global__ void __launch_bounds__(256) splitk_sum24_kernel(__half* __restrict__ out, const float* __restrict__ ws,
                                                           int64_t total, int splits) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    int s = 0;
    for (; s + 4 <= splits; s += 4) {
      const float v0 = ws[(int64_t)s * total + i], v1 = ws[(int64_t)(s + 1) * total + i];
      const float v2 = ws[(int64_t)(s + 2) * total + i], v3 = ws[(int64_t)(s + 3) * total + i];
      acc += v0, acc += v1, acc += v2, acc += v3;
    }
    for (; s < splits; ++s) acc += ws[(int64_t)s * total + i];
    out[i] = __float2half_rn(acc);
  }
}

template <int BM, int BITS>
int launch24(const b200_gptq_marlin_24_gemm_params& p, cudaStream_t st) {
  using P = Plan24<BM, BITS>;
  CUtensorMap tm_x, tm_w, tm_m, tm_s;
  int rc = make_tmap_2d(&tm_x, p.a, (uint64_t)p.size_m, (uint64_t)p.size_k, (uint64_t)p.size_k * 2, 2, BM);
  if (rc != B200_OK) return rc;
  const uint64_t rows = (uint64_t)p.size_k / 32, cols = (uint64_t)p.size_n * 2;
  const uint64_t wcols = (uint64_t)p.size_n * (BITS / 2);
  rc = make_tmap_2d_ex(&tm_w, p.b_q_weight, rows, wcols, wcols * 4, CU_TENSOR_MAP_DATA_TYPE_INT32, 256, 2,
                       CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_m, p.b_meta, rows, cols, cols * 2, CU_TENSOR_MAP_DATA_TYPE_UINT16, 256, 2,
                       CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_s, p.b_scales, (uint64_t)p.num_groups, (uint64_t)p.size_n, (uint64_t)p.size_n * 2,
                       CU_TENSOR_MAP_DATA_TYPE_UINT16, kTileN, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  Args24 a{};
  a.c = p.c;
  a.scales = reinterpret_cast<const uint16_t*>(p.b_scales);
  a.m = p.size_m, a.n = p.size_n, a.k = p.size_k;
  a.num_groups = p.num_groups;
  a.group_shift = p.num_groups > 1 ? __builtin_ctz(p.size_k / p.num_groups) : 30;
  a.num_kb = p.size_k / kStageK;
  const int m_tiles = (p.size_m + BM - 1) / BM, n_tiles = (p.size_n + kTileN - 1) / kTileN;
  const int tiles = m_tiles * n_tiles;
  const int target = (BM <= 32 ? 4 : 2) * num_sms();
  int splits = tiles >= target ? 1 : target / tiles;
  if (BM == 16 && p.size_m > 1 && splits > kMaxSplitsSmallM) splits = kMaxSplitsSmallM;
  static const int forced_splits = [] {  // debug knob: B200_M24_SPLITS
    const char* e = getenv("B200_M24_SPLITS");
    return e ? atoi(e) : 0;
  }();
  if (forced_splits > 0) splits = forced_splits;
  if (splits > a.num_kb / 4) splits = a.num_kb / 4;
  const int64_t per = (int64_t)p.size_m * p.size_n * 4;
  if (p.splitk_ws == nullptr) splits = 1;
  else if (splits > p.splitk_ws_bytes / per) splits = (int)(p.splitk_ws_bytes / per);
  if (splits < 1) splits = 1;
  a.kb_per_split = (a.num_kb + splits - 1) / splits;
  splits = (a.num_kb + a.kb_per_split - 1) / a.kb_per_split;
  a.partial = splits > 1 ? reinterpret_cast<float*>(p.splitk_ws) : nullptr;
  auto kern = marlin24_gemm_kernel<BM, BITS>;
  static size_t smem_configured[kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, smem_configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_24_gemm: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  launch_pdl(kern, dim3(m_tiles, n_tiles, splits), dim3(threads24<BM>()), P::kBytes, st, tm_x, tm_w, tm_m, tm_s, a);
  rc = check_launch("gptq_marlin_24_gemm");
  if (rc != B200_OK) return rc;
  if (splits > 1) {
    const int64_t total = (int64_t)p.size_m * p.size_n;
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    splitk_sum24_kernel<<<blocks, 256, 0, st>>>((__half*)p.c, a.partial, total, splits);
    rc = check_launch("gptq_marlin_24_gemm split-K sum");
  }
  return rc;
}

}  // namespace
}  // namespace b200

// split-K scratch that lets the 2:4 kernel fill the GPU for this shape (0 = none needed): the same
// arithmetic as launch24's split choice
extern "C" int64_t b200_gptq_marlin_24_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k) {
  using namespace b200;
  if (m <= 0 || n <= 0 || k <= 0) return 0;
  const int bm = m <= 16 ? 16 : m <= 32 ? 32 : 64;
  const int tiles = ((m + bm - 1) / bm) * ((n + kTileN - 1) / kTileN);
  const int target = (bm <= 32 ? 4 : 2) * num_sms();
  int splits = tiles >= target ? 1 : target / tiles;
  if (bm == 16 && m > 1 && splits > kMaxSplitsSmallM) splits = kMaxSplitsSmallM;
  if (splits > (k / kStageK) / 4) splits = (k / kStageK) / 4;
  return splits > 1 ? (int64_t)splits * m * n * 4 : 0;
}

extern "C" int b200_gptq_marlin_24_gemm(const b200_gptq_marlin_24_gemm_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "gptq_marlin_24_gemm: null params");
  const b200_gptq_marlin_24_gemm_params& p = *pp;
  // marlin_24_cuda_kernel.cu:1024-1116
  B200_REQUIRE(p.num_bits == 4 || p.num_bits == 8, B200_EINVAL, "num_bits must be 4 or 8. Got = %d", p.num_bits);
  B200_REQUIRE(p.size_m >= 0 && p.size_n > 0 && p.size_k > 0, B200_EINVAL, "gptq_marlin_24_gemm: bad sizes");
  if (p.size_m == 0) return B200_OK;
  B200_REQUIRE(p.size_k % 128 == 0, B200_EINVAL, "size_k = %d is not divisible by 128", p.size_k);
  B200_REQUIRE(p.size_n % 128 == 0, B200_EINVAL, "size_n = %d is not divisible by 128", p.size_n);
  B200_REQUIRE(p.c && p.a && p.b_q_weight && p.b_meta && p.b_scales, B200_EINVAL,
               "gptq_marlin_24_gemm: null pointer");
  B200_REQUIRE(p.num_groups >= 1, B200_EINVAL, "gptq_marlin_24_gemm: b_scales has no rows");
  B200_REQUIRE(p.num_groups == 1 || p.size_k / p.num_groups == 128, B200_EINVAL, "Unexpected groupsize = %d",
               p.size_k / p.num_groups);
  B200_REQUIRE(((uintptr_t)p.a & 15) == 0 && ((uintptr_t)p.b_q_weight & 15) == 0 &&
                   ((uintptr_t)p.b_meta & 15) == 0 && ((uintptr_t)p.b_scales & 15) == 0,
               B200_EINVAL, "gptq_marlin_24_gemm: operands must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.num_bits == 4) {
    if (p.size_m <= 16) return launch24<16, 4>(p, st);
    if (p.size_m <= 32) return launch24<32, 4>(p, st);
    return launch24<64, 4>(p, st);
  }
  if (p.size_m <= 16) return launch24<16, 8>(p, st);
  if (p.size_m <= 32) return launch24<32, 8>(p, st);
  return launch24<64, 8>(p, st);
}
