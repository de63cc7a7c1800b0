// gptq_marlin_gemm on tcgen05 / TMEM (sm_100a): C[M,N] = A[M,K] * dequant(W), W 4-bit or 8-bit
// Marlin-repacked with fp16 / bf16 group scales, symmetric (zero point 8 / 128), optional
// act-order.  Replaces csrc/quantization/gptq_marlin/gptq_marlin.cu:1726-1866 (mma.sync Marlin).
//
// One CTA computes a 128 (n) x BM (tokens) tile; the weights are the MMA *A* operand:
//   warp 0      bulk-async (TMA engine) producer: packed weights (4 x 1 KiB rows of the Marlin
//               tensor per 64-k stage), the group-scale rows of the stage, and the activation
//               tile through a 2-D tensor map (128-byte swizzle)
//   warps 2..9  dequant: LDS.128 packed words -> lop3 / hfma2 exact (q - zp) -> x group scale
//               (one rounding: the dequantised tile is bit-identical to the reference's w_ref)
//               -> stmatrix into the K-major swizzled operand tile -> fence.proxy.async
//   warp 1      one thread issues tcgen05.mma (M=128, N=BM, K=16 x4 per stage), fp32 in TMEM
//   warps 2..5  epilogue: tcgen05.ld -> fp16/bf16 -> C (or fp32 split-K partials)
// Decode (small M) is weight streaming: BM = 16..64, split-K fills the 148 SMs; HBM traffic is
// the packed bytes, read exactly once.
#include <cstdlib>

#include "marlin_dequant.cuh"
#include "umma.cuh"

namespace b200 {
// layer_glue.cu
bool splitk_add_rms_eligible(int hidden, const void* out, const void* residual, const void* weight);
int splitk_add_rms_norm(int dtype, void* out, const float* ws, int splits, void* residual, const void* weight,
                        float eps, int tokens, int hidden, cudaStream_t st);
namespace {

using namespace umma;
using namespace marlin;

constexpr int kTileN = 128;
constexpr int kStageK = 64;  // k per stage (128 bytes of fp16/bf16)
constexpr int kDequantWarps = 8;
constexpr int kThreads = 64 + kDequantWarps * 32;
constexpr int kAStages = 3;             // dequantised operand ring
constexpr int kSubTileBytes = kTileN * 128;  // 16 KiB: one 128(n) x 64(k) operand sub-tile
constexpr int kMaxScaleRows = 4;        // group_size >= 16 -> at most 4 groups per 64-k stage

struct MarlinArgs {
  void* c;
  float* partial;  // split-K partials [splits, M, N] or nullptr
  const uint32_t* w;
  const void* scales;
  const int32_t* g_idx;  // act-order: group of every (sorted) k row; nullptr otherwise
  int m, n, k;
  int num_groups, group_size;  // group_size = k when num_groups == 1
  int group_shift;             // log2(group_size) when it is a power of two, else -1
  int s_rows;                  // scale rows fetched per stage (0: scales are not staged)
  int kb_per_split, num_kb;
  int w_evict_first;
  int fp8;  // 8-bit words hold fp8 e4m3 values (fp8_marlin_gemm), channel-wise scales
  int32_t* locks;  // decode kernel, split-K: arrival counter per n tile (zero on entry, zero on exit)
};

// NT = 128-row sub-tiles of n per CTA: large-M tiles take 2 so every activation tile fetched from
// L2 feeds two MMAs (the X tile, not the packed weights, is what loads L2 at prefill)
template <int BITS, int BM, int NT>
struct Plan {
  static constexpr int kCtaN = kTileN * NT;
  static constexpr int kATileBytes = NT * kSubTileBytes;
  static constexpr int kPackedBytes = kStageK * kCtaN * BITS / 8;  // 4 KiB / 8 KiB per sub-tile
  static constexpr int kScaleBytes = kMaxScaleRows * kCtaN * 2;    // 1 KiB per sub-tile
  static constexpr int kXBytes = BM * 128;
  static constexpr int kStageBytes = kPackedBytes + kScaleBytes + kXBytes;  // multiple of 1 KiB
  static constexpr int kBudget = (BM <= 64 ? 100 : 225) * 1024 - kAStages * kATileBytes;
  static constexpr int kStages = kBudget / kStageBytes > 8 ? 8 : kBudget / kStageBytes;
  static constexpr int kBytes = kStages * kStageBytes + kAStages * kATileBytes + 1024 + 512;
  static_assert(kStages >= 2, "pipeline too shallow");
};

__device__ __forceinline__ int group_of(const MarlinArgs& a, int k) {
  return a.group_shift >= 0 ? (k >> a.group_shift) : (k / a.group_size);
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}

// packed weights arrive as TMA boxes of 4 k-tile rows x 1 KiB (256 int32: 128 columns at 4 bits,
// 64 at 8 bits), box l of a stage at l * 4 KiB.  Offset of the 16(k) x 64(n) Marlin tile
// (k-tile kt, 64-column chunk c of the CTA's n range) inside the stage's packed area:
template <int BITS>
__device__ __forceinline__ uint32_t chunk_off(int kt, int c) {
  return BITS == 4 ? (uint32_t)((c >> 1) * 4096 + kt * 1024 + (c & 1) * 512)
                   : (uint32_t)(c * 4096 + kt * 1024);
}

template <int BITS, int BM, int NT, typename T, bool ACT_ORDER>
__global__ void __launch_bounds__(kThreads, 1)
marlin_gemm_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                   const __grid_constant__ CUtensorMap tm_s, const MarlinArgs args) {
  using P = Plan<BITS, BM, NT>;
  constexpr int kStages = P::kStages;
  constexpr int kATileBytes = P::kATileBytes;
  constexpr int kCtaN = P::kCtaN;
  constexpr uint32_t kTmemCols = NT * BM < 32 ? 32 : NT * BM;
  static_assert(kTmemCols <= 512 && (kTmemCols & (kTmemCols - 1)) == 0, "TMEM columns");
  constexpr int KIND = DT<T>::id == B200_F16 ? KIND_F16 : KIND_BF16;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = smem;                               // kAStages x 16 KiB (1024-aligned)
  uint8_t* stages = smem + kAStages * kATileBytes;       // kStages x {X | packed | scales}
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kStages * P::kStageBytes);
  uint64_t* full_p = bars;
  uint64_t* empty_p = full_p + kStages;
  uint64_t* full_a = empty_p + kStages;
  uint64_t* empty_a = full_a + kAStages;
  uint64_t* tmem_full = empty_a + kAStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * kCtaN;
  const int split = blockIdx.z;
  const int kb_begin = split * args.kb_per_split;
  const int kb_end = min(kb_begin + args.kb_per_split, args.num_kb);
  const int n_valid = min(kCtaN, args.n - n0);  // multiple of 64
  const bool grouped = args.num_groups > 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_p[s], 1);
      mbar_init(&empty_p[s], kDequantWarps + 1);
    }
    for (int s = 0; s < kAStages; ++s) {
      mbar_init(&full_a[s], kDequantWarps);
      mbar_init(&empty_a[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    const bool leader = elect_one();  // issue under elect.sync: uniform-register operands (see scaled_mm.cu)
    // ======================= producer =======================
    // the whole warp runs the loop (uniform control flow keeps every operand in uniform
    // registers); lane 0 issues: one TMA per operand per stage
    const uint64_t pol_w = args.w_evict_first ? l2_policy_evict_first() : l2_policy_evict_last();
    const uint64_t pol_x = l2_policy_evict_last();
    constexpr int kWordsPerN = 16 * BITS / 32;
    constexpr int kLoadsW = kCtaN * kWordsPerN / 256;
    const uint32_t tx_bytes = P::kXBytes + P::kPackedBytes + (uint32_t)args.s_rows * (kCtaN * 2);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait_bounded(&empty_p[stage], phase ^ 1);
      if (leader) {
        uint8_t* x_dst = stages + stage * P::kStageBytes;
        uint8_t* w_dst = x_dst + P::kXBytes;
        uint8_t* s_dst = w_dst + P::kPackedBytes;
        mbar_arrive_expect_tx(&full_p[stage], tx_bytes);
        tma_load_2d(x_dst, &tm_x, kb * kStageK, m0, &full_p[stage], pol_x);
#pragma unroll
        for (int l = 0; l < kLoadsW; ++l)
          tma_load_2d(w_dst + l * 4096, &tm_w, n0 * kWordsPerN + l * 256, kb * 4, &full_p[stage], pol_w);
        if (args.s_rows > 0)
          tma_load_2d(s_dst, &tm_s, n0, group_of(args, kb * kStageK), &full_p[stage], pol_x);
      }
      __syncwarp();
      if (++stage == kStages) stage = 0, phase ^= 1;
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    // whole warp runs the loop (uniform operands); one elected lane issues (see scaled_mm.cu)
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = instr_desc<KIND>(kTileN, BM);
      int stage = 0, as = 0;
      uint32_t aphase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        // dequantised weights written; implies the stage's full_p (the dequant warps waited on it)
        mbar_wait_bounded(&full_a[as], aphase);
        fence_after_sync();
        const uint64_t dx = smem_desc_sw128(smem_u32(stages + stage * P::kStageBytes));
        if (leader) {
#pragma unroll
          for (int h = 0; h < NT; ++h) {
            const uint64_t dw = smem_desc_sw128(smem_u32(a_tiles + as * kATileBytes + h * kSubTileBytes));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              mma_ss<KIND>(tmem_base + h * BM, desc_advance(dw, k * 32), desc_advance(dx, k * 32), idesc,
                           (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          mma_commit(&empty_a[as]);
          mma_commit(&empty_p[stage]);
        }
        if (++stage == kStages) stage = 0;
        if (++as == kAStages) as = 0, aphase ^= 1;
      }
      if (leader) mma_commit(tmem_full);
    }
  } else {
    // ======================= dequant (8 warps) + epilogue (first 4 of them) ==================
    const int dw = warp - 2;      // 0..7
    const int kt = dw >> 1;       // k-tile of the stage (16 rows)
    const int nt0 = dw & 1;       // 64-column half of every 128-row sub-tile
    const int a_row = lane >> 2;  // n offset inside an 8-row group
    // stmatrix row address of this lane: matrix q = lane/8 (b = q>>1, high-k = q&1), row r = lane%8
    const int q = lane >> 3, r = lane & 7;
    const uint32_t st_off = (uint32_t)((nt0 * 64 + 8 * (q >> 1) + r) * 128 + (((kt * 2 + (q & 1)) ^ r) << 4));
    // single-group scales live in registers for the whole kernel
    uint32_t sc[NT][8];  // sc[h][2*j + b]: scale of n = 128 h + 64 nt0 + 16 j + a_row + 8 b, in both halves
#pragma unroll
    for (int h = 0; h < NT; ++h) {
#pragma unroll
      for (int i = 0; i < 8; ++i) sc[h][i] = DT<T>::pack2(1.f, 1.f);
      if (!grouped && (h * 128 + nt0 * 64) < n_valid) {
        const uint16_t* s = reinterpret_cast<const uint16_t*>(args.scales);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int n = n0 + h * 128 + nt0 * 64 + 16 * (i >> 1) + a_row + 8 * (i & 1);
          const uint32_t v = s[scale_pos_single(n)];
          sc[h][i] = v | (v << 16);
          if (BITS == 8 && args.fp8) sc[h][i] = fp8_scale_bias<T>(sc[h][i]);
        }
      }
    }
    int cur_group = -1;
    int stage = 0, as = 0;
    uint32_t phase = 0, aphase = 0;
    const uint32_t stages_u32 = smem_u32(stages);
    const uint32_t a_tiles_u32 = smem_u32(a_tiles) + st_off;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait_bounded(&full_p[stage], phase);
      const uint32_t w_src = stages_u32 + stage * P::kStageBytes + P::kXBytes;
      const uint32_t s_src = w_src + P::kPackedBytes;
      uint32_t words[NT][BITS == 4 ? 4 : 8];
#pragma unroll
      for (int h = 0; h < NT; ++h) {
        const uint32_t p = w_src + chunk_off<BITS>(kt, 2 * h + nt0) + lane * (BITS * 4);
        uint4 v0 = lds128(p);
        words[h][0] = v0.x, words[h][1] = v0.y, words[h][2] = v0.z, words[h][3] = v0.w;
        if constexpr (BITS == 8) {
          uint4 v1 = lds128(p + 16);
          words[h][4] = v1.x, words[h][5] = v1.y, words[h][6] = v1.z, words[h][7] = v1.w;
        }
      }
      if (grouped && !ACT_ORDER) {
        const int g = group_of(args, kb * kStageK + kt * 16);
        if (g != cur_group) {  // warp-uniform
          cur_group = g;
          const int g_lo = group_of(args, kb * kStageK);
#pragma unroll
          for (int h = 0; h < NT; ++h) {
            const uint4 v = lds128(s_src + (g - g_lo) * (kCtaN * 2) + (h * 128 + nt0 * 64 + 8 * a_row) * 2);
            sc[h][0] = dup16(v.x, 0), sc[h][1] = dup16(v.x, 1), sc[h][2] = dup16(v.y, 0), sc[h][3] = dup16(v.y, 1);
            sc[h][4] = dup16(v.z, 0), sc[h][5] = dup16(v.z, 1), sc[h][6] = dup16(v.w, 0), sc[h][7] = dup16(v.w, 1);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_p[stage]);  // packed words + scales are in registers
      mbar_wait_bounded(&empty_a[as], aphase ^ 1);
#pragma unroll
      for (int h = 0; h < NT; ++h) {
        const uint32_t a_base = a_tiles_u32 + as * kATileBytes + h * kSubTileBytes;
        const bool nt_valid = (h * 128 + nt0 * 64) < n_valid;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t rg[4];
          if constexpr (BITS == 4) {
            unpack4<T>(words[h][j], rg);
          } else {
            if (args.fp8) {
              unpack8_fp8<T>(words[h][2 * j], rg[0], rg[1]);
              unpack8_fp8<T>(words[h][2 * j + 1], rg[2], rg[3]);
            } else {
              unpack8<T>(words[h][2 * j], rg[0], rg[1]);
              unpack8<T>(words[h][2 * j + 1], rg[2], rg[3]);
            }
          }
          if constexpr (ACT_ORDER) {
            // every k row may belong to a different group: scale each element separately
            const uint16_t* s = reinterpret_cast<const uint16_t*>(args.scales);
            const int kg = kb * kStageK + kt * 16 + 2 * (lane & 3);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int kk = kg + ((e & 1) << 3);
              const int n = n0 + h * 128 + nt0 * 64 + 16 * j + a_row + 8 * (e >> 1);
              uint32_t s2 = 0;
              if (nt_valid) {
                const int pos = scale_pos_grouped(n);
                const uint32_t s_lo = s[(int64_t)args.g_idx[kk] * args.n + pos];
                const uint32_t s_hi = s[(int64_t)args.g_idx[kk + 1] * args.n + pos];
                s2 = s_lo | (s_hi << 16);
              }
              rg[e] = mul2<T>(rg[e], s2);
            }
          } else {
            rg[0] = mul2<T>(rg[0], sc[h][2 * j]);
            rg[1] = mul2<T>(rg[1], sc[h][2 * j]);
            rg[2] = mul2<T>(rg[2], sc[h][2 * j + 1]);
            rg[3] = mul2<T>(rg[3], sc[h][2 * j + 1]);
          }
          stmatrix_x4(a_base + j * (16 * 128), rg);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[as]);
      if (++stage == kStages) stage = 0, phase ^= 1;
      if (++as == kAStages) as = 0, aphase ^= 1;
    }
    if (dw < 4) {
      const int quarter = warp & 3;
      mbar_wait_bounded(tmem_full, 0);
      fence_after_sync();
#pragma unroll 1
      for (int h = 0; h < NT; ++h) {
        const int n = n0 + h * 128 + quarter * 32 + lane;
        const uint32_t taddr = tmem_base + h * BM + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
        for (int c = 0; c < BM / 16; ++c) {
          uint32_t rr[16];
          tmem_ld16(taddr + c * 16, rr);
          tmem_ld_wait();
          if (n < args.n) {
            if (args.partial != nullptr) {
              float* ws = args.partial + ((int64_t)split * args.m) * args.n;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int mm = m0 + c * 16 + j;
                if (mm < args.m) ws[(int64_t)mm * args.n + n] = __uint_as_float(rr[j]);
              }
            } else {
              T* out = reinterpret_cast<T*>(args.c);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int mm = m0 + c * 16 + j;
                if (mm < args.m) out[(int64_t)mm * args.n + n] = DT<T>::from_f(__uint_as_float(rr[j]));
              }
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
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// Decode kernel (BM <= 64, no act-order): the dequantised weights never touch shared memory.
// The dequant warps write their mma-fragment registers straight into TMEM (tcgen05.st 16x128b.x2,
// whose thread->(lane, column) map IS the Marlin fragment order) and the MMA takes A from TMEM:
// no stmatrix traffic, no proxy fence, half the shared-memory bandwidth per stage.
//   TMEM columns: [0, 96) = ring of 3 A stages (64 k = 32 columns each), [96, 96 + BM) = D.
//   warp w's TMEM quadrant is w % 4 = rows 32 q .. 32 q + 31 of the 128-row n tile, so the two
//   warps of a quadrant split the four 16-row k tiles of a stage between them.
// ------------------------------------------------------------------------------------------------
constexpr int kTmemAStageCols = kStageK / 2;  // 32
// dequant warps: one per TMEM quadrant; at BM = 64 -- two CTAs per SM instead of four -- two per
// quadrant, each taking half of the stage's k range, so that an SM still has 16 of them
#ifndef B200_TMEM_DW64
#define B200_TMEM_DW64 8
#endif
template <int BM>
__host__ __device__ constexpr int tmem_dequant_warps() { return BM >= 64 ? B200_TMEM_DW64 : 4; }
template <int BM>
__host__ __device__ constexpr int tmem_threads() { return 64 + tmem_dequant_warps<BM>() * 32; }
constexpr int kTmemDCol = kAStages * kTmemAStageCols;

template <int BITS, int BM>
struct PlanT {
  static constexpr int kPackedBytes = kStageK * kTileN * BITS / 8;
  static constexpr int kScaleBytes = kMaxScaleRows * kTileN * 2;
  static constexpr int kXBytes = BM * 128;
  static constexpr int kStageBytes = kPackedBytes + kScaleBytes + kXBytes;
  static constexpr int kStages = 6;
  static constexpr int kBytes = kStages * kStageBytes + 1024 + 512;
  static constexpr uint32_t kTmemCols = kTmemDCol + BM <= 128 ? 128 : 256;
};

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}

template <int BITS, int BM, typename T>
__global__ void __launch_bounds__(tmem_threads<BM>(), (BM <= 32 ? 4 : 2))
marlin_gemm_tmem_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                        const __grid_constant__ CUtensorMap tm_s, const MarlinArgs args) {
  using P = PlanT<BITS, BM>;
  constexpr int kStages = P::kStages;
  constexpr int kTDequantWarps = tmem_dequant_warps<BM>();
  constexpr int kKtPerWarp = 4 / (kTDequantWarps / 4);  // 16-k sub-steps of a stage per dequant warp
  constexpr int KIND = DT<T>::id == B200_F16 ? KIND_F16 : KIND_BF16;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stages = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kStages * P::kStageBytes);
  uint64_t* full_p = bars;
  uint64_t* empty_p = full_p + kStages;
  uint64_t* full_a = empty_p + kStages;
  uint64_t* empty_a = full_a + kAStages;
  uint64_t* tmem_full = empty_a + kAStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  // warp-uniform by construction: tell the compiler, so TMEM / barrier addresses live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * kTileN;
  const int split = blockIdx.z;
  const int kb_begin = split * args.kb_per_split;
  const int kb_end = min(kb_begin + args.kb_per_split, args.num_kb);
  const int n_valid = min(kTileN, args.n - n0);  // multiple of 64
  const bool grouped = args.num_groups > 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_w);
    if (args.s_rows > 0) tma_prefetch_desc(&tm_s);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_p[s], 1);
      mbar_init(&empty_p[s], kTDequantWarps + 1);
    }
    for (int s = 0; s < kAStages; ++s) {
      mbar_init(&full_a[s], kTDequantWarps);
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
    // the whole warp runs the loop (uniform control flow keeps every operand in uniform
    // registers); lane 0 issues: one TMA per operand per stage
    const uint64_t pol_w = args.w_evict_first ? l2_policy_evict_first() : l2_policy_evict_last();
    const uint64_t pol_x = l2_policy_evict_last();
    constexpr int kWordsPerN = 16 * BITS / 32;
    constexpr int kLoadsW = kTileN * kWordsPerN / 256;
    const uint32_t tx_bytes = P::kXBytes + P::kPackedBytes + (uint32_t)args.s_rows * (kTileN * 2);
    // Programmatic dependent launch: this grid may be resident while its predecessor (the kernel
    // that produces the activations) is still running.  Weights and scales do not depend on it, so
    // the first ring of stages streams them in right away; only the activation tiles wait.
    const int pre = min(kb_end - kb_begin, kStages);
    if (leader) {
      for (int i = 0; i < pre; ++i) {  // fresh ring: every stage is empty
        const int kb = kb_begin + i;
        uint8_t* w_dst = stages + i * P::kStageBytes + P::kXBytes;
        mbar_arrive_expect_tx(&full_p[i], tx_bytes);
#pragma unroll
        for (int l = 0; l < kLoadsW; ++l)
          tma_load_2d(w_dst + l * 4096, &tm_w, n0 * kWordsPerN + l * 256, kb * 4, &full_p[i], pol_w);
        if (args.s_rows > 0)
          tma_load_2d(w_dst + P::kPackedBytes, &tm_s, n0, ((kb * kStageK) >> args.group_shift), &full_p[i], pol_x);
      }
    }
    pdl_wait();  // the predecessor grid has completed: activations are visible
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
        uint8_t* s_dst = w_dst + P::kPackedBytes;
        mbar_arrive_expect_tx(&full_p[stage], tx_bytes);
        tma_load_2d(x_dst, &tm_x, kb * kStageK, m0, &full_p[stage], pol_x);
#pragma unroll
        for (int l = 0; l < kLoadsW; ++l)
          tma_load_2d(w_dst + l * 4096, &tm_w, n0 * kWordsPerN + l * 256, kb * 4, &full_p[stage], pol_w);
        if (args.s_rows > 0)
          tma_load_2d(s_dst, &tm_s, n0, ((kb * kStageK) >> args.group_shift), &full_p[stage], pol_x);
      }
      __syncwarp();
      if (++stage == kStages) stage = 0, phase ^= 1;
    }
  } else if (warp == 1) {
    // whole warp runs the loop (uniform operands); one elected lane issues (see scaled_mm.cu)
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = instr_desc<KIND>(kTileN, BM);
      int stage = 0, as = 0;
      uint32_t aphase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        // a full_a arrival implies the stage's full_p: the dequant warps waited on it first
        mbar_wait_bounded(&full_a[as], aphase);
        fence_after_sync();
        const uint64_t dx = smem_desc_sw128(smem_u32(stages + stage * P::kStageBytes));
        if (leader) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_ts_f16(tmem_base + kTmemDCol, tmem_base + as * kTmemAStageCols + k * 8,
                       desc_advance(dx, k * 32), idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          mma_commit(&empty_a[as]);
          mma_commit(&empty_p[stage]);
        }
        if (++stage == kStages) stage = 0;
        if (++as == kAStages) as = 0, aphase ^= 1;
      }
      if (leader) mma_commit(tmem_full);
      // main loop issued: the next kernel in the stream may become resident (its prologue and, for
      // a GEMM, its first weight stages overlap this grid's epilogue) and wait for our completion
      pdl_launch_dependents();
    }
  } else {
    const int qd = warp & 3;          // TMEM quadrant = rows 32 qd .. 32 qd + 31
    const int kt0 = ((warp - 2) >> 2) * kKtPerWarp;  // first 16-k sub-step of the stage this warp takes
    const int nt = qd >> 1;           // 64-column Marlin tile inside the 128-row n tile
    const int j0 = (qd & 1) * 2;      // first of its two 16-row fragments
    const int a_row = lane >> 2;
    const bool nt_valid = nt * 64 < n_valid;
    uint32_t sc[4];  // sc[2*jj + b]: scale of n = 64 nt + 16 (j0 + jj) + a_row + 8 b, in both halves
#pragma unroll
    for (int i = 0; i < 4; ++i) sc[i] = DT<T>::pack2(1.f, 1.f);
    if (!grouped && nt_valid) {
      const uint16_t* s = reinterpret_cast<const uint16_t*>(args.scales);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int n = n0 + nt * 64 + 16 * (j0 + (i >> 1)) + a_row + 8 * (i & 1);
        const uint32_t v = s[scale_pos_single(n)];
        sc[i] = v | (v << 16);
        if (BITS == 8 && args.fp8) sc[i] = fp8_scale_bias<T>(sc[i]);
      }
    }
    // one group per 64-k stage (group_size >= 64; smaller groups take the shared-memory kernel)
    int cur_group = -1;
    const uint32_t lane_src = smem_u32(stages) + P::kXBytes + chunk_off<BITS>(0, nt) + lane * (BITS * 4) +
                              (qd & 1) * (BITS * 2);
    const uint32_t scale_src = smem_u32(stages) + P::kXBytes + P::kPackedBytes + (nt * 64 + 8 * a_row) * 2 +
                               (qd & 1) * 8;
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
    // The k loop is unrolled over one full trip of the stage ring, so the stage index, the TMEM
    // A-ring slot and its phase are compile-time constants (kStages is a multiple of kAStages: every
    // outer iteration starts with the A ring at slot 0, phase 0) -- no ring bookkeeping, constant
    // shared-memory / TMEM offsets.
    static_assert(kStages % kAStages == 0, "the unrolled dequant loop needs whole A-ring wraps");
    uint32_t phase = 0;
    for (int kb0 = kb_begin; kb0 < kb_end; kb0 += kStages, phase ^= 1) {
#pragma unroll
     for (int u = 0; u < kStages; ++u) {
      const int kb = kb0 + u;
      if (kb >= kb_end) break;
      const int stage = u, as = u % kAStages;
      const uint32_t aphase = (u / kAStages) & 1;
      mbar_wait_bounded(&full_p[stage], phase);
      const uint32_t w_src = lane_src + stage * P::kStageBytes;
      uint32_t words[kKtPerWarp][BITS == 4 ? 2 : 4];
#pragma unroll
      for (int kk = 0; kk < kKtPerWarp; ++kk) {
        if constexpr (BITS == 4) {
          const uint2 v = lds64(w_src + (kt0 + kk) * 1024);
          words[kk][0] = v.x, words[kk][1] = v.y;
        } else {
          const uint4 v = lds128(w_src + (kt0 + kk) * 1024);
          words[kk][0] = v.x, words[kk][1] = v.y, words[kk][2] = v.z, words[kk][3] = v.w;
        }
      }
      if (grouped) {
        const int g = ((kb * kStageK) >> args.group_shift);
        if (g != cur_group) {  // warp-uniform
          cur_group = g;
          const uint2 v = lds64(scale_src + stage * P::kStageBytes);
          sc[0] = dup16(v.x, 0), sc[1] = dup16(v.x, 1), sc[2] = dup16(v.y, 0), sc[3] = dup16(v.y, 1);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_p[stage]);  // packed words + scales are in registers
      mbar_wait_bounded(&empty_a[as], aphase ^ 1);
      fence_after_sync();
      const uint32_t t_dst = t_lane + as * kTmemAStageCols;
#pragma unroll
      for (int kk = 0; kk < kKtPerWarp; ++kk) {
        const int kt = kt0 + kk;
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          uint32_t rg[4];
          if constexpr (BITS == 4) {
            unpack4<T>(words[kk][jj], rg);
          } else {
            if (args.fp8) {
              unpack8_fp8<T>(words[kk][2 * jj], rg[0], rg[1]);
              unpack8_fp8<T>(words[kk][2 * jj + 1], rg[2], rg[3]);
            } else {
              unpack8<T>(words[kk][2 * jj], rg[0], rg[1]);
              unpack8<T>(words[kk][2 * jj + 1], rg[2], rg[3]);
            }
          }
          rg[0] = mul2<T>(rg[0], sc[2 * jj]);
          rg[1] = mul2<T>(rg[1], sc[2 * jj]);
          rg[2] = mul2<T>(rg[2], sc[2 * jj + 1]);
          rg[3] = mul2<T>(rg[3], sc[2 * jj + 1]);
          // rg = {b0 k-lo, b0 k-hi, b1 k-lo, b1 k-hi} -> st order {b0 lo, b1 lo, b0 hi, b1 hi}
          tmem_st_16x128b_x2(t_dst + ((uint32_t)(16 * jj) << 16) + kt * 8, rg[0], rg[2], rg[1], rg[3]);
        }
      }
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[as]);
     }
    }
    {
      const int n = n0 + qd * 32 + lane;
      mbar_wait_bounded(tmem_full, 0);
      fence_after_sync();
      const uint32_t taddr = t_lane + kTmemDCol;
      const bool split_k = args.partial != nullptr;
#pragma unroll 1
      for (int c = (warp - 2) >> 2; c < BM / 16; c += kTDequantWarps / 4) {  // token chunks interleaved over the quadrant's warps
        if (m0 + c * 16 >= args.m) break;
        uint32_t rr[16];
        tmem_ld16(taddr + c * 16, rr);
        tmem_ld_wait();
        if (n < args.n) {
          if (split_k) {
            float* ws = args.partial + ((int64_t)split * args.m) * args.n;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int mm = m0 + c * 16 + j;
              if (mm < args.m) ws[(int64_t)mm * args.n + n] = __uint_as_float(rr[j]);
            }
          } else {
            T* out = reinterpret_cast<T*>(args.c);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int mm = m0 + c * 16 + j;
              if (mm < args.m) out[(int64_t)mm * args.n + n] = DT<T>::from_f(__uint_as_float(rr[j]));
            }
          }
        }
      }
      fence_before_sync();
      if (split_k && args.locks != nullptr) {
        // the last split to arrive at this n tile sums the partials (fixed order: deterministic)
        // and hands the counter back as zero -- no second kernel behind the GEMM
        uint32_t* flag = tmem_slot + 1;
        __threadfence();
        asm volatile("bar.sync 1, %0;" ::"n"(kTDequantWarps * 32) : "memory");
        if (warp == 2 && lane == 0) {
          const int old = atomicAdd(&args.locks[blockIdx.y], 1);
          *flag = (old == (int)gridDim.z - 1) ? 1u : 0u;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kTDequantWarps * 32) : "memory");
        if (*flag != 0 && n < args.n) {
          __threadfence();
          const int splits = gridDim.z;
          const int64_t plane = (int64_t)args.m * args.n;
          T* out = reinterpret_cast<T*>(args.c);
          const int m_end = min(m0 + BM, args.m);
          for (int mm = m0; mm < m_end; ++mm) {
            const float* src = args.partial + (int64_t)mm * args.n + n;
            float acc = 0.f;
            int sidx = 0;
            for (; sidx + 8 <= splits; sidx += 8) {
              float v[8];
#pragma unroll
              for (int u = 0; u < 8; ++u) v[u] = __ldcg(src + (int64_t)(sidx + u) * plane);
#pragma unroll
              for (int u = 0; u < 8; ++u) acc += v[u];
            }
            for (; sidx < splits; ++sidx) acc += __ldcg(src + (int64_t)sidx * plane);
            out[(int64_t)mm * args.n + n] = DT<T>::from_f(acc);
          }
          if (warp == 2 && lane == 0) args.locks[blockIdx.y] = 0;
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, P::kTmemCols);
  }
}

// act-order: a_tmp[m, k'] = a[m, perm[k']] (gptq_marlin.cu:345-394 permute_cols_kernel)
template <typename T>
__global__ void __launch_bounds__(256) permute_cols_kernel(T* __restrict__ dst, const T* __restrict__ src,
                                                           const int32_t* __restrict__ perm, int m,
                                                           int k, int64_t lda) {
  const int row = blockIdx.x;
  if (row >= m) return;
  for (int c = threadIdx.x; c < k; c += blockDim.x) dst[(int64_t)row * k + c] = src[(int64_t)row * lda + perm[c]];
}

template <typename T>
__global__ void __launch_bounds__(256) splitk_sum_kernel(T* __restrict__ out, const float* __restrict__ ws,
                                                         int64_t total, int splits) {
  pdl_wait();                // launched as a programmatic dependent of the GEMM: partials complete
  pdl_launch_dependents();   // (a no-op wait / trigger for a normal launch)
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < total;
       i += (int64_t)gridDim.x * blockDim.x * 4) {
    float4 acc = *reinterpret_cast<const float4*>(ws + i);
    for (int s = 1; s < splits; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(ws + (int64_t)s * total + i);
      acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
    }
    uint2 o;
    o.x = DT<T>::pack2(acc.x, acc.y);
    o.y = DT<T>::pack2(acc.z, acc.w);
    *reinterpret_cast<uint2*>(out + i) = o;
  }
}

// test hook: materialise the dequantised weight [K, N] with the SAME unpack/scale code the GEMM
// uses, so tests can check it bit-for-bit against the reference's w_ref.
template <int BITS, typename T>
__global__ void __launch_bounds__(32) marlin_dequant_kernel(T* __restrict__ out, const uint32_t* __restrict__ w,
                                                            const uint16_t* __restrict__ scales, int k,
                                                            int n, int num_groups) {
  const int kt = blockIdx.y, nt = blockIdx.x, lane = threadIdx.x;
  constexpr int kWords = 16 * 64 * BITS / 32;
  const uint32_t* src = w + (int64_t)kt * (n * 16 * BITS / 32) + (int64_t)nt * kWords + lane * (BITS == 4 ? 4 : 8);
  const int group_size = k / num_groups;
  const int g = (kt * 16) / group_size;
  const int a_row = lane >> 2;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t rg[4];
    if constexpr (BITS == 4) {
      unpack4<T>(src[j], rg);
    } else {
      unpack8<T>(src[2 * j], rg[0], rg[1]);
      unpack8<T>(src[2 * j + 1], rg[2], rg[3]);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int nn = nt * 64 + 16 * j + a_row + 8 * (e >> 1);
      const int kk = kt * 16 + 2 * (lane & 3) + ((e & 1) << 3);
      const uint32_t s = num_groups > 1 ? scales[(int64_t)g * n + scale_pos_grouped(nn)]
                                        : scales[scale_pos_single(nn)];
      const uint32_t v = mul2<T>(rg[e], s | (s << 16));
      reinterpret_cast<uint16_t*>(out)[(int64_t)kk * n + nn] = (uint16_t)(v & 0xffff);
      reinterpret_cast<uint16_t*>(out)[(int64_t)(kk + 1) * n + nn] = (uint16_t)(v >> 16);
    }
  }
}

constexpr int kLockMaxM = 1;  // largest M whose split-K partials are reduced in-kernel by the last CTA to arrive

inline int pick_bm(int m) { return m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : m <= 128 ? 128 : 256; }

inline int pick_nt(int bm) { return bm >= 128 ? 2 : 1; }

// 64 < M <= 256 may still take the decode kernel (registers -> TMEM dequant, BM = 64, several m
// tiles) instead of the prefill kernel (shared-memory-staged dequant, 256-wide tiles, one CTA per
// SM).  The decode kernel re-reads and re-dequantises the weights once per m tile, so it wins
// while K x N x m_tiles is small -- measured on the Llama-3-8B layers (us, prefill -> decode kernel):
// M = 128: qkv 18.6 -> 13.8, o 17.0 -> 11.6, down 25.7 -> 25.3, gate_up 40.4 -> 47.1 (kept on the
// prefill kernel); M = 256: qkv 29.2 -> 24.4, o 26.6 -> 17.2, down 39.5 -> 43.3 (kept), gate_up
// 59.8 -> 84.4 (kept).  B200_MARLIN_TMEM_MAX_WORK (weights x m tiles, in millions) overrides.
inline int decode_bm(int m, int n, int k) {
  if (m > 64 && m <= 256) {
    static int knob[2] = {0, 0};
    const int64_t limit = (int64_t)env_int_once("B200_MARLIN_TMEM_MAX_WORK", knob, 200) * 1000000;
    if ((int64_t)k * n * ((m + 63) / 64) <= limit) return 64;
    return 0;  // the prefill kernel (a BM = 128 instantiation of the decode kernel measured the same: 39.6 vs 39.7 us)
  }
  return m <= 64 ? pick_bm(m) : 0;
}

inline int pick_splits(int m, int n, int num_kb, int64_t ws_bytes, int bm_used = 0) {
  const int bm = bm_used > 0 ? bm_used : pick_bm(m);
  const int cta_n = kTileN * pick_nt(bm);
  const int tiles = ((m + bm - 1) / bm) * ((n + cta_n - 1) / cta_n);
  // CTAs that are resident at once: 4 per SM for the decode kernel at BM <= 32, 2 at BM = 64, 1 for
  // the large tiles.  Split K until one resident wave covers the grid (never past it: a second,
  // partial wave costs more than idle SM slots).
  // Measured exception (benchmarks/small_gemm_sweep.py, Llama-3-8B qkv / o / down): at 1 < M <= 16
  // -- the split-K sum kernel instead of the in-kernel reduce of M = 1, short k ranges per CTA --
  // half as many splits win (M = 16: qkv 10.6 -> 9.0 us, o 10.0 -> 7.6, down 16.5 -> 14.8).
  const int per_sm = bm <= 32 ? ((bm == 16 && m > kLockMaxM) ? 2 : 4) : bm <= 64 ? 2 : 1;
  const int target = per_sm * num_sms();
  if (tiles >= target || tiles == 0) return 1;
  int splits = target / tiles;
  {  // tuning hook (benchmarks only), read once per process
    static int knob[2] = {0, 0};
    const int v = env_int_once("B200_MARLIN_SPLITS", knob, 0);
    if (v > 0) splits = v;
  }
  if (splits > num_kb / 4) splits = num_kb / 4;
  const int64_t per = (int64_t)m * n * 4;
  if (ws_bytes >= 0 && per > 0 && splits > ws_bytes / per) splits = (int)(ws_bytes / per);
  return splits < 1 ? 1 : splits;
}

template <int BITS, int BM, int NT, typename T, bool ACT>
int launch(const b200_gptq_marlin_gemm_params& p, const void* a_used, int64_t lda, cudaStream_t st) {
  using P = Plan<BITS, BM, NT>;
  CUtensorMap tm_x;
  int rc = make_tmap_2d(&tm_x, a_used, (uint64_t)p.size_m, (uint64_t)p.size_k, (uint64_t)lda * 2, 2, BM);
  if (rc != B200_OK) return rc;
  MarlinArgs a{};
  a.c = p.c;
  a.w = p.b_q_weight;
  a.scales = p.b_scales;
  a.g_idx = p.g_idx;
  a.m = p.size_m, a.n = p.size_n, a.k = p.size_k;
  a.num_groups = p.num_groups;
  a.group_size = p.num_groups > 1 ? p.size_k / p.num_groups : p.size_k;
  a.group_shift = (a.group_size & (a.group_size - 1)) == 0 ? __builtin_ctz(a.group_size) : -1;
  a.num_kb = p.size_k / kStageK;
  a.fp8 = p.weight_format == 1;
  CUtensorMap tm_w, tm_s;
  rc = make_weight_maps(p, p.num_groups > 1 && !ACT, P::kCtaN, &tm_w, &tm_s, &a);
  if (rc != B200_OK) return rc;
  int splits = pick_splits(p.size_m, p.size_n, a.num_kb, p.splitk_ws ? p.splitk_ws_bytes : 0);
  if ((int64_t)p.size_m * p.size_n % 4 != 0) splits = 1;
  a.kb_per_split = (a.num_kb + splits - 1) / splits;
  splits = (a.num_kb + a.kb_per_split - 1) / a.kb_per_split;
  a.partial = splits > 1 ? reinterpret_cast<float*>(p.splitk_ws) : nullptr;
  const int m_tiles = (p.size_m + BM - 1) / BM, n_tiles = (p.size_n + P::kCtaN - 1) / P::kCtaN;
  a.w_evict_first = (m_tiles == 1);
  auto kern = marlin_gemm_kernel<BITS, BM, NT, T, ACT>;
  static size_t smem_configured[kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, smem_configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_gemm: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  kern<<<dim3(m_tiles, n_tiles, splits), kThreads, P::kBytes, st>>>(tm_x, tm_w, tm_s, a);
  rc = check_launch("gptq_marlin_gemm");
  if (rc != B200_OK) return rc;
  if (splits > 1) {
    const int64_t total = (int64_t)p.size_m * p.size_n;
    int blocks = (int)((total / 4 + 255) / 256);
    if (blocks > num_sms() * 4) blocks = num_sms() * 4;
    if (blocks < 1) blocks = 1;
    splitk_sum_kernel<T><<<blocks, 256, 0, st>>>((T*)p.c, a.partial, total, splits);
    rc = check_launch("gptq_marlin_gemm split-K sum");
  }
  return rc;
}

// tensor maps of the packed weights ([K/16, N * words] int32, box 4 x 256) and of the group scales
// ([groups, N] 16-bit, box s_rows x cta_n); s_rows = scale rows a 64-k stage can touch
inline int make_weight_maps(const b200_gptq_marlin_gemm_params& p, bool staged_scales, int cta_n,
                            CUtensorMap* tm_w, CUtensorMap* tm_s, MarlinArgs* a) {
  const uint64_t words = (uint64_t)p.size_n * (16 * p.num_bits / 32);
  int rc = make_tmap_2d_ex(tm_w, p.b_q_weight, (uint64_t)p.size_k / 16, words, words * 4,
                           CU_TENSOR_MAP_DATA_TYPE_INT32, 256, 4, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc != B200_OK) return rc;
  a->s_rows = 0;
  if (staged_scales) {
    const int gs = a->group_size;
    a->s_rows = a->group_shift >= 0 ? (gs >= kStageK ? 1 : kStageK / gs) : (kStageK / gs + 2 > 4 ? 4 : kStageK / gs + 2);
    return make_tmap_2d_ex(tm_s, p.b_scales, (uint64_t)p.num_groups, (uint64_t)p.size_n, (uint64_t)p.size_n * 2,
                           CU_TENSOR_MAP_DATA_TYPE_UINT16, (uint32_t)cta_n, (uint32_t)a->s_rows,
                           CU_TENSOR_MAP_SWIZZLE_NONE);
  }
  *tm_s = *tm_w;  // never dereferenced
  return B200_OK;
}

// gptq_marlin_gemm_add_rms_norm: when set (by that entry point, for the duration of one call), a
// launch that would run the split-K sum kernel leaves the fp32 partials in the workspace instead
// and reports them here; the caller sums them inside the norm kernel
struct SplitKHandOff {
  const float* partial = nullptr;
  int splits = 0;
};
thread_local SplitKHandOff* tl_splitk_handoff = nullptr;

template <int BITS, int BM, typename T>
int launch_tmem(const b200_gptq_marlin_gemm_params& p, cudaStream_t st) {
  using P = PlanT<BITS, BM>;
  CUtensorMap tm_x;
  int rc = make_tmap_2d(&tm_x, p.a, (uint64_t)p.size_m, (uint64_t)p.size_k, (uint64_t)p.lda * 2, 2, BM);
  if (rc != B200_OK) return rc;
  MarlinArgs a{};
  a.c = p.c;
  a.w = p.b_q_weight;
  a.scales = p.b_scales;
  a.m = p.size_m, a.n = p.size_n, a.k = p.size_k;
  a.num_groups = p.num_groups;
  a.group_size = p.num_groups > 1 ? p.size_k / p.num_groups : p.size_k;
  a.group_shift = (a.group_size & (a.group_size - 1)) == 0 ? __builtin_ctz(a.group_size) : -1;
  a.num_kb = p.size_k / kStageK;
  a.fp8 = p.weight_format == 1;
  CUtensorMap tm_w, tm_s;
  rc = make_weight_maps(p, p.num_groups > 1, kTileN, &tm_w, &tm_s, &a);
  if (rc != B200_OK) return rc;
  int splits = pick_splits(p.size_m, p.size_n, a.num_kb, p.splitk_ws ? p.splitk_ws_bytes : 0, BM);
  if ((int64_t)p.size_m * p.size_n % 4 != 0) splits = 1;
  a.kb_per_split = (a.num_kb + splits - 1) / splits;
  splits = (a.num_kb + a.kb_per_split - 1) / a.kb_per_split;
  a.partial = splits > 1 ? reinterpret_cast<float*>(p.splitk_ws) : nullptr;
  const int m_tiles = (p.size_m + BM - 1) / BM, n_tiles = (p.size_n + kTileN - 1) / kTileN;
  a.w_evict_first = (m_tiles == 1);
  auto kern = marlin_gemm_tmem_kernel<BITS, BM, T>;
  static size_t smem_configured[kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, smem_configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_gemm: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  // M = 1: the last split to arrive reduces in-kernel; otherwise a grid-wide reduce kernel (measured:
  // at M = 2 .. 4 the in-kernel reduce loses -- qkv 11.4 vs 8.4 us, o 8.8 vs 7.0, down 15.9 vs 14.2 at M = 4)
  static int lock_knob[2] = {0, 0};
  const int lock_max_m = env_int_once("B200_MARLIN_LOCK_MAX_M", lock_knob, kLockMaxM);
  a.locks = (splits > 1 && m_tiles == 1 && p.size_m <= lock_max_m) ? p.workspace : nullptr;
  // Programmatic dependent launch: behind a short producer kernel (RMSNorm, SiLU*mul, the attention
  // reduce) this grid's prologue and first ring of weight stages overlap the producer.  (Back to
  // back behind ANOTHER GEMM it was measured to lose -- early-resident dependents take TMEM /
  // shared-memory slots from the running grid, 29 -> 38 us on gate_up -- which the decode chain
  // never does: the glue kernels in between release their dependents only after their own wait.)
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(m_tiles, n_tiles, splits);
  cfg.blockDim = dim3(tmem_threads<BM>());
  cfg.dynamicSmemBytes = P::kBytes;
  cfg.stream = st;
  cudaLaunchAttribute pdl_attr[1];
  pdl_attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  pdl_attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = pdl_attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaError_t le = cudaLaunchKernelEx(&cfg, kern, tm_x, tm_w, tm_s, a);
  // (the split-K sum kernel below is launched as a programmatic dependent too: it waits for the
  // GEMM grid at its top, so only its launch latency overlaps the GEMM's tail)
  if (le != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_gemm: launch failed: %s", cudaGetErrorString(le));
    return B200_ECUDA;
  }
  if (splits > 1 && a.locks == nullptr && tl_splitk_handoff != nullptr) {
    tl_splitk_handoff->partial = a.partial;  // the caller's kernel sums the planes
    tl_splitk_handoff->splits = splits;
    return B200_OK;
  }
  if (splits > 1 && a.locks == nullptr) {
    const int64_t total = (int64_t)p.size_m * p.size_n;
    int blocks = (int)((total / 4 + 255) / 256);
    if (blocks > num_sms() * 4) blocks = num_sms() * 4;
    if (blocks < 1) blocks = 1;
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = 0;
    le = cudaLaunchKernelEx(&cfg, splitk_sum_kernel<T>, (T*)p.c, (const float*)a.partial, total, splits);
    if (le != cudaSuccess) {
      cudaGetLastError();
      set_error("gptq_marlin_gemm split-K sum: launch failed: %s", cudaGetErrorString(le));
      return B200_ECUDA;
    }
  }
  return B200_OK;
}

#include "marlin_gemm_stream.cuh"

// which decode kernel: 0 = stream-K persistent (default), 1 = first-generation split-K kernel
int g_decode_choice_override = -1;  // tests: b200_debug_set_marlin_decode_kernel
inline int decode_kernel_choice() {
  if (g_decode_choice_override >= 0) return g_decode_choice_override;
  static int knob[2] = {0, 0};
  return env_int_once("B200_MARLIN_DECODE_V1", knob, 0);
}

// The stream-K kernel pays where every n tile is shared by at most ~3 CTAs (the split-K fix-up is
// then one or two slot reads by a CTA that is done anyway): layers with at least num_sms / 2 tiles
// of 256 columns, e.g. Llama-3-8B gate_up (20.3 vs 21.6 us at M = 1).  Narrow layers (qkv, o, down:
// 16-24 tiles, each split over ~10 CTAs) stay with the first-generation kernel -- measured,
// profiles/r2_marlin_stream.md.
// B200_MARLIN_DECODE_V1 = 1 forces the first-generation kernel, 2 forces the stream-K kernel.
inline bool stream_kernel_pays(const b200_gptq_marlin_gemm_params& p) {
  const int64_t n_tiles = (p.size_n + kSCtaN - 1) / kSCtaN;
  if ((n_tiles * (p.size_k / kStageK) + 1) * num_sms() >= (1ll << 31)) return false;
  const int choice = decode_kernel_choice();
  if (choice == 1) return false;
  if (choice == 2) return true;
  // (at M = 64 the fp32 pieces are 64 KiB each and the first-generation kernel wins: measured)
  return n_tiles * 2 >= num_sms() && p.size_m <= 32;
}

template <int BITS, typename T, bool ACT>
int dispatch_bm(const b200_gptq_marlin_gemm_params& p, const void* a_used, int64_t lda, cudaStream_t st) {
  // decode kernel: ungrouped, or groups that never change inside a 64-k stage
  // (and power-of-two group sizes, so the group index is a shift: every GPTQ group size is)
  const int gsz = p.num_groups > 1 ? p.size_k / p.num_groups : 0;
  const bool one_group_per_stage = p.num_groups == 1 || (gsz % kStageK == 0 && (gsz & (gsz - 1)) == 0);
  if (!ACT && one_group_per_stage && p.size_m <= 64 && p.workspace != nullptr && p.splitk_ws != nullptr &&
      p.splitk_ws_bytes >= stream_ws_bytes(p.size_m) && stream_kernel_pays(p)) {
    switch (pick_bm(p.size_m)) {
      case 16: return launch_stream<BITS, 16, T>(p, st);
      case 32: return launch_stream<BITS, 32, T>(p, st);
      default: return launch_stream<BITS, 64, T>(p, st);
    }
  }
  if (!ACT && one_group_per_stage) {
    switch (decode_bm(p.size_m, p.size_n, p.size_k)) {
      case 16: return launch_tmem<BITS, 16, T>(p, st);
      case 32: return launch_tmem<BITS, 32, T>(p, st);
      case 64: return launch_tmem<BITS, 64, T>(p, st);
      default: break;
    }
  }
  switch (pick_bm(p.size_m)) {
    case 16: return launch<BITS, 16, 1, T, ACT>(p, a_used, lda, st);
    case 32: return launch<BITS, 32, 1, T, ACT>(p, a_used, lda, st);
    case 64: return launch<BITS, 64, 1, T, ACT>(p, a_used, lda, st);
    case 128: return launch<BITS, 128, 2, T, ACT>(p, a_used, lda, st);
    default: return launch<BITS, 256, 2, T, ACT>(p, a_used, lda, st);
  }
}

template <int BITS, typename T>
int run(const b200_gptq_marlin_gemm_params& p, cudaStream_t st) {
  const bool act = p.g_idx != nullptr;
  const void* a_used = p.a;
  int64_t lda = p.lda;
  if (act) {
    B200_REQUIRE(p.perm != nullptr && p.a_tmp != nullptr, B200_EINVAL,
                 "gptq_marlin_gemm: act-order needs perm and a_tmp");
    permute_cols_kernel<T><<<p.size_m, 256, 0, st>>>((T*)p.a_tmp, (const T*)p.a, p.perm, p.size_m,
                                                     p.size_k, p.lda);
    int rc = check_launch("gptq_marlin_gemm permute_cols");
    if (rc != B200_OK) return rc;
    a_used = p.a_tmp;
    lda = p.size_k;
    return dispatch_bm<BITS, T, true>(p, a_used, lda, st);
  }
  return dispatch_bm<BITS, T, false>(p, a_used, lda, st);
}

}  // namespace
}  // namespace b200

extern "C" {

int64_t b200_gptq_marlin_gemm_workspace_bytes(int32_t m, int32_t n, int32_t k) {
  int splits = b200::pick_splits(m, n, k / b200::kStageK, -1);
  if (m > 64 && b200::decode_bm(m, n, k) > 0) {  // the decode kernel serves this shape
    const int s2 = b200::pick_splits(m, n, k / b200::kStageK, -1, b200::decode_bm(m, n, k));
    if (s2 > splits) splits = s2;
  }
  const int64_t v1 = splits > 1 ? (int64_t)splits * m * n * 4 : 0;
  const int64_t v2 = m <= 64 ? b200::stream_ws_bytes(m) : 0;  // stream-K decode kernel: 2 slots per CTA
  return v1 > v2 ? v1 : v2;
}

int b200_gptq_marlin_gemm(const b200_gptq_marlin_gemm_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "gptq_marlin_gemm: null params");
  const b200_gptq_marlin_gemm_params& p = *pp;
  B200_REQUIRE(p.num_bits == 4 || p.num_bits == 8, B200_EINVAL, "num_bits must be 4 or 8. Got = %d",
               p.num_bits);
  B200_REQUIRE(p.weight_format == 0 || (p.weight_format == 1 && p.num_bits == 8 && p.num_groups == 1 &&
                                        p.g_idx == nullptr),
               B200_EINVAL, "gptq_marlin_gemm: fp8 weights need num_bits 8, channel-wise scales, no act-order");
  B200_REQUIRE(p.size_m >= 0 && p.size_n > 0 && p.size_k > 0, B200_EINVAL, "gptq_marlin_gemm: bad sizes");
  if (p.size_m == 0) return B200_OK;
  B200_REQUIRE(p.size_n % 64 == 0, B200_EINVAL, "size_n = %d, is not divisible by min_thread_n = 64", p.size_n);
  B200_REQUIRE(p.size_k % 64 == 0, B200_ENOTSUP, "size_k = %d is not divisible by 64", p.size_k);
  B200_REQUIRE(p.c && p.a && p.b_q_weight && p.b_scales, B200_EINVAL, "gptq_marlin_gemm: null pointer");
  B200_REQUIRE(p.num_groups >= 1, B200_EINVAL, "gptq_marlin_gemm: num_groups must be >= 1");
  if (p.g_idx == nullptr && p.num_groups > 1) {
    B200_REQUIRE(p.size_k % p.num_groups == 0, B200_EINVAL,
                 "size_k = %d, is not divisible by b_scales.size(0) = %d", p.size_k, p.num_groups);
    B200_REQUIRE((p.size_k / p.num_groups) % 16 == 0, B200_ENOTSUP,
                 "group_size = %d must be a multiple of 16", p.size_k / p.num_groups);
  }
  B200_REQUIRE(((uintptr_t)p.a & 15) == 0 && (p.lda * 2) % 16 == 0, B200_EINVAL,
               "gptq_marlin_gemm: a must be 16-byte aligned with a 16-byte aligned row stride");
  B200_REQUIRE(((uintptr_t)p.b_q_weight & 15) == 0 && ((uintptr_t)p.b_scales & 15) == 0, B200_EINVAL,
               "gptq_marlin_gemm: b_q_weight / b_scales must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.dtype == B200_F16) return p.num_bits == 4 ? run<4, __half>(p, st) : run<8, __half>(p, st);
  if (p.dtype == B200_BF16)
    return p.num_bits == 4 ? run<4, __nv_bfloat16>(p, st) : run<8, __nv_bfloat16>(p, st);
  set_error("gpt_marlin_gemm only supports bfloat16 and float16");
  return B200_EINVAL;
}

// test hook: 0 = heuristic, 1 = first-generation decode kernel, 2 = stream-K kernel wherever it is
// eligible, -1 = back to the environment (B200_MARLIN_DECODE_V1)
void b200_debug_set_marlin_decode_kernel(int32_t choice) { b200::g_decode_choice_override = choice; }

// debug: per-CTA cycle accounting of the stream-K decode kernel (all zeros unless the library was
// built with -DB200_STREAM_PROF).  dst = 160 x 32 uint64 on the host.
// extension (fused decode-side glue): the row-parallel linear followed by
// fused_add_rms_norm (layernorm_kernels.cu:199-249) -- c = norm(gemm(a, w) + residual) * weight,
// residual += gemm(a, w) in place.  Where the GEMM splits K, its fp32 partial planes are summed
// inside the norm kernel (one kernel less, and the GEMM output never goes to HBM in 16 bits);
// otherwise the two operators run back to back.  Bit-identical to the unfused pair either way.
int b200_gptq_marlin_gemm_add_rms_norm(const b200_gptq_marlin_gemm_params* pp, void* residual, const void* norm_weight,
                                       float epsilon, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr && residual != nullptr && norm_weight != nullptr, B200_EINVAL,
               "gptq_marlin_gemm_add_rms_norm: null pointer");
  SplitKHandOff h;
  const bool can_fuse = pp->g_idx == nullptr && splitk_add_rms_eligible(pp->size_n, pp->c, residual, norm_weight);
  tl_splitk_handoff = can_fuse ? &h : nullptr;
  const int rc = b200_gptq_marlin_gemm(pp, stream);
  tl_splitk_handoff = nullptr;
  if (rc != B200_OK || pp->size_m == 0) return rc;
  if (h.splits > 1)
    return splitk_add_rms_norm(pp->dtype, pp->c, h.partial, h.splits, residual, norm_weight, epsilon, pp->size_m,
                               pp->size_n, static_cast<cudaStream_t>(stream));
  return b200_fused_add_rms_norm(pp->c, residual, norm_weight, epsilon, pp->size_m, pp->size_n, pp->dtype, stream);
}

int b200_debug_stream_prof(unsigned long long* dst, int32_t n) {
#ifdef B200_STREAM_PROF
  if (n > 160 * 32) n = 160 * 32;
  return cudaMemcpyFromSymbol(dst, b200::g_stream_prof, (size_t)n * 8) == cudaSuccess ? n : -1;
#else
  (void)dst;
  (void)n;
  return 0;
#endif
}

// fp8_marlin_gemm (csrc/quantization/fp8/fp8_marlin.cu:1212-1308): fp8 e4m3 weights in the Marlin
// 8-bit layout, channel-wise scales; same kernels, other unpack
int b200_fp8_marlin_gemm(const b200_gptq_marlin_gemm_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "fp8_marlin_gemm: null params");
  B200_REQUIRE(pp->num_bits == 8, B200_EINVAL, "num_bits must be 8. Got = %d", pp->num_bits);
  B200_REQUIRE(pp->num_groups == 1, B200_EINVAL, "fp8_marlin_gemm: b_scales must have one row");
  b200_gptq_marlin_gemm_params p = *pp;
  p.weight_format = 1;
  p.g_idx = nullptr;
  p.perm = nullptr;
  return b200_gptq_marlin_gemm(&p, stream);
}

int b200_marlin_dequant(void* out, const uint32_t* b_q_weight, const void* b_scales, int32_t size_k,
                        int32_t size_n, int32_t num_groups, int32_t num_bits, int32_t dtype,
                        void* stream) {
  using namespace b200;
  B200_REQUIRE(size_k % 16 == 0 && size_n % 64 == 0 && num_groups >= 1 && size_k % num_groups == 0 &&
                   (size_k / num_groups) % 16 == 0,
               B200_EINVAL, "marlin_dequant: bad shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid(size_n / 64, size_k / 16);
  const uint16_t* s = reinterpret_cast<const uint16_t*>(b_scales);
  if (dtype == B200_F16 && num_bits == 4)
    marlin_dequant_kernel<4, __half><<<grid, 32, 0, st>>>((__half*)out, b_q_weight, s, size_k, size_n, num_groups);
  else if (dtype == B200_F16 && num_bits == 8)
    marlin_dequant_kernel<8, __half><<<grid, 32, 0, st>>>((__half*)out, b_q_weight, s, size_k, size_n, num_groups);
  else if (dtype == B200_BF16 && num_bits == 4)
    marlin_dequant_kernel<4, __nv_bfloat16><<<grid, 32, 0, st>>>((__nv_bfloat16*)out, b_q_weight, s, size_k, size_n, num_groups);
  else if (dtype == B200_BF16 && num_bits == 8)
    marlin_dequant_kernel<8, __nv_bfloat16><<<grid, 32, 0, st>>>((__nv_bfloat16*)out, b_q_weight, s, size_k, size_n, num_groups);
  else {
    set_error("marlin_dequant: unsupported dtype/bits");
    return B200_EINVAL;
  }
  return check_launch("marlin_dequant");
}

}  // extern "C"
