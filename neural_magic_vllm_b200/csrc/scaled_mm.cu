// W8A8 GEMM with scale epilogue on tcgen05 / TMEM, fed by TMA (sm_100a).
//   out[m, n] = a_scales[m | 0] * (b_scales[n | 0] * sum_k a[m, k] * b[k, n]) (+ bias[n])
// int8 x int8 -> s32 (kind::i8) or e4m3 x e4m3 -> f32 (kind::f8f6f4); out fp16 / bf16.
// Replaces cutlass_scaled_mm (csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:48-100,
// scaled_mm_c3x.cu:100-237, 539-556), whose sm_90a wgmma kernel cannot load on sm_100.
//
// Tile: 128 output features (MMA M, TMEM lanes) x BM tokens (MMA N, TMEM columns) x 128 bytes of
// K per stage.  b is column-major [K, N] = row-major [N, K]: both operands are K-major and go
// straight from TMA (128-byte swizzle) to the tensor core.  Warp roles: 0 = TMA producer,
// 1 = MMA issuer + TMEM owner, 2..5 = epilogue (TMEM -> registers -> scale/bias -> global).
// Small M (decode) is weight streaming: BM shrinks to 16, the grid is filled by split-K
// (fp32 / s32 partials in a caller-provided workspace, reduced by a second small kernel).
#include "umma.cuh"

namespace b200 {
namespace {

using namespace umma;

constexpr int kTileN = 128;      // output features per CTA (MMA M)
constexpr int kStageKBytes = 128;  // bytes of K per pipeline stage (one swizzle atom)
constexpr int kThreads = 192;

struct MmArgs {
  void* out;
  const float* a_scales;
  const float* b_scales;
  const void* bias;
  void* partial;  // split-K workspace [splits, M, N] 32-bit, or nullptr
  int m, n, k;    // k in elements
  int64_t ldc;
  int a_scales_per_token, b_scales_per_channel;
  int kb_per_split, num_kb;
  int w_evict_first;
};

template <int BM>
struct Smem {
  static constexpr int kStageBytes = kTileN * kStageKBytes + BM * kStageKBytes;
  static constexpr int kStages = (200 * 1024 / kStageBytes) > 8 ? 8 : (200 * 1024 / kStageBytes);
  static constexpr int kBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/ + BM * 4;
};

template <typename OutT>
__device__ __forceinline__ OutT cvt_out(float v) {
  return DT<OutT>::from_f(v);
}

template <int KIND, int BM, typename OutT>
__global__ void __launch_bounds__(kThreads, 1)
scaled_mm_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                 const MmArgs args) {
  using S = Smem<BM>;
  constexpr int kStages = S::kStages;
  constexpr uint32_t kTmemCols = BM < 32 ? 32 : BM;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * S::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  float* s_ascale = reinterpret_cast<float*>(smem + kStages * S::kStageBytes + 256);

  // warp-uniform by construction: say so, so the role loops keep their operands in uniform registers
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * kTileN;
  const int split = blockIdx.z;
  const int kb_begin = split * args.kb_per_split;
  const int kb_end = min(kb_begin + args.kb_per_split, args.num_kb);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_x);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  // Producer and MMA issuer run their loops with the whole warp (uniform control flow); only the
  // TMA / MMA instructions themselves sit under elect_one().  Issued from inside an `if (lane == 0)`
  // region their uniform-register operands were rebuilt per instruction (R2UR + a one-trip
  // waterfall loop, ~100 cycles per tcgen05.mma): measured, profiles/r2_marlin_stream.md.
  if (warp == 0) {
    const bool leader = elect_one();
    {
      const uint64_t pol_w = args.w_evict_first ? l2_policy_evict_first() : l2_policy_evict_last();
      const uint64_t pol_x = l2_policy_evict_last();
      // programmatic dependent launch: weights do not depend on the predecessor (the activation
      // quantiser), so the first ring of weight tiles streams in before the wait
      const int pre = min(kb_end - kb_begin, kStages);
      for (int i = 0; i < pre; ++i) {
        if (leader) {
          mbar_arrive_expect_tx(&full_bar[i], S::kStageBytes);
          tma_load_2d(tiles + i * S::kStageBytes, &tm_w, (kb_begin + i) * kStageKBytes, n0, &full_bar[i], pol_w);
        }
      }
      pdl_wait();
      for (int i = 0; i < pre; ++i)
        if (leader)
          tma_load_2d(tiles + i * S::kStageBytes + kTileN * kStageKBytes, &tm_x, (kb_begin + i) * kStageKBytes, m0,
                      &full_bar[i], pol_x);
      int stage = pre == kStages ? 0 : pre;
      uint32_t phase = pre == kStages ? 1 : 0;
      for (int kb = kb_begin + pre; kb < kb_end; ++kb) {
        mbar_wait_bounded(&empty_bar[stage], phase ^ 1);
        uint8_t* w_dst = tiles + stage * S::kStageBytes;
        uint8_t* x_dst = w_dst + kTileN * kStageKBytes;
        if (leader) {
          mbar_arrive_expect_tx(&full_bar[stage], S::kStageBytes);
          tma_load_2d(w_dst, &tm_w, kb * kStageKBytes, n0, &full_bar[stage], pol_w);
          tma_load_2d(x_dst, &tm_x, kb * kStageKBytes, m0, &full_bar[stage], pol_x);
        }
        if (++stage == kStages) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    {
      constexpr uint32_t idesc = instr_desc<KIND>(kTileN, BM);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait_bounded(&full_bar[stage], phase);
        fence_after_sync();
        const uint32_t w_addr = smem_u32(tiles + stage * S::kStageBytes);
        const uint64_t dw = smem_desc_sw128(w_addr);
        const uint64_t dx = smem_desc_sw128(w_addr + kTileN * kStageKBytes);
        if (leader) {
#pragma unroll
          for (int k = 0; k < kStageKBytes / 32; ++k)
            mma_ss<KIND>(tmem_base, desc_advance(dw, k * 32), desc_advance(dx, k * 32), idesc,
                         (kb > kb_begin || k > 0) ? 1u : 0u);
          mma_commit(&empty_bar[stage]);
        }
        if (++stage == kStages) stage = 0, phase ^= 1;
      }
      if (leader) mma_commit(tmem_full);
      pdl_launch_dependents();  // main loop issued: the next kernel may become resident and wait
    }
  } else {
    // ---- epilogue: warp w owns TMEM lanes [32 * (w % 4), +32) ----
    pdl_wait();  // the per-token activation scales are the predecessor's output
    const int quarter = warp & 3;
    const int n = n0 + quarter * 32 + lane;
    const int et = threadIdx.x - 64;  // 0..127
    for (int i = et; i < BM; i += 128) {
      const int mm = m0 + i;
      s_ascale[i] = (mm < args.m) ? args.a_scales[args.a_scales_per_token ? mm : 0] : 0.f;
    }
    float bs = 0.f, bias = 0.f;
    if (n < args.n) {
      bs = args.b_scales[args.b_scales_per_channel ? n : 0];
      if (args.bias != nullptr) bias = DT<OutT>::to_f(reinterpret_cast<const OutT*>(args.bias)[n]);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");  // s_ascale visible to the 4 epilogue warps
    mbar_wait_bounded(tmem_full, 0);
    fence_after_sync();
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BM / 16; ++c) {
      uint32_t r[16];
      tmem_ld16(taddr + c * 16, r);
      tmem_ld_wait();
      if (n < args.n) {
        if (args.partial != nullptr) {
          uint32_t* ws = reinterpret_cast<uint32_t*>(args.partial) +
                         ((int64_t)split * args.m) * args.n;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int mm = m0 + c * 16 + j;
            if (mm < args.m) ws[(int64_t)mm * args.n + n] = r[j];
          }
        } else {
          OutT* out = reinterpret_cast<OutT*>(args.out);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int mm = m0 + c * 16 + j;
            if (mm < args.m) {
              const float acc = (KIND == KIND_I8) ? (float)(int32_t)r[j] : __uint_as_float(r[j]);
              float v = s_ascale[c * 16 + j] * (bs * acc);
              if (args.bias != nullptr) v += bias;
              out[(int64_t)mm * args.ldc + n] = cvt_out<OutT>(v);
            }
          }
        }
      }
    }
    fence_before_sync();
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// split-K second pass: sum the partials in split order, apply the epilogue
template <bool IS_INT, typename OutT>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(OutT* __restrict__ out, const uint32_t* __restrict__ ws,
                     const float* __restrict__ a_scales, const float* __restrict__ b_scales,
                     const OutT* __restrict__ bias, int m, int n, int64_t ldc, int splits,
                     int a_per_token, int b_per_channel) {
  const int64_t total = (int64_t)m * n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int mm = (int)(i / n), nn = (int)(i % n);
    float acc;
    if constexpr (IS_INT) {
      int32_t s = 0;
      for (int sp = 0; sp < splits; ++sp) s += (int32_t)ws[(int64_t)sp * total + i];
      acc = (float)s;
    } else {
      float s = 0.f;
      for (int sp = 0; sp < splits; ++sp) s += __uint_as_float(ws[(int64_t)sp * total + i]);
      acc = s;
    }
    float v = a_scales[a_per_token ? mm : 0] * (b_scales[b_per_channel ? nn : 0] * acc);
    if (bias != nullptr) v += DT<OutT>::to_f(bias[nn]);
    out[(int64_t)mm * ldc + nn] = DT<OutT>::from_f(v);
  }
}

template <int KIND, int BM, typename OutT>
int launch(const b200_scaled_mm_params& p, cudaStream_t st) {
  using S = Smem<BM>;
  const int elem = 1;
  CUtensorMap tm_w, tm_x;
  int rc = make_tmap_2d(&tm_w, p.b, (uint64_t)p.n, (uint64_t)p.k, (uint64_t)p.ldb * elem, elem, kTileN);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d(&tm_x, p.a, (uint64_t)p.m, (uint64_t)p.k, (uint64_t)p.lda * elem, elem, BM);
  if (rc != B200_OK) return rc;
  MmArgs a{};
  a.out = p.out;
  a.a_scales = p.a_scales;
  a.b_scales = p.b_scales;
  a.bias = p.bias;
  a.m = p.m, a.n = p.n, a.k = p.k;
  a.ldc = p.ldc;
  a.a_scales_per_token = p.a_scales_numel > 1;
  a.b_scales_per_channel = p.b_scales_numel > 1;
  a.num_kb = (p.k * elem + kStageKBytes - 1) / kStageKBytes;
  const int m_tiles = (p.m + BM - 1) / BM, n_tiles = (p.n + kTileN - 1) / kTileN;
  const int tiles = m_tiles * n_tiles;
  // split-K only to fill the machine for weight streaming; each split keeps >= 4 k-blocks
  int splits = 1;
  if (p.workspace != nullptr && tiles < num_sms()) {
    splits = (2 * num_sms() + tiles - 1) / tiles / 1;
    if (splits > a.num_kb / 4) splits = a.num_kb / 4;
    const int64_t per_split = (int64_t)p.m * p.n * 4;
    if (per_split > 0 && splits > p.workspace_bytes / per_split) splits = (int)(p.workspace_bytes / per_split);
    if (splits < 1) splits = 1;
  }
  a.kb_per_split = (a.num_kb + splits - 1) / splits;
  splits = (a.num_kb + a.kb_per_split - 1) / a.kb_per_split;
  if (splits < 1) splits = 1;
  a.partial = splits > 1 ? p.workspace : nullptr;
  a.w_evict_first = (m_tiles == 1);
  auto kern = scaled_mm_kernel<KIND, BM, OutT>;
  static size_t smem_configured[kMaxDevices] = {};  // per template instantiation, per device
  if (ensure_dyn_smem(kern, S::kBytes, smem_configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("scaled_mm: cannot reserve %d B of shared memory", (int)S::kBytes);
    return B200_ECUDA;
  }
  dim3 grid(m_tiles, n_tiles, splits);
  launch_pdl(kern, grid, dim3(kThreads), S::kBytes, st, tm_w, tm_x, a);
  rc = check_launch("scaled_mm");
  if (rc != B200_OK) return rc;
  if (splits > 1) {
    const int64_t total = (int64_t)p.m * p.n;
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    splitk_reduce_kernel<KIND == KIND_I8, OutT><<<blocks, 256, 0, st>>>(
        (OutT*)p.out, (const uint32_t*)p.workspace, p.a_scales, p.b_scales, (const OutT*)p.bias, p.m,
        p.n, p.ldc, splits, a.a_scales_per_token, a.b_scales_per_channel);
    rc = check_launch("scaled_mm split-K reduce");
  }
  return rc;
}

template <int KIND, typename OutT>
int dispatch_bm(const b200_scaled_mm_params& p, cudaStream_t st) {
  if (p.m <= 16) return launch<KIND, 16, OutT>(p, st);
  if (p.m <= 32) return launch<KIND, 32, OutT>(p, st);
  if (p.m <= 64) return launch<KIND, 64, OutT>(p, st);
  if (p.m <= 128) return launch<KIND, 128, OutT>(p, st);
  return launch<KIND, 256, OutT>(p, st);
}

}  // namespace
}  // namespace b200

extern "C" {

int b200_scaled_mm_supports_fp8(int32_t cuda_device_capability) {
  return cuda_device_capability >= 89 ? 1 : 0;  // scaled_mm_entry.cu:32-46
}

int64_t b200_scaled_mm_workspace_bytes(int32_t m, int32_t n, int32_t k) {
  // enough for the split-K partials of the weight-streaming regime; 0 when the grid is full
  const int bm = m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : m <= 128 ? 128 : 256;
  const int tiles = ((m + bm - 1) / bm) * ((n + 127) / 128);
  if (tiles >= b200::num_sms() || tiles == 0) return 0;
  int splits = (2 * b200::num_sms() + tiles - 1) / tiles;
  const int num_kb = (k + 127) / 128;
  if (splits > num_kb / 4) splits = num_kb / 4;
  if (splits <= 1) return 0;
  return (int64_t)splits * m * n * 4;
}

int b200_scaled_mm(const b200_scaled_mm_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "scaled_mm: null params");
  const b200_scaled_mm_params& p = *pp;
  B200_REQUIRE(p.m >= 0 && p.n >= 0 && p.k >= 0, B200_EINVAL, "scaled_mm: negative size");
  if (p.m == 0 || p.n == 0) return B200_OK;
  B200_REQUIRE(p.out && p.a && p.b && p.a_scales && p.b_scales, B200_EINVAL, "scaled_mm: null pointer");
  B200_REQUIRE(p.k > 0, B200_EINVAL, "scaled_mm: k must be positive");
  // scaled_mm_entry.cu:66-68: 16-byte aligned rows
  B200_REQUIRE(p.lda % 16 == 0 && p.ldb % 16 == 0 && p.ldc % 8 == 0, B200_EINVAL,
               "scaled_mm: a.stride(0) %% 16, b.stride(1) %% 16 and c.stride(0) %% 16 bytes must be 0");
  B200_REQUIRE(((uintptr_t)p.a & 15) == 0 && ((uintptr_t)p.b & 15) == 0, B200_EINVAL,
               "scaled_mm: a and b must be 16-byte aligned");
  B200_REQUIRE(p.a_scales_numel == 1 || p.a_scales_numel == p.m, B200_EINVAL,
               "scaled_mm: a_scales must have 1 or m elements");
  B200_REQUIRE(p.b_scales_numel == 1 || p.b_scales_numel == p.n, B200_EINVAL,
               "scaled_mm: b_scales must have 1 or n elements");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.qtype == B200_Q_INT8) {
    if (p.out_dtype == B200_F16) return dispatch_bm<KIND_I8, __half>(p, st);
    if (p.out_dtype == B200_BF16) return dispatch_bm<KIND_I8, __nv_bfloat16>(p, st);
  } else if (p.qtype == B200_Q_FP8_E4M3) {
    if (p.out_dtype == B200_F16) return dispatch_bm<KIND_E4M3, __half>(p, st);
    if (p.out_dtype == B200_BF16) return dispatch_bm<KIND_E4M3, __nv_bfloat16>(p, st);
  }
  set_error("scaled_mm: unsupported qtype %d / out dtype %d", p.qtype, p.out_dtype);
  return B200_EINVAL;
}

}  // extern "C"
