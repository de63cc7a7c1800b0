// Variable-length causal attention for prefill on tcgen05 / TMEM (sm_100a) -- SURVEY.md 8(f) rank 3.
// Stands in for the library call the reference's prefill path makes: flash_attn_varlen_func
// (vllm/attention/backends/flash_attn.py:320-346) / xformers memory_efficient_attention_forward
// (backends/xformers.py:340-424); semantics = per-sequence softmax(scale * Q K^T + causal mask) V as
// the reference's CPU backend spells it out (backends/torch_sdpa.py:203-218).  GQA: query head h
// reads kv head h / (H / KVH).  Causal alignment is bottom-right (query i of a sequence with
// q_len <= k_len sees keys <= i + k_len - q_len), as flash-attn does.
//
// One CTA = 256 queries of one head: two 128-row query tiles (A, B) that share every K / V tile and
// ping-pong on the tensor pipe.  Both GEMMs run on tcgen05 with fp32 accumulators in TMEM:
//   S = Q K^T   A = Q tile (shared memory, K-major, TMA 128B swizzle), B = K tile (same layout)
//   O += P V    A = P (fp16/bf16, written by the softmax warps straight into TMEM, over the S
//               columns it came from), B = V tile in its natural [key][d] layout = MN-major operand
//               (descriptor: 64-d column blocks LBO apart, 8-key groups 1 KiB apart)
//   warp 0      TMA producer: both Q tiles once, then K / V tiles of 128 keys through two-stage rings
//   warp 1      MMA issuer (one elected lane), per step j: P V of tile A, Q K^T of tile A for step
//               j + 1, then the same for tile B -- the tensor pipe works for one tile while the
//               other tile's softmax warps are busy
//   warps 2-5 / 6-9  softmax of tile A / B, one thread per query row (TMEM lane = row: row max /
//               sum need no shuffles): the S row is read once into registers (max, then exp2 +
//               pack into P).  O accumulates in TMEM over the kv tiles; it is rescaled -- by these
//               same warps, between two P V of their tile -- only when a row maximum has grown
//               by more than 2^8 (lazy rescale), which after the first tiles is rare.
// TMEM: S_A, S_B (2 x 128 columns, P aliased inside) + O of A and B (2 x 128) = 512 columns.
#include "umma.cuh"

namespace b200 {
namespace {

using namespace umma;

// Per-role cycle accounting (debug builds with -DB200_FA_PROF, read back through b200_debug_fa_prof):
// totals over all CTAs of where the MMA issuer and one softmax warp of tile A spend their time.
#ifdef B200_FA_PROF
__device__ unsigned long long g_fa_prof[32];
#define FPROF_DECL long long fp_t = clock64(), fp_n; unsigned long long fp_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define FPROF(i) (fp_n = clock64(), fp_acc[i] += (unsigned long long)(fp_n - fp_t), fp_t = fp_n)
#define FPROF_FLUSH(base)                                                             \
  do {                                                                                \
    for (int fp_i = 0; fp_i < 8; ++fp_i) atomicAdd(&g_fa_prof[(base) + fp_i], fp_acc[fp_i]); \
  } while (0)
#else
#define FPROF_DECL
#define FPROF(i)
#define FPROF_FLUSH(base)
#endif

constexpr int kBQ = 128;   // queries per CTA
constexpr int kBK = 128;   // keys per tile
constexpr int kFaThreads = 10 * 32;  // producer, MMA issuer, 2 x 4 softmax warps

struct FaArgs {
  void* out;
  const int32_t* cu_q;
  const int32_t* cu_k;
  int num_heads, num_kv_heads;
  int64_t out_stride;  // elements between consecutive tokens of out
  float scale_log2;    // softmax scale * log2(e)
  int causal;
};

template <int D>
struct FaPlan {
  static constexpr int kTileBytes = kBQ * D * 2;  // one Q / K / V tile: D / 64 boxes of 128 x 128 B
  static constexpr int kBoxBytes = kBQ * 128;
  static constexpr int kBytes = 6 * kTileBytes + 1024 + 256;  // 2 Q + 2 K + 2 V
};

// MN-major operand tile (rows = k index, 128 bytes = 64 elements of the N dimension per row, 8-row
// swizzle atoms 1 KiB apart, 64-element column blocks `lbo_bytes` apart), SWIZZLE_128B.
// Canonical form ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units (cute/atom/mma_traits_sm100.hpp).
__device__ __forceinline__ uint64_t smem_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((1024u >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
// tcgen05.wait::ld that also names the destination registers of the load it waits for: with the
// next chunk's load already in flight, nothing but a data dependency keeps the compiler from
// scheduling a consumer of these registers above the wait
__device__ __forceinline__ void tmem_ld_wait_on(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait_on(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
// packed fp32 pairs (FFMA2 / FADD2): one issue slot for two elements
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{\n.reg .b64 ra, rb, rc, rd;\n"
      "mov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %4};\nmov.b64 rc, {%5, %5};\n"
      "fma.rn.ftz.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1) {  // (d0, d1) += (a0, a1)
  asm volatile("{\n.reg .b64 ra, rd;\n"
      "mov.b64 ra, {%2, %3};\nmov.b64 rd, {%0, %1};\n"
      "add.rn.ftz.f32x2 rd, rd, ra;\nmov.b64 {%0, %1}, rd;\n}"
      : "+f"(d0), "+f"(d1)
      : "f"(a0), "f"(a1));
}
__device__ __forceinline__ void fma2v(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  asm("{\n.reg .b64 ra, rb, rc, rd;\n"
      "mov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmov.b64 rc, {%6, %7};\n"
      "fma.rn.ftz.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void add2v(float& d0, float& d1, float a0, float a1, float b) {
  asm("{\n.reg .b64 ra, rb, rd;\n"
      "mov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %4};\n"
      "add.rn.ftz.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b));
}
// exp2 of a pair on the FMA / ALU pipes (the MUFU unit does 4 lanes per clock per sub-partition and
// is the softmax bottleneck): round-to-nearest split x = n + f through the 1.5 * 2^23 magic add,
// degree-3 minimax polynomial for 2^f on [-0.5, 0.5] (max relative error 7.5e-5, below the fp16 /
// bf16 rounding of P), n added into the exponent field.  x <= 8 by construction (lazy rescale).
__device__ __forceinline__ void ex2_poly2(float& p0, float& p1, float x0, float x1) {
  constexpr float kMagic = 12582912.f;
  x0 = fmaxf(x0, -125.f);
  x1 = fmaxf(x1, -125.f);
  float t0, t1, n0, n1, f0, f1, r0, r1;
  add2v(t0, t1, x0, x1, kMagic);
  add2v(n0, n1, t0, t1, -kMagic);
  fma2v(f0, f1, n0, n1, -1.f, -1.f, x0, x1);
  fma2v(r0, r1, f0, f1, 0.0551716685f, 0.0551716685f, 0.2426111251f, 0.2426111251f);
  fma2v(r0, r1, r0, r1, f0, f1, 0.6932609677f, 0.6932609677f);
  fma2v(r0, r1, r0, r1, f0, f1, 0.9999280572f, 0.9999280572f);
  p0 = __int_as_float(__float_as_int(r0) + (__float_as_int(t0) << 23));
  p1 = __int_as_float(__float_as_int(r1) + (__float_as_int(t1) << 23));
}
#ifndef B200_FA_POLY_PAIRS
#define B200_FA_POLY_PAIRS 0  // of every 16 column pairs, this many take the polynomial instead of MUFU
#endif
// volatile twins: keep the hand-staged order of the unmasked softmax pass
__device__ __forceinline__ float ex2v(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void tmem_st16v(uint32_t taddr, const uint32_t (&r)[16]) { tmem_st16(taddr, r); }
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T, int D, int POLY>
// (10 warps are allocated as 12: 168 registers per thread is the ceiling for this block size)
__global__ void __launch_bounds__(kFaThreads, 1)
varlen_attention_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                        const __grid_constant__ CUtensorMap tm_v, const FaArgs args) {
  using P = FaPlan<D>;
  constexpr int KIND = DT<T>::id == B200_F16 ? KIND_F16 : KIND_BF16;
  constexpr int kBoxes = D / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* q_s = smem;                      // [2] query tiles A, B
  uint8_t* k_s = q_s + 2 * P::kTileBytes;   // [2] stages
  uint8_t* v_s = k_s + 2 * P::kTileBytes;   // [2] stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(v_s + 2 * P::kTileBytes);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;    // [2] stage
  uint64_t* v_full = bars + 3;    // [2] stage
  uint64_t* k_empty = bars + 5;   // [2] stage
  uint64_t* v_empty = bars + 7;   // [2] stage
  uint64_t* s_full = bars + 9;    // [2] query tile
  uint64_t* p_full = bars + 11;   // [2][2] query tile, half of the P row (keys 0-63 / 64-127)
  uint64_t* o_full = bars + 15;   // [2] query tile
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int seq = blockIdx.y, head = blockIdx.x;
  const int q_begin = args.cu_q[seq], q_len = args.cu_q[seq + 1] - q_begin;
  const int k_begin = args.cu_k[seq], k_len = args.cu_k[seq + 1] - k_begin;
  // first query of tile A inside the sequence (tile B = q0 + 128).  Causal work grows with q0 and
  // CTAs are dispatched x-fastest, then y, then z: with the query tile in z (reversed), the longest
  // tiles of ALL heads and sequences go first (longest-processing-time order over the whole grid,
  // not per head), and the heads of one kv group run side by side and share K / V in L2
  const int q0 = (int)(gridDim.z - 1 - blockIdx.z) * (2 * kBQ);
  if (q0 >= q_len) return;                // uniform over the CTA: nothing allocated yet
  const int shift = k_len - q_len;        // bottom-right causal alignment
  // kv tiles each query tile walks: keys [0, k_end_t)
  int nt[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int qt = q0 + t * kBQ;
    const int k_end = qt >= q_len ? 0 : (args.causal ? max(0, min(k_len, min(qt + kBQ, q_len) + shift)) : k_len);
    nt[t] = (k_end + kBK - 1) / kBK;
  }
  const int ntmax = max(nt[0], nt[1]);
  const int kv_head = head / (args.num_heads / args.num_kv_heads);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[2 * i], 4);
      mbar_init(&p_full[2 * i + 1], 4);
      mbar_init(&o_full[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  // TMEM columns: S of tile t (P aliased inside) at 128 t, O_t of tile t at 256 + 128 t
  constexpr uint32_t kOCol = 256;

  if (warp == 0) {
    // ======================= TMA producer =======================
    const bool leader = elect_one();
    const uint64_t pol = l2_policy_evict_last();  // K / V are re-read by the other heads of the group
    pdl_wait();
    if (leader) {
      mbar_arrive_expect_tx(q_full, 2 * P::kTileBytes);  // rows past the tensor end arrive as zeros
#pragma unroll
      for (int t = 0; t < 2; ++t)
#pragma unroll
        for (int b = 0; b < kBoxes; ++b)
          tma_load_2d(q_s + t * P::kTileBytes + b * P::kBoxBytes, &tm_q, head * D + b * 64,
                      q_begin + q0 + t * kBQ, q_full, pol);
    }
    for (int j = 0; j < ntmax; ++j) {
      const int st = j & 1;
      const uint32_t ph = ((j >> 1) & 1) ^ 1;
      mbar_wait_bounded(&k_empty[st], ph);
      if (leader) {
        mbar_arrive_expect_tx(&k_full[st], P::kTileBytes);
#pragma unroll
        for (int b = 0; b < kBoxes; ++b)
          tma_load_2d(k_s + st * P::kTileBytes + b * P::kBoxBytes, &tm_k, kv_head * D + b * 64,
                      k_begin + j * kBK, &k_full[st], pol);
      }
      mbar_wait_bounded(&v_empty[st], ph);
      if (leader) {
        mbar_arrive_expect_tx(&v_full[st], P::kTileBytes);
#pragma unroll
        for (int b = 0; b < kBoxes; ++b)
          tma_load_2d(v_s + st * P::kTileBytes + b * P::kBoxBytes, &tm_v, kv_head * D + b * 64,
                      k_begin + j * kBK, &v_full[st], pol);
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    // Ping-pong over the two query tiles: while the softmax warps of one tile work on S_t(j), the
    // tensor pipe runs P V of the other tile's step j and its Q K^T of step j + 1.
    const bool leader = elect_one();
    FPROF_DECL;
    constexpr uint32_t idesc_qk = instr_desc<KIND>(kBQ, kBK);             // S[128 q, 128 keys]
    constexpr uint32_t idesc_pv = instr_desc<KIND>(kBQ, D) | (1u << 16);  // O_t[128 q, D], B MN-major
    auto issue_qk = [&](int t, int j) {
      const int st = j & 1;
      FPROF(0);
      mbar_wait_bounded(&k_full[st], (j >> 1) & 1);
      FPROF(1);
      fence_after_sync();
      const uint32_t qa = smem_u32(q_s + t * P::kTileBytes), ka = smem_u32(k_s + st * P::kTileBytes);
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * P::kBoxBytes + (kk & 3) * 32;
          mma_ss<KIND>(tmem_base + t * kBK, smem_desc_sw128(qa + off), smem_desc_sw128(ka + off), idesc_qk,
                       kk > 0 ? 1u : 0u);
        }
        mma_commit(&s_full[t]);
      }
    };
    auto issue_pv = [&](int t, int j) {
      const int st = j & 1;
      FPROF(0);
      // P arrives in two halves: the P V MMAs of keys 0-63 run while the softmax warps are still
      // in the exponentials of keys 64-127 (the O rescale, when the row maxima moved, precedes both)
      mbar_wait_bounded(&p_full[2 * t], j & 1);
      FPROF(2 + t);
      mbar_wait_bounded(&v_full[st], (j >> 1) & 1);
      FPROF(1);
      fence_after_sync();
      const uint32_t va = smem_u32(v_s + st * P::kTileBytes);
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < kBK / 32; ++kk)  // 16 keys per MMA: 8 packed P columns, 16 V rows
          mma_ts_f16(tmem_base + kOCol + t * D, tmem_base + t * kBK + kk * 8,
                     smem_desc_mn_sw128(va + kk * 2048, P::kBoxBytes), idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
      }
      FPROF(0);
      mbar_wait_bounded(&p_full[2 * t + 1], j & 1);
      FPROF(2 + t);
      fence_after_sync();
      if (leader) {
#pragma unroll
        for (int kk = kBK / 32; kk < kBK / 16; ++kk)
          mma_ts_f16(tmem_base + kOCol + t * D, tmem_base + t * kBK + kk * 8,
                     smem_desc_mn_sw128(va + kk * 2048, P::kBoxBytes), idesc_pv, 1u);
        if (j + 1 == nt[t]) mma_commit(&o_full[t]);
      }
    };
    mbar_wait_bounded(q_full, 0);
    // the key tile of step j is free once BOTH tiles' Q K^T of step j are issued (commit covers
    // every MMA issued before it); likewise the value tile after both P V
    for (int t = 0; t < 2; ++t)
      if (nt[t] > 0) issue_qk(t, 0);
    if (ntmax > 0 && leader) mma_commit(&k_empty[0]);
    for (int j = 0; j < ntmax; ++j) {
      for (int t = 0; t < 2; ++t) {
        if (j < nt[t]) issue_pv(t, j);
        if (j + 1 < nt[t]) issue_qk(t, j + 1);
      }
      if (leader) {
        mma_commit(&v_empty[j & 1]);
        if (j + 1 < ntmax) mma_commit(&k_empty[(j + 1) & 1]);
      }
    }
    FPROF(0);
    if (lane == 0) FPROF_FLUSH(0);
    pdl_launch_dependents();
  } else {
    // ======================= softmax + output accumulation (thread = query row) =======================
    // warps 2-5 own query tile A, warps 6-9 tile B; a warp reaches the TMEM lanes of quadrant warp % 4
    const int t = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;           // TMEM lane
    const int qi = q0 + t * kBQ + row;        // query index inside the sequence
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t s_addr = t_lane + t * kBK;
    const uint32_t o_addr = t_lane + kOCol + t * D;
    const float c = args.scale_log2;
    const int my_nt = nt[t];
    // keys this row may see: [0, row_end)
    const int row_end = args.causal ? min(k_len, qi + shift + 1) : k_len;
    // O accumulates in TMEM across the kv tiles (P V with the accumulate flag).  The exponent
    // reference m_ref follows the running row maximum lazily: it moves only when the maximum has
    // grown by more than 2^8 in the exp2 domain, and only then is O (and l) rescaled -- p stays
    // <= 256, exact enough for fp16 / bf16 P and fp32 sums, and o / l is the same quotient.
    // S_t(j) arriving implies P V of step j - 1 has finished (the tensor pipe runs in issue order and
    // the commit covers everything issued before it), so O may be touched right after s_full.
    float m_ref = -1e30f, l = 0.f;
    FPROF_DECL;
    for (int j = 0; j < my_nt; ++j) {
      const int lim = row_end - j * kBK;  // columns < lim are visible to this row
      // no masked column in this tile for any row of the warp (all but the diagonal / tail tiles);
      // warp-uniform, because the TMEM loads / stores below are warp-collective
      const bool warp_full = __all_sync(0xffffffffu, lim >= kBK);
      FPROF(0);
      mbar_wait_bounded(&s_full[t], j & 1);
      FPROF(1);
      fence_after_sync();
      // the S row in registers (one TMEM read)
      uint32_t sr[4][32];
      float mx0 = -1e30f, mx1 = -1e30f, mx2 = -1e30f, mx3 = -1e30f;
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) tmem_ld32(s_addr + cc * 32, sr[cc]);
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        tmem_ld_wait_on(sr[cc]);  // the first one waits for all four loads
        if (warp_full) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {  // four chains of 3-input max
            mx0 = fmaxf(mx0, fmaxf(__uint_as_float(sr[cc][i]), __uint_as_float(sr[cc][i + 1])));
            mx1 = fmaxf(mx1, fmaxf(__uint_as_float(sr[cc][i + 2]), __uint_as_float(sr[cc][i + 3])));
            mx2 = fmaxf(mx2, fmaxf(__uint_as_float(sr[cc][i + 4]), __uint_as_float(sr[cc][i + 5])));
            mx3 = fmaxf(mx3, fmaxf(__uint_as_float(sr[cc][i + 6]), __uint_as_float(sr[cc][i + 7])));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (cc * 32 + i < lim) mx0 = fmaxf(mx0, __uint_as_float(sr[cc][i]));
        }
      }
      FPROF(2);
      const float m_new = fmaxf(fmaxf(m_ref, mx0), fmaxf(mx1, fmaxf(mx2, mx3)));
      const bool moved = (m_new - m_ref) * c > 8.f;
      if (__any_sync(0xffffffffu, moved)) {
        const float alpha = moved ? ex2((m_ref - m_new) * c) : 1.f;
        if (moved) m_ref = m_new;
        l *= alpha;
        if (j > 0) {  // at j = 0 the first P V overwrites O
#pragma unroll 1
          for (int cc = 0; cc < D / 16; ++cc) {
            uint32_t r[16];
            tmem_ld16(o_addr + cc * 16, r);
            tmem_ld_wait_on(r);
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st16(o_addr + cc * 16, r);
          }
        }
      }
      FPROF(3);
      const float mc = m_ref * c;
      auto p_arrive = [&](int half) {  // this warp's P columns of one half are in TMEM
        tmem_st_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[2 * t + half]);
      };
      // p = exp2(s c - m_ref c), row sum, packed 16-bit P over the S columns
      float sum0 = 0.f, sum1 = 0.f;
      if (warp_full) {
        // Staged by hand so that a MUFU result is consumed a whole 32-column chunk (>= 256 MUFU
        // cycles) after it was issued: with one softmax warp per tile on a sub-partition, a
        // consumer right behind its ex2 stalls for the MUFU latency on every pair.
        float sumv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        auto exps = [&](int cc) {  // sr[cc] <- exp2(sr[cc] * c - mc), in place
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float x0, x1;
            fma2(x0, x1, __uint_as_float(sr[cc][i]), __uint_as_float(sr[cc][i + 1]), c, -mc);
            if (POLY < 0) {
            } else if ((i >> 1) * POLY % 16 < POLY) {  // evenly spread over the chunk
              ex2_poly2(x0, x1, x0, x1);
            } else {
              x0 = ex2v(x0);
              x1 = ex2v(x1);
            }
            sr[cc][i] = __float_as_uint(x0);
            sr[cc][i + 1] = __float_as_uint(x1);
          }
        };
        auto pack_store = [&](int cc) {  // row sum, 16-bit P over the S columns
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = __uint_as_float(sr[cc][i]), p1 = __uint_as_float(sr[cc][i + 1]);
            add2(sumv[i & 6], sumv[(i & 6) + 1], p0, p1);  // four independent chains: FADD2 latency is ~8 cycles
            pk[i >> 1] = DT<T>::pack2(p0, p1);
          }
          tmem_st16v(s_addr + cc * 16, pk);
        };
        exps(0);
        exps(1);
        pack_store(0);
        pack_store(1);
        p_arrive(0);
        exps(2);
        exps(3);
        pack_store(2);
        pack_store(3);
        sum0 = (sumv[0] + sumv[2]) + (sumv[4] + sumv[6]);
        sum1 = (sumv[1] + sumv[3]) + (sumv[5] + sumv[7]);
      } else {
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float p0 = ex2(fmaf(__uint_as_float(sr[cc][i]), c, -mc));
            float p1 = ex2(fmaf(__uint_as_float(sr[cc][i + 1]), c, -mc));
            if (cc * 32 + i >= lim) p0 = 0.f;
            if (cc * 32 + i + 1 >= lim) p1 = 0.f;
            sum0 += p0;
            sum1 += p1;
            pk[i >> 1] = DT<T>::pack2(p0, p1);
          }
          tmem_st16(s_addr + cc * 16, pk);
        }
      }
      l += sum0 + sum1;
      FPROF(4);
      if (!warp_full) p_arrive(0);
      p_arrive(1);
      FPROF(5);
    }
    // o / l -> out[token][head][:]
    if (my_nt > 0) {
      mbar_wait_bounded(&o_full[t], 0);
      fence_after_sync();
      const float inv = l > 0.f ? 1.f / l : 0.f;
      T* dst = reinterpret_cast<T*>(args.out) + (int64_t)(q_begin + qi) * args.out_stride + (int64_t)head * D;
      uint32_t ra[16], rb[16];
      tmem_ld16(o_addr, ra);
#pragma unroll
      for (int cc = 0; cc < D / 16; ++cc) {
        uint32_t (&cur)[16] = (cc & 1) ? rb : ra;
        uint32_t (&nxt)[16] = (cc & 1) ? ra : rb;
        tmem_ld_wait_on(cur);
        if (cc + 1 < D / 16) tmem_ld16(o_addr + (cc + 1) * 16, nxt);
        if (qi < q_len) {
#pragma unroll
          for (int i = 0; i < 16; i += 8) {
            uint4 v;
            v.x = DT<T>::pack2(__uint_as_float(cur[i]) * inv, __uint_as_float(cur[i + 1]) * inv);
            v.y = DT<T>::pack2(__uint_as_float(cur[i + 2]) * inv, __uint_as_float(cur[i + 3]) * inv);
            v.z = DT<T>::pack2(__uint_as_float(cur[i + 4]) * inv, __uint_as_float(cur[i + 5]) * inv);
            v.w = DT<T>::pack2(__uint_as_float(cur[i + 6]) * inv, __uint_as_float(cur[i + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + cc * 16 + i) = v;
          }
        }
      }
      FPROF(6);
    } else if (qi < q_len) {  // a query row that sees no key: zeros
      T* dst = reinterpret_cast<T*>(args.out) + (int64_t)(q_begin + qi) * args.out_stride + (int64_t)head * D;
#pragma unroll
      for (int i = 0; i < D; i += 8) *reinterpret_cast<uint4*>(dst + i) = make_uint4(0, 0, 0, 0);
    }
#ifdef B200_FA_PROF
    if ((warp == 2 || warp == 6) && lane == 0) FPROF_FLUSH(warp == 2 ? 8 : 16);
#endif
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
}

template <typename T, int D, int POLY>
int launch_fa_poly(const b200_varlen_attention_params& p, cudaStream_t st) {
  using P = FaPlan<D>;
  CUtensorMap tm_q, tm_k, tm_v;
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_UINT16;
  int rc = make_tmap_2d_ex(&tm_q, p.q, (uint64_t)p.total_q, (uint64_t)p.num_heads * D, (uint64_t)p.q_stride * 2, dt, 64,
                           kBQ, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_k, p.k, (uint64_t)p.total_k, (uint64_t)p.num_kv_heads * D, (uint64_t)p.k_stride * 2, dt, 64,
                       kBK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != B200_OK) return rc;
  rc = make_tmap_2d_ex(&tm_v, p.v, (uint64_t)p.total_k, (uint64_t)p.num_kv_heads * D, (uint64_t)p.v_stride * 2, dt, 64,
                       kBK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != B200_OK) return rc;
  FaArgs a{};
  a.out = p.out;
  a.cu_q = p.cu_seqlens_q;
  a.cu_k = p.cu_seqlens_k;
  a.num_heads = p.num_heads;
  a.num_kv_heads = p.num_kv_heads;
  a.out_stride = p.out_stride;
  a.scale_log2 = p.softmax_scale * 1.4426950408889634f;
  a.causal = p.causal;
  auto kern = varlen_attention_kernel<T, D, POLY>;
  static size_t configured[kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, configured) != cudaSuccess) {
    cudaGetLastError();
    set_error("varlen_attention_fwd: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  const int q_tiles = (p.max_seqlen_q + 2 * kBQ - 1) / (2 * kBQ);  // a CTA owns two query tiles
  cudaError_t le = launch_pdl(kern, dim3(p.num_heads, p.num_seqs, q_tiles), dim3(kFaThreads), (size_t)P::kBytes, st,
                              tm_q, tm_k, tm_v, a);
  if (le != cudaSuccess) {
    cudaGetLastError();
    set_error("varlen_attention_fwd: launch failed: %s", cudaGetErrorString(le));
    return B200_ECUDA;
  }
  return B200_OK;
}

template <typename T, int D>
int launch_fa(const b200_varlen_attention_params& p, cudaStream_t st) {
  static const int poly = [] {
    const char* e = getenv("B200_FA_POLY");
    return e ? atoi(e) : B200_FA_POLY_PAIRS;
  }();
  if constexpr (D == 128) {
    switch (poly) {
      case 4: return launch_fa_poly<T, D, 4>(p, st);
      case 8: return launch_fa_poly<T, D, 8>(p, st);
#ifdef B200_FA_PROF
      case 16: return launch_fa_poly<T, D, 16>(p, st);
      case -1: return launch_fa_poly<T, D, -1>(p, st);  // no exp2 at all (wrong results): the floor of the rest
#endif
      default: break;
    }
  }
  return launch_fa_poly<T, D, 0>(p, st);
}

}  // namespace
}  // namespace b200

extern "C" int b200_debug_fa_prof(unsigned long long* dst, int32_t n, int32_t reset) {
#ifdef B200_FA_PROF
  if (n > 32) n = 32;
  if (cudaMemcpyFromSymbol(dst, b200::g_fa_prof, (size_t)n * 8) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[32] = {};
    cudaMemcpyToSymbol(b200::g_fa_prof, z, sizeof(z));
  }
  return n;
#else
  (void)dst;
  (void)n;
  (void)reset;
  return 0;
#endif
}

extern "C" int b200_varlen_attention_fwd(const b200_varlen_attention_params* pp, void* stream) {
  using namespace b200;
  B200_REQUIRE(pp != nullptr, B200_EINVAL, "varlen_attention_fwd: null params");
  const b200_varlen_attention_params& p = *pp;
  B200_REQUIRE(p.num_seqs >= 0 && p.total_q >= 0 && p.total_k >= 0, B200_EINVAL, "varlen_attention_fwd: negative size");
  if (p.num_seqs == 0 || p.total_q == 0 || p.max_seqlen_q == 0) return B200_OK;
  B200_REQUIRE(p.out && p.q && p.k && p.v && p.cu_seqlens_q && p.cu_seqlens_k, B200_EINVAL,
               "varlen_attention_fwd: null pointer");
  B200_REQUIRE(p.head_size == 128 || p.head_size == 64, B200_ENOTSUP,
               "varlen_attention_fwd: head_size %d (64 and 128 are built)", p.head_size);
  B200_REQUIRE(p.num_kv_heads > 0 && p.num_heads % p.num_kv_heads == 0, B200_EINVAL,
               "varlen_attention_fwd: num_heads must be a multiple of num_kv_heads");
  B200_REQUIRE(((uintptr_t)p.q & 15) == 0 && ((uintptr_t)p.k & 15) == 0 && ((uintptr_t)p.v & 15) == 0 &&
                   ((uintptr_t)p.out & 15) == 0 && p.q_stride % 8 == 0 && p.k_stride % 8 == 0 &&
                   p.v_stride % 8 == 0 && p.out_stride % 8 == 0,
               B200_EINVAL, "varlen_attention_fwd: tensors must be 16-byte aligned with 16-byte aligned token strides");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.dtype == B200_F16) return p.head_size == 128 ? launch_fa<__half, 128>(p, st) : launch_fa<__half, 64>(p, st);
  if (p.dtype == B200_BF16)
    return p.head_size == 128 ? launch_fa<__nv_bfloat16, 128>(p, st) : launch_fa<__nv_bfloat16, 64>(p, st);
  set_error("varlen_attention_fwd: unsupported dtype %d (fp16 / bf16)", p.dtype);
  return B200_EINVAL;
}
