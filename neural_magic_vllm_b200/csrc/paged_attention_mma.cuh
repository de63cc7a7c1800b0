// Tensor-core paged decode attention kernel body (see paged_attention.cuh for the design).
// Included by paged_attention_f16.cu / paged_attention_bf16.cu, which instantiate it per dtype.
//
// Template axes:
//   T      __half / __nv_bfloat16 (query, output, P.V operand type)
//   KV     B200_KV_AUTO (cache holds T), B200_KV_FP8_E4M3, B200_KV_FP8_E5M2 (cache holds bytes)
//   NT     n-tiles of 8 query heads handled per CTA (GQA group <= 8 -> 1, <= 16 -> 2)
//   HMAX   register budget for the head dimension (128 or 256)
//   FAST   head_size == HMAX and block_size == 16: every shared-memory offset is an immediate
#pragma once

#include "paged_attention.cuh"

namespace b200 {

template <typename T>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                          uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&d)[4], const uint32_t (&a)[4],
                                                  uint32_t b0, uint32_t b1) {
  mma_16816_f16(d, a, b0, b1);
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4],
                                                         uint32_t b0, uint32_t b1) {
  mma_16816_bf16(d, a, b0, b1);
}

__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t v) {
  uint32_t o;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(o) : "r"(v));
  return o;
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}

// two packed fp8 (low 16 bits of v) -> two packed T. Exact for both fp8 flavours.
template <typename T, int KV>
__device__ __forceinline__ uint32_t fp8x2_to_T2(uint32_t v) {
  uint32_t h;
  if constexpr (KV == B200_KV_FP8_E5M2) {
    h = __byte_perm(v, 0, 0x1404);  // e5m2 is the top byte of an fp16: (b1 b0) -> (b1,0,b0,0)
  } else {
    unsigned short s = (unsigned short)(v & 0xffffu);
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h) : "h"(s));
  }
  if constexpr (DT<T>::id == B200_BF16) {
    __half2 hh = *reinterpret_cast<__half2*>(&h);
    float2 f = __half22float2(hh);
    __nv_bfloat162 b = __floats2bfloat162_rn(f.x, f.y);  // exact: e4m3/e5m2 fit in bf16
    return *reinterpret_cast<uint32_t*>(&b);
  } else {
    return h;
  }
}

__device__ __forceinline__ bool umma_elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

template <typename T, int KV>
__device__ __forceinline__ void fp8x4_to_frag(uint32_t w, uint32_t& lo, uint32_t& hi) {
  lo = fp8x2_to_T2<T, KV>(w);
  hi = fp8x2_to_T2<T, KV>(w >> 16);
}

// NW: warps per CTA (4; 3 where that lets the whole grid be resident at once, see paged_attention.cu)
template <typename T, int KV, int NT, int HMAX, bool FAST, int NW>
__global__ void __launch_bounds__(NW * 32)
    paged_attn_mma_kernel(const AttnArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr bool FP8 = KV != B200_KV_AUTO;
  constexpr int KS = HMAX / 16;  // k-steps of QK == d-tiles of PV
  constexpr int ESZ = FP8 ? 1 : 2;
  // fp8 K rows are read 16 B at a time (4 k-steps per load) when the head dim is a multiple of 64
  constexpr bool K128 = FP8 && FAST && (HMAX % 64 == 0);
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int r = lane >> 2, t = lane & 3;
  const int G = a.num_heads / a.num_kv_heads;
  const int kvh = blockIdx.x / a.q_tiles;
  const int qtile = blockIdx.x - kvh * a.q_tiles;
  const int seq = blockIdx.y;
  // a cluster of C CTAs (consecutive in z) shares one (kv head, sequence, partition): each takes a
  // contiguous slice of the partition's 16-token items and the slices are merged through
  // distributed shared memory at the end -- more CTAs per sequence without touching the
  // caller-owned partition layout (paged_attn.py:145-156 fixes it at 512 tokens)
  const int C = a.cluster;
  const int part = C > 1 ? blockIdx.z / C : blockIdx.z;
  const int crank = C > 1 ? blockIdx.z - part * C : 0;
  const int BS = FAST ? 16 : a.block_size;
  const int D = FAST ? HMAX : a.head_size;
  const int nks = D >> 4;
  // programmatic dependent launch: this grid may be resident before the kernel that wrote the
  // query / the newest KV slot (rotary + cache write) has finished; every CTA waits for it here
  pdl_wait();
  const int seq_len = a.sl[seq];
  const bool partitioned = a.partition_size > 0;
  const int tok_begin = partitioned ? part * a.partition_size : 0;
  const int tok_end = partitioned ? min(seq_len, tok_begin + a.partition_size) : seq_len;
  if (partitioned && tok_begin >= seq_len) return;  // attention_kernels.cu:128-131

  const int IT = BS < 16 ? 16 : BS;  // tokens per pipeline item (stage)
  int item_begin = tok_begin / IT;
  int item_end = (tok_end + IT - 1) / IT;
  if (C > 1) {
    const int per = (max(item_end - item_begin, 0) + C - 1) / C;
    item_begin += crank * per;
    item_end = min(item_end, item_begin + per);
  }
  const int n_total = max(item_end - item_begin, 0);
  const int n_items = n_total > warp ? (n_total - warp + NW - 1) / NW : 0;
  const int num_seq_blocks = (seq_len + BS - 1) / BS;

  uint64_t* full = reinterpret_cast<uint64_t*>(smem);  // [NW][kAttnMaxStages]
  const int S = a.stages;
  const uint32_t stage_bytes = FAST ? 2u * HMAX * 16 * ESZ : (uint32_t)a.stage_bytes;
  const uint32_t tile_bytes = (uint32_t)D * BS * ESZ;  // one (block, kv head) K or V tile
  uint8_t* wbase = smem + 256 + (size_t)warp * S * stage_bytes;
  const uint32_t wbase_s = smem_u32(wbase);
  uint64_t* wbar = &full[warp * kAttnMaxStages];

  if (tid == 0) {
    for (int i = 0; i < NW * kAttnMaxStages; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  __syncthreads();

  // ---- block-table window: lane i holds the physical block(s) of this warp's item 32*grp+i ----
  const int32_t* bt_row = a.bt + (int64_t)seq * a.max_blocks_per_seq;
  int bt0 = 0, bt1 = 0, bt_grp = -1;
  auto load_bt = [&](int grp) {
    const int item = item_begin + warp + NW * (grp * 32 + lane);
    if (BS >= 16) {
      bt0 = item < num_seq_blocks ? __ldg(bt_row + item) : 0;
    } else {
      bt0 = 2 * item < num_seq_blocks ? __ldg(bt_row + 2 * item) : 0;
      bt1 = 2 * item + 1 < num_seq_blocks ? __ldg(bt_row + 2 * item + 1) : 0;
    }
    bt_grp = grp;
  };
  const uint8_t* kc = reinterpret_cast<const uint8_t*>(a.kc) + (int64_t)kvh * a.kv_head_stride * ESZ;
  const uint8_t* vc = reinterpret_cast<const uint8_t*>(a.vc) + (int64_t)kvh * a.kv_head_stride * ESZ;
  const int64_t blk_bytes = a.kv_block_stride * ESZ;

  // producer state: item index -> ring slot, without divisions
  const bool leader = umma_elect_one();
  int iss_i = 0, iss_s = 0;
  auto issue_next = [&]() {  // whole warp calls; lane 0 issues
    const int i = iss_i;
    if ((i >> 5) != bt_grp) load_bt(i >> 5);
    const int p0 = __shfl_sync(0xffffffffu, bt0, i & 31);
    int p1 = 0;
    if (!FAST && BS < 16) p1 = __shfl_sync(0xffffffffu, bt1, i & 31);
    if (leader) {  // elect.sync: the copy's operands stay in uniform registers (no waterfall loop)
      uint8_t* dst = wbase + (size_t)iss_s * stage_bytes;
      uint64_t* bar = &wbar[iss_s];
      if (FAST || BS >= 16) {
        mbar_arrive_expect_tx(bar, 2 * tile_bytes);
        bulk_g2s(dst, kc + p0 * blk_bytes, tile_bytes, bar);
        bulk_g2s(dst + (stage_bytes >> 1), vc + p0 * blk_bytes, tile_bytes, bar);
      } else {
        const int item = item_begin + warp + NW * i;
        const bool two = 2 * item + 1 < num_seq_blocks;
        mbar_arrive_expect_tx(bar, two ? 4 * tile_bytes : 2 * tile_bytes);
        bulk_g2s(dst, kc + p0 * blk_bytes, tile_bytes, bar);
        bulk_g2s(dst + (stage_bytes >> 1), vc + p0 * blk_bytes, tile_bytes, bar);
        if (two) {
          bulk_g2s(dst + tile_bytes, kc + p1 * blk_bytes, tile_bytes, bar);
          bulk_g2s(dst + (stage_bytes >> 1) + tile_bytes, vc + p1 * blk_bytes, tile_bytes, bar);
        }
      }
    }
    ++iss_i;
    if (++iss_s == S) iss_s = 0;
  };

  // prologue: get the KV stream going before anything else
  for (int i = 0; i < S && i < n_items; ++i) issue_next();
  // programmatic dependent launch: the v2 reduce grid may become resident from here on (it
  // still waits for this grid to finish before reading tmp_out)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  // ---- Q^T fragments (B operand): heads in N, head dim in K ----
  // k-slot -> head-dim maps (the same permutation is applied to the K operand):
  //   16-bit cache : slots {2t,2t+1 | 2t+8,2t+9} of step kk <-> d = 16kk + {2t,2t+1 | 2t+8,2t+9}
  //   fp8, LDS.32  : d = 16kk + {4t,4t+1 | 4t+2,4t+3}
  //   fp8, LDS.128 : step kk = 4g+j <-> d = 16(4g+t) + 4j + {0,1 | 2,3}
  const int head_in_group0 = qtile * (8 * NT);
  uint32_t qf[NT][KS][2];
  float alibi2[NT][2];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int hg = head_in_group0 + 8 * j + r;  // B fragment: n = lane/4
    const bool hv = hg < G;
    const T* qp = reinterpret_cast<const T*>(a.q) + (int64_t)seq * a.q_stride +
                  (int64_t)(kvh * G + hg) * D;
#pragma unroll
    for (int kk = 0; kk < KS; ++kk) {
      uint32_t b0 = 0, b1 = 0;
      if (hv && (FAST || kk < nks)) {
        if constexpr (!FP8) {
          b0 = *reinterpret_cast<const uint32_t*>(qp + 16 * kk + 2 * t);
          b1 = *reinterpret_cast<const uint32_t*>(qp + 16 * kk + 8 + 2 * t);
        } else {
          const int d0 = K128 ? 16 * (4 * (kk >> 2) + t) + 4 * (kk & 3) : 16 * kk + 4 * t;
          const uint2 v = *reinterpret_cast<const uint2*>(qp + d0);
          b0 = v.x;
          b1 = v.y;
        }
      }
      qf[j][kk][0] = b0;
      qf[j][kk][1] = b1;
    }
    // softmax state is indexed by the C fragment: heads 2t, 2t+1
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int hc = head_in_group0 + 8 * j + 2 * t + e;
      alibi2[j][e] = (a.alibi != nullptr && hc < G) ? a.alibi[kvh * G + hc] * kLog2e : 0.f;
    }
  }
  const bool has_alibi = a.alibi != nullptr;
  const float scale2 = a.scale * a.k_scale * kLog2e;

  float m_run[NT][2], l_run[NT][2];
  float acc[KS][NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    m_run[j][0] = m_run[j][1] = -INFINITY;
    l_run[j][0] = l_run[j][1] = 0.f;
  }
#pragma unroll
  for (int dt = 0; dt < KS; ++dt)
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[dt][j][e] = 0.f;

  // token owned by S^T row r (rho0) and row r+8 (rho1) inside a 16-token chunk
  const int rho0 = FP8 ? (4 * (r >> 1) + (r & 1)) : r;
  const int rho1 = FP8 ? rho0 + 2 : r + 8;

  // ---- main loop over this warp's items ----
  int cs = 0;
  uint32_t cphase = 0;
  for (int it = 0; it < n_items; ++it) {
    mbar_wait(&wbar[cs], cphase);
    const uint32_t st = wbase_s + (uint32_t)cs * stage_bytes;
    const int item = item_begin + warp + NW * it;
    const int n_chunks = (FAST || BS <= 16) ? 1 : (BS >> 4);
    for (int c = 0; c < n_chunks; ++c) {
      const int tok0 = item * IT + 16 * c;
      if (tok0 >= tok_end) break;
      const bool tail = tok0 + 16 > tok_end;

      // ---- S^T[16 tok][8*NT heads] = K . Q^T ----
      float sacc[2][NT][4];
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) sacc[h][j][e] = 0.f;

      if constexpr (!FP8) {
        // ldmatrix.x4: lanes 0-7 rows of (tok 0-7, g), 8-15 (tok 8-15, g), 16-23 (tok 0-7, g+1), ...
        const int o = (lane >> 3) & 1;
        uint32_t kaddr;
        uint32_t gstride;  // bytes between consecutive x-groups (x = 8 elements)
        if (FAST || BS >= 16) {
          gstride = (uint32_t)BS * 16;
          kaddr = st + (uint32_t)(16 * c + 8 * o + (lane & 7)) * 16 + (lane >> 4) * gstride;
        } else {
          gstride = 8 * 16;
          kaddr = st + o * tile_bytes + (uint32_t)(lane & 7) * 16 + (lane >> 4) * gstride;
        }
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
          if (FAST || kk < nks) {
            uint32_t af[4];
            ldmatrix_x4(af[0], af[1], af[2], af[3], kaddr + (uint32_t)kk * 2 * gstride);
#pragma unroll
            for (int j = 0; j < NT; ++j) mma_16816<T>(sacc[kk & 1][j], af, qf[j][kk][0], qf[j][kk][1]);
          }
        }
      } else if constexpr (K128) {
        // one 16-byte row (token, x-group 4g+t) feeds four k-steps
        const uint32_t k0 = st + (uint32_t)rho0 * 16 + (uint32_t)t * (16 * 16);
        const uint32_t k1 = st + (uint32_t)rho1 * 16 + (uint32_t)t * (16 * 16);
#pragma unroll
        for (int g = 0; g < KS / 4; ++g) {
          const uint4 w0 = lds128(k0 + g * (4 * 16 * 16));
          const uint4 w1 = lds128(k1 + g * (4 * 16 * 16));
          const uint32_t w0a[4] = {w0.x, w0.y, w0.z, w0.w};
          const uint32_t w1a[4] = {w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const int kk = 4 * g + jj;
            uint32_t af[4];
            fp8x4_to_frag<T, KV>(w0a[jj], af[0], af[2]);
            fp8x4_to_frag<T, KV>(w1a[jj], af[1], af[3]);
#pragma unroll
            for (int j = 0; j < NT; ++j) mma_16816<T>(sacc[kk & 1][j], af, qf[j][kk][0], qf[j][kk][1]);
          }
        }
      } else {
        // fp8: x = 16, one x-group per k-step; row = token, 16 bytes of head dim
        uint32_t k0, k1, gstride;
        if (BS >= 16) {
          gstride = (uint32_t)BS * 16;
          k0 = st + (uint32_t)(16 * c + rho0) * 16 + 4 * t;
          k1 = st + (uint32_t)(16 * c + rho1) * 16 + 4 * t;
        } else {
          gstride = 8 * 16;
          k0 = st + (rho0 >> 3) * tile_bytes + (uint32_t)(rho0 & 7) * 16 + 4 * t;
          k1 = st + (rho1 >> 3) * tile_bytes + (uint32_t)(rho1 & 7) * 16 + 4 * t;
        }
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
          if (FAST || kk < nks) {
            uint32_t af[4];
            fp8x4_to_frag<T, KV>(lds32(k0 + (uint32_t)kk * gstride), af[0], af[2]);
            fp8x4_to_frag<T, KV>(lds32(k1 + (uint32_t)kk * gstride), af[1], af[3]);
#pragma unroll
            for (int j = 0; j < NT; ++j) mma_16816<T>(sacc[kk & 1][j], af, qf[j][kk][0], qf[j][kk][1]);
          }
        }
      }

      // ---- online softmax (per head; heads 2t, 2t+1 of every n-tile live in this thread) ----
      const int tp0 = tok0 + rho0, tp1 = tok0 + rho1;
      uint32_t pb0[NT], pb1[NT];
      float corr[NT][2];
      bool changed = false;
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        float sv[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) sv[e] = (sacc[0][j][e] + sacc[1][j][e]) * scale2;
        if (has_alibi) {
          const float d0 = (float)(tp0 - seq_len + 1), d1 = (float)(tp1 - seq_len + 1);
          sv[0] = fmaf(alibi2[j][0], d0, sv[0]);
          sv[1] = fmaf(alibi2[j][1], d0, sv[1]);
          sv[2] = fmaf(alibi2[j][0], d1, sv[2]);
          sv[3] = fmaf(alibi2[j][1], d1, sv[3]);
        }
        if (tail) {
          if (tp0 >= tok_end) sv[0] = sv[1] = -INFINITY;
          if (tp1 >= tok_end) sv[2] = sv[3] = -INFINITY;
        }
        float mx0 = fmaxf(sv[0], sv[2]), mx1 = fmaxf(sv[1], sv[3]);
#pragma unroll
        for (int off = 4; off <= 16; off <<= 1) {
          mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
          mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
        }
        const float mn0 = fmaxf(m_run[j][0], mx0), mn1 = fmaxf(m_run[j][1], mx1);
        changed |= (mn0 != m_run[j][0]) | (mn1 != m_run[j][1]);
        // every processed chunk has >= 1 valid token, so mn is finite
        corr[j][0] = exp2f(m_run[j][0] - mn0);
        corr[j][1] = exp2f(m_run[j][1] - mn1);
        m_run[j][0] = mn0;
        m_run[j][1] = mn1;
        const float p0 = exp2f(sv[0] - mn0), p1 = exp2f(sv[1] - mn1);
        const float p2 = exp2f(sv[2] - mn0), p3 = exp2f(sv[3] - mn1);
        l_run[j][0] = fmaf(l_run[j][0], corr[j][0], p0 + p2);
        l_run[j][1] = fmaf(l_run[j][1], corr[j][1], p1 + p3);
        // probabilities are rounded to the model dtype before P.V (attention_kernels.cu:398-400)
        pb0[j] = movmatrix_trans(DT<T>::pack2(p0, p1));
        pb1[j] = movmatrix_trans(DT<T>::pack2(p2, p3));
      }
      // rescale the running output only when some head's maximum moved (rare after the first
      // few blocks); the vote keeps the branch warp-uniform
      if (__any_sync(0xffffffffu, changed)) {
#pragma unroll
        for (int dt = 0; dt < KS; ++dt)
#pragma unroll
          for (int j = 0; j < NT; ++j) {
            acc[dt][j][0] *= corr[j][0];
            acc[dt][j][1] *= corr[j][1];
            acc[dt][j][2] *= corr[j][0];
            acc[dt][j][3] *= corr[j][1];
          }
      }

      // ---- O^T[D][8*NT heads] += V^T . P^T ----
      uint32_t vaddr, dstride;  // 16-bit: ldmatrix lane address; fp8: row-r address
      if constexpr (!FP8) {
        const int o = lane >> 4;              // token octet of this lane's matrix
        const int dsub = (lane >> 3) & 1;     // d octet inside the 16-row tile
        if (FAST || BS >= 16) {
          dstride = (uint32_t)BS * 2;
          vaddr = st + (stage_bytes >> 1) + (uint32_t)(16 * c + 8 * o) * 2 +
                  (uint32_t)(8 * dsub + (lane & 7)) * dstride;
        } else {
          dstride = 8 * 2;
          vaddr = st + (stage_bytes >> 1) + o * tile_bytes +
                  (uint32_t)(8 * dsub + (lane & 7)) * dstride;
        }
      } else {
        if (FAST || BS >= 16) {
          dstride = (uint32_t)BS;
          vaddr = st + (stage_bytes >> 1) + (uint32_t)(16 * c) + 4 * t + (uint32_t)r * dstride;
        } else {
          dstride = 8;
          vaddr = st + (stage_bytes >> 1) + (t >> 1) * tile_bytes + 4 * (t & 1) + (uint32_t)r * dstride;
        }
      }
      // tail: zero V of tokens past the end (they may hold NaN; attention_kernels.cu:420-430)
      uint32_t vm0 = 0xffffffffu, vm2 = 0xffffffffu;
      if (tail) {
        const int ta = tok0 + (FP8 ? 4 * t : 2 * t);          // tokens of af[0]/af[1]: ta, ta+1
        const int tb = tok0 + (FP8 ? 4 * t + 2 : 2 * t + 8);  // tokens of af[2]/af[3]: tb, tb+1
        vm0 = (ta < tok_end ? 0x0000ffffu : 0u) | (ta + 1 < tok_end ? 0xffff0000u : 0u);
        vm2 = (tb < tok_end ? 0x0000ffffu : 0u) | (tb + 1 < tok_end ? 0xffff0000u : 0u);
      }
#pragma unroll
      for (int dt = 0; dt < KS; ++dt) {
        if (FAST || dt < nks) {
          uint32_t af[4];
          if constexpr (!FP8) {
            ldmatrix_x4(af[0], af[1], af[2], af[3], vaddr + (uint32_t)dt * 16 * dstride);
          } else {
            fp8x4_to_frag<T, KV>(lds32(vaddr + (uint32_t)(16 * dt) * dstride), af[0], af[2]);
            fp8x4_to_frag<T, KV>(lds32(vaddr + (uint32_t)(16 * dt + 8) * dstride), af[1], af[3]);
          }
          if (tail) {
            af[0] &= vm0;
            af[1] &= vm0;
            af[2] &= vm2;
            af[3] &= vm2;
          }
#pragma unroll
          for (int j = 0; j < NT; ++j) mma_16816<T>(acc[dt][j], af, pb0[j], pb1[j]);
        }
      }
    }
    __syncwarp();
    if (iss_i < n_items) issue_next();
    if (++cs == S) {
      cs = 0;
      cphase ^= 1;
    }
  }

  // ---- merge the 4 warps' partial softmax states through shared memory ----
  // per-thread l is partial over the 8 lanes that share t: finish it
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int off = 4; off <= 16; off <<= 1)
        l_run[j][e] += __shfl_xor_sync(0xffffffffu, l_run[j][e], off);

  constexpr int NH = 8 * NT;
  const int DP = D + 4;  // padded row: conflict-free fragment stores
  // this warp's own ring is drained; reuse it: [NH][DP] fp32 acc, then m[NH], l[NH]
  float* wacc = reinterpret_cast<float*>(wbase);
  float* wm = wacc + NH * DP;
  float* wl = wm + NH;
#pragma unroll
  for (int dt = 0; dt < KS; ++dt) {
    if (FAST || dt < nks) {
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int h = 8 * j + 2 * t;
        const int d = 16 * dt + r;
        wacc[h * DP + d] = acc[dt][j][0];
        wacc[(h + 1) * DP + d] = acc[dt][j][1];
        wacc[h * DP + d + 8] = acc[dt][j][2];
        wacc[(h + 1) * DP + d + 8] = acc[dt][j][3];
      }
    }
  }
  if (r == 0) {
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        wm[8 * j + 2 * t + e] = m_run[j][e];
        wl[8 * j + 2 * t + e] = l_run[j][e];
      }
  }
  __syncthreads();

  const size_t wstride = (size_t)S * stage_bytes / sizeof(float);  // floats between warps' areas
  const float* base = reinterpret_cast<const float*>(smem + 256);
  const int half_d = D >> 1;
  // cluster exchange area behind the rings: [NH][D] unnormalised output, then m[NH], l[NH]
  float* xo = reinterpret_cast<float*>(smem + 256 + (size_t)NW * S * stage_bytes);
  float* xm = xo + NH * D;
  float* xl = xm + NH;
  auto finish = [&](int h, int d2, float M, float L, float o0, float o1) {
    const int hg = head_in_group0 + h;
    // attention_kernels.cu:342: inv_sum = 1 / (exp_sum + 1e-6)
    const float inv = a.v_scale * __fdividef(1.f, L + 1e-6f);
    const int head = kvh * G + hg;
    const uint32_t packed = DT<T>::pack2(o0 * inv, o1 * inv);
    if (partitioned && a.max_num_partitions > 1) {
      T* dst = reinterpret_cast<T*>(a.tmp_out) +
               (((int64_t)seq * a.num_heads + head) * a.max_num_partitions + part) * D;
      *reinterpret_cast<uint32_t*>(dst + 2 * d2) = packed;
      if (d2 == 0) {
        const int64_t si = ((int64_t)seq * a.num_heads + head) * a.max_num_partitions + part;
        a.exp_sums[si] = L;
        a.max_logits[si] = M * kLn2;  // back to natural-log units
      }
    } else {
      T* dst = reinterpret_cast<T*>(a.out) + ((int64_t)seq * a.num_heads + head) * D;
      *reinterpret_cast<uint32_t*>(dst + 2 * d2) = packed;
    }
  };
  for (int idx = tid; idx < NH * half_d; idx += (NW * 32)) {
    const int h = idx / half_d;
    const int d2 = idx - h * half_d;
    if (head_in_group0 + h >= G) continue;
    float mw[NW], M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      mw[w] = base[w * wstride + NH * DP + h];
      M = fmaxf(M, mw[w]);
    }
    float L = 0.f, o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float f = mw[w] == -INFINITY ? 0.f : exp2f(mw[w] - M);
      L = fmaf(base[w * wstride + NH * DP + NH + h], f, L);
      const float2 v = *reinterpret_cast<const float2*>(&base[w * wstride + h * DP + 2 * d2]);
      o0 = fmaf(v.x, f, o0);
      o1 = fmaf(v.y, f, o1);
    }
    if (C == 1) {
      finish(h, d2, M, L, o0, o1);
    } else {
      *reinterpret_cast<float2*>(&xo[h * D + 2 * d2]) = make_float2(o0, o1);
      if (d2 == 0) xm[h] = M, xl[h] = L;
    }
  }
  if (C > 1) {
    // every CTA of the cluster has published its slice; rank 0 merges them in rank order
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (crank == 0) {
      const uint32_t xo_s = smem_u32(xo), xm_s = smem_u32(xm), xl_s = smem_u32(xl);
      auto remote = [](uint32_t local, int rank) {
        uint32_t r;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
        return r;
      };
      auto ld_f = [](uint32_t addr) {
        float v;
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr));
        return v;
      };
      for (int idx = tid; idx < NH * half_d; idx += (NW * 32)) {
        const int h = idx / half_d;
        const int d2 = idx - h * half_d;
        if (head_in_group0 + h >= G) continue;
        float M = -INFINITY;
        for (int c = 0; c < C; ++c) M = fmaxf(M, ld_f(remote(xm_s + h * 4, c)));
        float L = 0.f, o0 = 0.f, o1 = 0.f;
        for (int c = 0; c < C; ++c) {
          const float mc = ld_f(remote(xm_s + h * 4, c));
          const float f = mc == -INFINITY ? 0.f : exp2f(mc - M);
          L = fmaf(ld_f(remote(xl_s + h * 4, c)), f, L);
          const uint32_t oa = remote(xo_s + (h * D + 2 * d2) * 4, c);
          o0 = fmaf(ld_f(oa), f, o0);
          o1 = fmaf(ld_f(oa + 4), f, o1);
        }
        finish(h, d2, M, L, o0, o1);
      }
    }
    // nobody leaves (and frees its shared memory) before rank 0 has read it
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
}

template <typename T, int KV, int NT, int HMAX, bool FAST, int NW = kAttnWarps>
static int launch_one(const AttnArgs& a, dim3 grid, size_t smem, cudaStream_t stream) {
  auto kern = paged_attn_mma_kernel<T, KV, NT, HMAX, FAST, NW>;
  static size_t configured[kMaxDevices] = {};  // per instantiation and device; grows monotonically
  cudaError_t e = ensure_dyn_smem(kern, smem, configured);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("paged_attention: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
    return B200_ECUDA;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(NW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (a.cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 1;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = (unsigned)a.cluster;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaLaunchKernelEx(&cfg, kern, a);
  return check_launch("paged_attention");
}

template <typename T, int KV>
static int launch_kv(const AttnArgs& a, dim3 grid, size_t smem, cudaStream_t stream) {
  const int G = a.num_heads / a.num_kv_heads;
  const bool fast = a.head_size == 128 && a.block_size == 16;
  if (G > 8) {
    if (fast) return launch_one<T, KV, 2, 128, true>(a, grid, smem, stream);
    if (a.head_size <= 128) return launch_one<T, KV, 2, 128, false>(a, grid, smem, stream);
    return launch_one<T, KV, 2, 256, false>(a, grid, smem, stream);
  }
  if (fast && a.warps == 3) return launch_one<T, KV, 1, 128, true, 3>(a, grid, smem, stream);
  if (fast) return launch_one<T, KV, 1, 128, true>(a, grid, smem, stream);
  if (a.head_size <= 128) return launch_one<T, KV, 1, 128, false>(a, grid, smem, stream);
  return launch_one<T, KV, 1, 256, false>(a, grid, smem, stream);
}

template <typename T>
int launch_paged_attn_mma(const AttnArgs& a, int kv_dtype, dim3 grid, size_t smem,
                          cudaStream_t stream) {
  switch (kv_dtype) {
    case B200_KV_AUTO: return launch_kv<T, B200_KV_AUTO>(a, grid, smem, stream);
    case B200_KV_FP8_E4M3: return launch_kv<T, B200_KV_FP8_E4M3>(a, grid, smem, stream);
    default: return launch_kv<T, B200_KV_FP8_E5M2>(a, grid, smem, stream);
  }
}

}  // namespace b200
