// Host entry points of paged attention + the two kernels that are not dtype-specialised tensor
// kernels: the split-KV reduce (replaces paged_attention_v2_reduce_kernel,
// csrc/attention/attention_kernels.cu:567-669) and a SIMT kernel that covers what the tensor path
// does not (fp32 activations, block-sparse masks, unaligned queries).
#include <cfloat>
#include <cstdlib>

#include "paged_attention.cuh"

namespace b200 {

// -----------------------------------------------------------------------------------------------
// v2 reduce: out = sum_p tmp_out[p] * exp_sums[p] * exp(max[p] - gmax) / (sum_p(...) + 1e-6)
// One warp per (sequence, head) row; lanes own element pairs; the partition loop is unrolled so
// all of a row's loads are in flight together.  Launched with programmatic dependent launch: the
// grid is resident (indices computed, seq_len loaded) before the split-KV kernel drains and only
// waits at griddepcontrol.wait for its results.
// -----------------------------------------------------------------------------------------------
constexpr int kReduceWarps = 4;

template <typename T>
__global__ void __launch_bounds__(kReduceWarps * 32)
    paged_attn_reduce_kernel(T* __restrict__ out, const float* __restrict__ exp_sums,
                             const float* __restrict__ max_logits, const T* __restrict__ tmp_out,
                             const int32_t* __restrict__ seq_lens, const int num_rows,
                             const int num_heads, const int head_size, const int max_num_partitions,
                             const int partition_size) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kReduceWarps + (threadIdx.x >> 5);
  if (row >= num_rows) return;
  const int seq = row / num_heads;
  const int seq_len = seq_lens[seq];
  const int np = min((seq_len + partition_size - 1) / partition_size, max_num_partitions);
  const float* ml = max_logits + (int64_t)row * max_num_partitions;
  const float* es = exp_sums + (int64_t)row * max_num_partitions;
  const uint32_t* tp =
      reinterpret_cast<const uint32_t*>(tmp_out + (int64_t)row * max_num_partitions * head_size);
  uint32_t* op = reinterpret_cast<uint32_t*>(out + (int64_t)row * head_size);
  const int half_d = head_size >> 1;  // element pairs per partition row
  constexpr int MAXP = 4;             // pairs per lane: head_size <= 256

  asm volatile("griddepcontrol.wait;" ::: "memory");  // results of the split-KV kernel are visible
  pdl_launch_dependents();  // the kernel behind the reduce (o_proj GEMM) may take its slots now

  float gmax = -FLT_MAX;
  for (int p = lane; p < np; p += 32) gmax = fmaxf(gmax, ml[p]);
  gmax = warp_max(gmax);
  float acc[MAXP][2];
#pragma unroll
  for (int i = 0; i < MAXP; ++i) acc[i][0] = acc[i][1] = 0.f;
  float gsum = 0.f;
  for (int p0 = 0; p0 < np; p0 += 32) {
    const int pl = p0 + lane;
    const float w = pl < np ? es[pl] * __expf(ml[pl] - gmax) : 0.f;
    gsum += w;
    const int cnt = min(32, np - p0);
    for (int q0 = 0; q0 < cnt; q0 += 8) {
      uint32_t v[8][MAXP];
#pragma unroll
      for (int q = 0; q < 8; ++q)
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
          const int d2 = lane + 32 * i;
          v[q][i] = (q0 + q < cnt && d2 < half_d) ? tp[(int64_t)(p0 + q0 + q) * half_d + d2] : 0u;
        }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float wq = __shfl_sync(0xffffffffu, w, (q0 + q) & 31);
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
          if constexpr (sizeof(T) == 2) {
            const T* e = reinterpret_cast<const T*>(&v[q][i]);
            acc[i][0] = fmaf(DT<T>::to_f(e[0]), wq, acc[i][0]);
            acc[i][1] = fmaf(DT<T>::to_f(e[1]), wq, acc[i][1]);
          }
        }
      }
    }
  }
  gsum = warp_sum(gsum);
  const float inv = np > 0 ? __fdividef(1.f, gsum + 1e-6f) : 0.f;
#pragma unroll
  for (int i = 0; i < MAXP; ++i) {
    const int d2 = lane + 32 * i;
    if constexpr (sizeof(T) == 2) {
      if (d2 < half_d) op[d2] = DT<T>::pack2(acc[i][0] * inv, acc[i][1] * inv);
    }
  }
}

// fp32 rows (SIMT path only): one thread per element
__global__ void paged_attn_reduce_f32_kernel(float* __restrict__ out,
                                             const float* __restrict__ exp_sums,
                                             const float* __restrict__ max_logits,
                                             const float* __restrict__ tmp_out,
                                             const int32_t* __restrict__ seq_lens,
                                             const int head_size, const int max_num_partitions,
                                             const int partition_size) {
  const int head = blockIdx.x, seq = blockIdx.y, num_heads = gridDim.x;
  const int seq_len = seq_lens[seq];
  const int np = min((seq_len + partition_size - 1) / partition_size, max_num_partitions);
  const int64_t row = (int64_t)seq * num_heads + head;
  const float* ml = max_logits + row * max_num_partitions;
  const float* es = exp_sums + row * max_num_partitions;
  const float* tp = tmp_out + row * max_num_partitions * head_size;
  float gmax = -FLT_MAX;
  for (int p = 0; p < np; ++p) gmax = fmaxf(gmax, ml[p]);
  float gsum = 0.f;
  for (int p = 0; p < np; ++p) gsum += es[p] * __expf(ml[p] - gmax);
  const float inv = np > 0 ? __fdividef(1.f, gsum + 1e-6f) : 0.f;
  for (int d = threadIdx.x; d < head_size; d += blockDim.x) {
    float acc = 0.f;
    for (int p = 0; p < np; ++p)
      acc = fmaf(tp[(int64_t)p * head_size + d], es[p] * __expf(ml[p] - gmax), acc);
    out[row * head_size + d] = acc * inv;
  }
}

// -----------------------------------------------------------------------------------------------
// SIMT kernel: one CTA per (query head, sequence, partition). Correctness-first: handles every
// dtype / head size / block size / alibi / block-sparse combination of the reference.
// -----------------------------------------------------------------------------------------------
template <typename T, bool FP8>
__device__ __forceinline__ float load_cache(const void* p, int64_t idx, int e5m2) {
  if constexpr (FP8) {
    const uint8_t v = reinterpret_cast<const uint8_t*>(p)[idx];
    return e5m2 ? fp8_to_float<B200_KV_FP8_E5M2>(v) : fp8_to_float<B200_KV_FP8_E4M3>(v);
  } else {
    return DT<T>::to_f(reinterpret_cast<const T*>(p)[idx]);
  }
}

template <typename T, bool FP8>
__global__ void __launch_bounds__(128) paged_attn_simt_kernel(const AttnArgs a) {
  extern __shared__ float sm[];  // [head_size] q | [tokens] logits
  __shared__ float red[4];
  const int head = blockIdx.x, seq = blockIdx.y, part = blockIdx.z;
  const int tid = threadIdx.x;
  const int D = a.head_size, BS = a.block_size;
  const int G = a.num_heads / a.num_kv_heads;
  const int kvh = head / G;
  const int seq_len = a.sl[seq];
  const bool partitioned = a.partition_size > 0;
  const int tok_begin = partitioned ? part * a.partition_size : 0;
  const int tok_end = partitioned ? min(seq_len, tok_begin + a.partition_size) : seq_len;
  if (partitioned && tok_begin >= seq_len) return;
  const int x = FP8 ? 16 : 16 / (int)sizeof(T);
  float* qs = sm;
  float* logits = sm + D;
  const T* qp = reinterpret_cast<const T*>(a.q) + (int64_t)seq * a.q_stride + (int64_t)head * D;
  for (int d = tid; d < D; d += blockDim.x) qs[d] = DT<T>::to_f(qp[d]);
  __syncthreads();

  const int32_t* bt_row = a.bt + (int64_t)seq * a.max_blocks_per_seq;
  const bool sparse = a.bs_vert_stride > 1;
  int bs_offset = 0, q_bs_block = 0;
  if (sparse) {  // attention_kernels.cu:209-224
    q_bs_block = (seq_len - 1) / a.bs_block_size;
    bs_offset = a.bs_head_sliding_step >= 0
                    ? (a.tp_rank * a.num_heads + head) * a.bs_head_sliding_step + 1
                    : (a.tp_rank * a.num_kv_heads + kvh) * (-a.bs_head_sliding_step) + 1;
  }
  auto attended = [&](int block_idx) {
    if (!sparse) return true;
    const int kb = block_idx * BS / a.bs_block_size;
    return ((kb + bs_offset) % a.bs_vert_stride == 0) || (kb > q_bs_block - a.bs_local_blocks);
  };
  const float alibi = a.alibi ? a.alibi[head] : 0.f;
  const float sc = a.scale * a.k_scale;

  float mx = -FLT_MAX;
  for (int tok = tok_begin + tid; tok < tok_end; tok += blockDim.x) {
    const int bi = tok / BS, off = tok - bi * BS;
    float lg = -FLT_MAX;
    if (attended(bi)) {
      const int64_t kb = (int64_t)bt_row[bi] * a.kv_block_stride + (int64_t)kvh * a.kv_head_stride;
      float dot = 0.f;
      for (int d = 0; d < D; ++d) {
        const int64_t idx = kb + ((int64_t)(d / x) * BS + off) * x + (d % x);
        dot = fmaf(qs[d], load_cache<T, FP8>(a.kc, idx, a.e5m2), dot);
      }
      lg = dot * sc + (a.alibi ? alibi * (float)(tok - seq_len + 1) : 0.f);
      mx = fmaxf(mx, lg);
    }
    logits[tok - tok_begin] = lg;
  }
  mx = warp_max(mx);
  if ((tid & 31) == 0) red[tid >> 5] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float sum = 0.f;
  for (int i = tid; i < tok_end - tok_begin; i += blockDim.x) {
    const float p = __expf(logits[i] - mx);
    logits[i] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  if ((tid & 31) == 0) red[tid >> 5] = sum;
  __syncthreads();
  sum = red[0] + red[1] + red[2] + red[3];
  const float inv = __fdividef(1.f, sum + 1e-6f);

  T* dst;
  if (partitioned && a.max_num_partitions > 1) {
    const int64_t si = ((int64_t)seq * a.num_heads + head) * a.max_num_partitions + part;
    dst = reinterpret_cast<T*>(a.tmp_out) + si * D;
    if (tid == 0) {
      a.exp_sums[si] = sum;
      a.max_logits[si] = mx;
    }
  } else {
    dst = reinterpret_cast<T*>(a.out) + ((int64_t)seq * a.num_heads + head) * D;
  }
  for (int d = tid; d < D; d += blockDim.x) {
    float acc = 0.f;
    for (int tok = tok_begin; tok < tok_end; ++tok) {
      const int bi = tok / BS, off = tok - bi * BS;
      if (!attended(bi)) continue;
      const int64_t vb = (int64_t)bt_row[bi] * a.kv_block_stride + (int64_t)kvh * a.kv_head_stride;
      // probabilities are rounded to the model dtype first (attention_kernels.cu:398-400)
      const float p = DT<T>::to_f(DT<T>::from_f(logits[tok - tok_begin] * inv));
      acc = fmaf(p, load_cache<T, FP8>(a.vc, vb + (int64_t)d * BS + off, a.e5m2), acc);
    }
    dst[d] = DT<T>::from_f(acc * a.v_scale);
  }
}

template <typename T>
static int launch_simt(const AttnArgs& a, bool fp8, dim3 grid, size_t smem, cudaStream_t stream) {
  auto set = [&](auto kern) -> int {
    if (smem > 48 * 1024) {
      cudaError_t e =
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("paged_attention: max_seq_len needs %zu B of shared memory: %s", smem,
                  cudaGetErrorString(e));
        return B200_ENOTSUP;
      }
    }
    kern<<<grid, 128, smem, stream>>>(a);
    return check_launch("paged_attention (simt)");
  };
  return fp8 ? set(paged_attn_simt_kernel<T, true>) : set(paged_attn_simt_kernel<T, false>);
}

static int run_paged_attention(const b200_paged_attention_params* p, bool v2, cudaStream_t stream) {
  B200_REQUIRE(p != nullptr, B200_EINVAL, "paged_attention: null params");
  if (p->num_seqs == 0) return B200_OK;
  B200_REQUIRE(p->num_seqs > 0 && p->num_heads > 0 && p->num_kv_heads > 0 &&
                   p->num_heads % p->num_kv_heads == 0,
               B200_EINVAL, "paged_attention: bad head counts (%d q heads, %d kv heads)",
               p->num_heads, p->num_kv_heads);
  // the reference's supported sets (attention_kernels.cu:763-766, 800-802 / 899-927, 950-964)
  const int hs = p->head_size;
  B200_REQUIRE(hs == 64 || hs == 80 || hs == 96 || hs == 112 || hs == 128 || hs == 192 || hs == 256,
               B200_ENOTSUP, "Unsupported head size: %d", hs);
  B200_REQUIRE(p->block_size == 8 || p->block_size == 16 || p->block_size == 32, B200_ENOTSUP,
               "Unsupported block size: %d", p->block_size);
  B200_REQUIRE(p->kv_dtype >= B200_KV_AUTO && p->kv_dtype <= B200_KV_FP8_E5M2, B200_EINVAL,
               "Unsupported data type of kv cache: %d", p->kv_dtype);
  B200_REQUIRE(p->dtype >= B200_F16 && p->dtype <= B200_F32, B200_EINVAL,
               "Unsupported data type: %d", p->dtype);
  const bool fp8 = p->kv_dtype != B200_KV_AUTO;
  if (v2) {
    B200_REQUIRE(p->exp_sums && p->max_logits && p->tmp_out, B200_EINVAL,
                 "paged_attention_v2: exp_sums / max_logits / tmp_out are required");
  }

  AttnArgs a{};
  a.out = p->out;
  a.q = p->query;
  a.kc = p->key_cache;
  a.vc = p->value_cache;
  a.bt = p->block_tables;
  a.sl = p->seq_lens;
  a.alibi = p->alibi_slopes;
  a.exp_sums = p->exp_sums;
  a.max_logits = p->max_logits;
  a.tmp_out = p->tmp_out;
  a.num_seqs = p->num_seqs;
  a.num_heads = p->num_heads;
  a.num_kv_heads = p->num_kv_heads;
  a.head_size = hs;
  a.block_size = p->block_size;
  a.max_blocks_per_seq = p->max_num_blocks_per_seq;
  a.partition_size = v2 ? B200_PARTITION_SIZE : 0;
  a.max_num_partitions =
      v2 ? (p->max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE : 1;
  if (a.max_num_partitions < 1) a.max_num_partitions = 1;
  a.e5m2 = p->kv_dtype == B200_KV_FP8_E5M2;
  a.q_stride = p->q_stride;
  a.kv_block_stride = p->kv_block_stride;
  a.kv_head_stride = p->kv_head_stride;
  a.scale = p->scale;
  a.k_scale = fp8 ? p->kv_scale : 1.f;
  a.v_scale = fp8 ? p->kv_scale : 1.f;
  a.tp_rank = p->tp_rank;
  a.bs_local_blocks = p->blocksparse_local_blocks;
  a.bs_vert_stride = p->blocksparse_vert_stride;
  a.bs_block_size = p->blocksparse_block_size;
  a.bs_head_sliding_step = p->blocksparse_head_sliding_step;

  const bool sparse = p->blocksparse_vert_stride > 1;
  const int esz_q = p->dtype == B200_F32 ? 4 : 2;
  const int esz_c = fp8 ? 1 : esz_q;
  const bool aligned = (reinterpret_cast<uintptr_t>(p->query) % 8 == 0) &&
                       ((p->q_stride * esz_q) % 8 == 0) &&
                       (reinterpret_cast<uintptr_t>(p->key_cache) % 16 == 0) &&
                       (reinterpret_cast<uintptr_t>(p->value_cache) % 16 == 0) &&
                       ((p->kv_block_stride * esz_c) % 16 == 0) &&
                       ((p->kv_head_stride * esz_c) % 16 == 0) &&
                       (reinterpret_cast<uintptr_t>(p->out) % 4 == 0) &&
                       (!v2 || reinterpret_cast<uintptr_t>(p->tmp_out) % 4 == 0);
  const bool tensor_path = p->dtype != B200_F32 && !sparse && aligned;

  int rc;
  if (tensor_path) {
    const int G = p->num_heads / p->num_kv_heads;
    const int nt = G > 8 ? 2 : 1;
    a.q_tiles = (G + 8 * nt - 1) / (8 * nt);
    const int it_tokens = p->block_size < 16 ? 16 : p->block_size;
    a.stage_bytes = 2 * hs * it_tokens * esz_c;
    const int merge_bytes = 8 * nt * (hs + 4) * 4 + 8 * nt * 8;
    // <= 64 KiB of ring per CTA so that 3-4 CTAs (12-16 warps) share an SM: measured faster than
    // deeper rings with fewer resident warps (profiles/r01_attention.md)
    int stages = (12 * 1024) / a.stage_bytes;
    if (stages < 2) stages = 2;
    if (stages > kAttnMaxStages) stages = kAttnMaxStages;
    {  // tuning knob (bench sweeps), read once per process
      static int knob[2] = {0, 0};
      const int v = env_int_once("B200_ATTN_STAGES", knob, 0);
      if (v >= 2 && v <= kAttnMaxStages) stages = v;
    }
    while (stages * a.stage_bytes < merge_bytes) ++stages;
    // 227 KiB of shared memory per CTA: very large tiles (head 256 x block 32) run single-stage,
    // the 4 warps still overlap each other's copies
    constexpr size_t kMaxSmem = 227 * 1024;
    while (stages > 1 && 256 + (size_t)kAttnWarps * stages * a.stage_bytes > kMaxSmem) --stages;
    B200_REQUIRE(stages <= kAttnMaxStages && stages * a.stage_bytes >= merge_bytes, B200_ENOTSUP,
                 "paged_attention: internal stage sizing");
    a.stages = stages;
    // Fewer CTAs than SMs (batch 1-4 decode): split every partition over a cluster of 2 / 4 / 8 CTAs,
    // merged through distributed shared memory, until the grid covers the SMs or a CTA would get
    // fewer than 128 tokens.  (Measured, profiles/r2_decode_tokens.md: batch 1 2.43 -> 2.34 ms per
    // step; with >= 1 CTA per SM already the split only adds cluster-launch and merge cost --
    // batch 16 2.79 -> 3.33 ms when forced -- so it stops there.)
    const int base_ctas = p->num_kv_heads * a.q_tiles * p->num_seqs * a.max_num_partitions;
    const int part_tokens = v2 ? (p->max_seq_len < B200_PARTITION_SIZE ? p->max_seq_len : B200_PARTITION_SIZE)
                               : p->max_seq_len;
    int cluster = 1;
    while (cluster < 8 && base_ctas * cluster < num_sms() && part_tokens / (cluster * 2) >= 128) cluster *= 2;
    {  // tuning knob (bench sweeps), read once per process: 1 / 2 / 4 / 8 forces the cluster size
      static int knob[2] = {0, 0};
      const int v = env_int_once("B200_ATTN_CLUSTER", knob, 0);
      if (v == 1 || v == 2 || v == 4 || v == 8) cluster = v;
    }
    a.cluster = cluster;
    // A CTA is latency-bound (one 16-token item per warp per memory round trip: ~25 us for 1024
    // tokens whatever the load), so a grid of slightly more than one wave of 4-warp CTAs (3 per SM)
    // takes two full CTA times -- batch 64 x 8 kv heads = 512 CTAs on 444 slots: 52 us, where batch
    // 96 takes 63.  With 3 warps a CTA needs 48 KiB and four fit an SM: when that makes the whole
    // grid resident at once, it runs as one HBM-bound wave instead.
    int warps = kAttnWarps;
    {
      static int knob_w[2] = {0, 0};
      const bool fast = hs == 128 && p->block_size == 16 && G <= 8;
      const int64_t ctas = (int64_t)base_ctas * cluster;
      const int64_t slots4 = 3 * num_sms(), slots3 = 4 * num_sms();  // resident CTAs with 4 / 3 warps
      const int64_t waves4 = (ctas + slots4 - 1) / slots4, waves3 = (ctas + slots3 - 1) / slots3;
      const int cap = env_int_once("B200_ATTN_WARPS3_MAX_WAVES", knob_w, 2);
      if (fast && cluster == 1 && waves3 < waves4 && waves3 <= cap) warps = 3;
      static int knob[2] = {0, 0};  // tuning knob (bench sweeps): 3 / 4 forces the CTA shape
      const int v = env_int_once("B200_ATTN_WARPS", knob, 0);
      if (fast && (v == 3 || v == 4)) warps = v;
    }
    a.warps = warps;
    const int nh = 8 * nt;
    const size_t smem = 256 + (size_t)warps * stages * a.stage_bytes +
                        (cluster > 1 ? (size_t)nh * (hs + 2) * sizeof(float) : 0);
    dim3 grid(p->num_kv_heads * a.q_tiles, p->num_seqs, a.max_num_partitions * cluster);
    rc = p->dtype == B200_F16 ? launch_paged_attn_mma<__half>(a, p->kv_dtype, grid, smem, stream)
                              : launch_paged_attn_mma<__nv_bfloat16>(a, p->kv_dtype, grid, smem, stream);
  } else {
    const int toks = v2 ? B200_PARTITION_SIZE : ((p->max_seq_len + 15) / 16) * 16;
    const size_t smem = (size_t)(hs + toks) * sizeof(float);
    dim3 grid(p->num_heads, p->num_seqs, a.max_num_partitions);
    rc = p->dtype == B200_F16    ? launch_simt<__half>(a, fp8, grid, smem, stream)
         : p->dtype == B200_BF16 ? launch_simt<__nv_bfloat16>(a, fp8, grid, smem, stream)
                                 : launch_simt<float>(a, fp8, grid, smem, stream);
  }
  if (rc != B200_OK) return rc;

  if (v2 && a.max_num_partitions > 1) {
    if (p->dtype == B200_F32) {
      dim3 grid(p->num_heads, p->num_seqs);
      paged_attn_reduce_f32_kernel<<<grid, 128, 0, stream>>>(
          (float*)p->out, p->exp_sums, p->max_logits, (const float*)p->tmp_out, p->seq_lens, hs,
          a.max_num_partitions, B200_PARTITION_SIZE);
      return check_launch("paged_attention_v2 reduce");
    }
    const int rows = p->num_seqs * p->num_heads;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((rows + kReduceWarps - 1) / kReduceWarps);
    cfg.blockDim = dim3(kReduceWarps * 32);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaError_t le;
    if (p->dtype == B200_F16) {
      le = cudaLaunchKernelEx(&cfg, paged_attn_reduce_kernel<__half>, (__half*)p->out,
                              (const float*)p->exp_sums, (const float*)p->max_logits,
                              (const __half*)p->tmp_out, p->seq_lens, rows, (int)p->num_heads, hs,
                              a.max_num_partitions, (int)B200_PARTITION_SIZE);
    } else {
      le = cudaLaunchKernelEx(&cfg, paged_attn_reduce_kernel<__nv_bfloat16>, (__nv_bfloat16*)p->out,
                              (const float*)p->exp_sums, (const float*)p->max_logits,
                              (const __nv_bfloat16*)p->tmp_out, p->seq_lens, rows,
                              (int)p->num_heads, hs, a.max_num_partitions, (int)B200_PARTITION_SIZE);
    }
    if (le != cudaSuccess) {
      cudaGetLastError();
      set_error("paged_attention_v2 reduce: launch failed: %s", cudaGetErrorString(le));
      return B200_ECUDA;
    }
    return check_launch("paged_attention_v2 reduce");
  }
  return B200_OK;
}

}  // namespace b200

extern "C" {

int b200_paged_attention_v1(const b200_paged_attention_params* p, void* stream) {
  return b200::run_paged_attention(p, false, (cudaStream_t)stream);
}

int b200_paged_attention_v2(const b200_paged_attention_params* p, void* stream) {
  return b200::run_paged_attention(p, true, (cudaStream_t)stream);
}

}  // extern "C"
