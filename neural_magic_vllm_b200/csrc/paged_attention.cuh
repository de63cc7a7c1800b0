// Paged decode attention for B200 -- shared declarations of the kernel family.
// Replaces csrc/attention/attention_kernels.cu of the reference (paged_attention_kernel :90-496,
// paged_attention_v2_reduce_kernel :567-669).
//
// Design: one CTA per (kv head, sequence, 512-token partition).  All query
// heads of the GQA group ride in the N dimension of the tensor-core tiles, so every 16-token KV
// block is fetched from HBM exactly once (the reference launches one CTA per *query* head and
// re-reads each block num_heads/num_kv_heads times through L2).  Each of the 4 warps owns a private
// ring of shared-memory stages; lane 0 gathers its blocks through the block table with 1-D bulk
// async copies (TMA engine, UBLKCP) completing on per-stage mbarriers, so there is no CTA-wide
// barrier in the main loop.  Math is S^T = K.Q^T and O^T = V^T.P^T on mma.sync m16n8k16
// (fp32 accumulate): the paged layouts [D/x][tok][x] and [D][tok] are exactly row-major A operands
// with the token / head-dim index on the M rows, which makes ldmatrix conflict-free for K and lets
// the 1..16 query heads sit in the (cheap) N=8 direction instead of padding M to 16.
#pragma once

#include "common.cuh"

namespace b200 {

constexpr int kAttnWarps = 4;
constexpr int kAttnThreads = kAttnWarps * 32;
constexpr int kAttnMaxStages = 6;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct AttnArgs {
  void* out;
  const void* q;
  const void* kc;
  const void* vc;
  const int32_t* bt;
  const int32_t* sl;
  const float* alibi;
  float* exp_sums;
  float* max_logits;
  void* tmp_out;
  int num_seqs, num_heads, num_kv_heads, head_size, block_size;
  int max_blocks_per_seq, max_num_partitions;
  int partition_size;  // 0: one partition spans the sequence (v1)
  int q_tiles;         // CTAs per kv head (group size split into tiles of 8*NT heads)
  int stages, stage_bytes;
  int cluster;         // CTAs (thread-block cluster along z) that split one partition's tokens; 1 = none
  int warps;           // warps per CTA of the tensor-core kernel: kAttnWarps, or 3 (one-wave grids)
  int e5m2;  // fp8 flavour of the cache (only read when FP8)
  int64_t q_stride, kv_block_stride, kv_head_stride;
  float scale;        // softmax scale
  float k_scale;      // kv_scale folded into the logits
  float v_scale;      // kv_scale folded into the output
  // block-sparse (generic kernel only)
  int tp_rank, bs_local_blocks, bs_vert_stride, bs_block_size, bs_head_sliding_step;
};

// tensor-core kernels, instantiated in paged_attention_{f16,bf16}.cu
template <typename T, int KV, int NT, int HMAX, bool FAST>
__global__ void paged_attn_mma_kernel(const AttnArgs a);

template <typename T>
int launch_paged_attn_mma(const AttnArgs& a, int kv_dtype, dim3 grid, size_t smem,
                          cudaStream_t stream);

}  // namespace b200
