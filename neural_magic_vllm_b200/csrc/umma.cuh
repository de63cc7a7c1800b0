// tcgen05 / TMEM / TMA building blocks for the quantized GEMMs (sm_100a only).
//
// Conventions used by every GEMM in this library ("weights as the MMA A operand"):
//   D[n, m] (TMEM, 128 lanes = 128 output features n, one fp32/s32 column per token m)
//     += W[n, k] (shared memory, K-major, 128-byte swizzle)  x  X[m, k] (same layout)
// so the accumulator tile is 128 (n) x BM (tokens), the MMA shape is M=128, N=BM, and small-M
// decode wastes nothing: BM shrinks to 16 while every MMA still streams 128 weight rows.
#pragma once

#include <cuda.h>

#include "common.cuh"

namespace b200 {
namespace umma {

// ---- bounded mbarrier wait: a protocol bug traps instead of hanging the GPU ----------------------
// The bound is WALL TIME (globaltimer), not a spin count: under programmatic dependent launch a
// GEMM CTA can be resident -- its MMA / dequant warps already waiting here -- while its predecessor
// (e.g. a cross-GPU all-reduce polling a peer that is seconds late: first graph replay, a stalled
// host) is still running.  kMbarTimeoutNs is longer than any skew the all-reduce itself tolerates.
constexpr uint64_t kMbarTimeoutNs = 120ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
  // try_wait itself suspends the thread for a HW-defined interval; the clock is read only once
  // every 4096 failed probes
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0xfffu) == 0) {
      const uint64_t t = global_timer_ns();
      if (t0 == 0)
        t0 = t;
      else if (t - t0 > kMbarTimeoutNs)
        __trap();
    }
  }
}

// one lane of a converged warp (elect.sync): role loops run warp-uniform and only the issue of
// TMA / MMA instructions sits under this predicate, so their operands stay in uniform registers
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ---- TMEM allocation (one warp, power-of-two columns >= 32) ---------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
               : "memory");
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// ---- descriptors ----------------------------------------------------------------------------------
// K-major operand tile, rows of exactly 128 bytes (one swizzle atom along K), 8-row groups 1024 B
// apart, SWIZZLE_128B.  Field layout: cute/arch/mma_sm100_desc.hpp (SmemDescriptor): start>>4 in
// [0,14), LBO>>4 in [16,30) (unused for swizzled K-major), SBO>>4 in [32,46), version=1 at [46,48),
// layout type at [61,64) (2 = SWIZZLE_128B).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((1024u >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// advance along K inside the 128-byte atom: +bytes>>4 on the start-address field
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (bytes >> 4); }

enum Kind { KIND_F16 = 0, KIND_BF16 = 1, KIND_E4M3 = 2, KIND_I8 = 3 };

// InstrDescriptor (same header): c_format [4,6), a_format [7,10), b_format [10,13),
// a_major bit 15, b_major bit 16 (0 = K-major), n>>3 at [17,23), m>>4 at [24,29).
template <int KIND>
__host__ __device__ constexpr uint32_t instr_desc(int m, int n) {
  uint32_t c_fmt = (KIND == KIND_I8) ? 2u : 1u;           // S32 : F32
  uint32_t ab = (KIND == KIND_BF16 || KIND == KIND_I8) ? 1u : 0u;  // F16=0 BF16=1 ; E4M3=0 ; INT8=1
  return (c_fmt << 4) | (ab << 7) | (ab << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; one thread issues for the CTA
template <int KIND>
__device__ __forceinline__ void mma_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                       uint32_t idesc, uint32_t accumulate) {
  if constexpr (KIND == KIND_F16 || KIND == KIND_BF16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else if constexpr (KIND == KIND_E4M3) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}

// D[tmem] (+)= A[tmem] * B[smem] (kind::f16): A is M x K with row m in TMEM lane m and the 16-bit
// elements packed two per 32-bit column (K = 16 -> 8 columns at `tmem_a`); A cannot be transposed
__device__ __forceinline__ void mma_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// registers -> TMEM, shape 16x128b.x2: 16 lanes x 8 columns starting at `taddr`.  Thread t supplies
//   r0 = (lane t/4,     col t%4)   r1 = (lane t/4 + 8, col t%4)
//   r2 = (lane t/4,     col 4+t%4) r3 = (lane t/4 + 8, col 4+t%4)
// (cute/atom/copy_traits_sm100.hpp, SM100_TMEM_*_16dp128b2x) -- the m16n8k16 fragment order.
__device__ __forceinline__ void tmem_st_16x128b_x2(uint32_t taddr, uint32_t r0, uint32_t r1,
                                                   uint32_t r2, uint32_t r3) {
  asm volatile("tcgen05.st.sync.aligned.16x128b.x2.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(r0),
               "r"(r1), "r"(r2), "r"(r3)
               : "memory");
}
// shape 16x128b.x8: 16 lanes x 32 columns; unit u (4 columns) takes r[2u] = (lane t/4, col 4u + t%4)
// and r[2u+1] = (lane t/4 + 8, col 4u + t%4)
__device__ __forceinline__ void tmem_st_16x128b_x8(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// all previously issued MMAs of this thread arrive on `bar` when they complete
// (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

// ---- TMEM -> registers: 32 lanes x 16 consecutive 32-bit columns per warp -------------------------
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---- TMA: 2-D tiled tensor-map load into shared memory, completion on an mbarrier -----------------
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1,
                                            uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

// ---- host: tensor maps -------------------------------------------------------------------------------
// generic 2-D tiled map over a row-major [rows, cols] matrix; out-of-bounds elements read as zero
// (tails in every dimension need no special code) and still count towards the mbarrier's tx bytes.
inline int make_tmap_2d_ex(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                           uint64_t row_stride_bytes, CUtensorMapDataType dt, uint32_t box_cols,
                           uint32_t box_rows, CUtensorMapSwizzle swz) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  // the driver entry point is resolved through the runtime: the library has no link-time
  // dependency on libcuda.so.1, so it loads (and exports its symbols) on a box without a driver
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      set_error("cuTensorMapEncodeTiled is not available from this driver");
      return B200_ECUDA;
    }
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  CUresult r = encode(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%llu cols=%llu stride=%llu box=%ux%u)",
              (int)r, (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)row_stride_bytes, box_cols, box_rows);
    return B200_ECUDA;
  }
  return B200_OK;
}

// K-major MMA operand: box = box_rows x 128 bytes (one swizzle atom along the contiguous
// dimension), SWIZZLE_128B
inline int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                        uint64_t row_stride_bytes, int elem_bytes, uint32_t box_rows) {
  return make_tmap_2d_ex(map, base, rows, cols, row_stride_bytes,
                         elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16,
                         (uint32_t)(128 / elem_bytes), box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace umma
}  // namespace b200
