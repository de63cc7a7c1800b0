// Register-level unpack / dequant of Marlin-repacked weights (4-bit and 8-bit, symmetric).
//
// Layout facts (derived from marlin_perms.py:16-41 of the reference; checked bit-exactly by
// tests against marlin_weights()): a 16(k) x 64(n) tile is 128 (4-bit) or 256 (8-bit) int32.
// Lane i of a warp owns, for j = 0..3, the values
//     n = 16*j + i/4 + 8*b   (b = 0,1),   k = 2*(i%4) + {0,1} (+8 for the "high" pair).
//   4-bit: word [4*i + j], nibbles p0..p7 = (b0,k0) (b0,k0+8) (b1,k0) (b1,k0+8)
//                                           (b0,k0+1) (b0,k0+9) (b1,k0+1) (b1,k0+9)
//   8-bit: word [8*i + 2*j + b], bytes = k0, k0+8, k0+1, k0+9
// so one word (4-bit) / word pair (8-bit) yields four packed 16-bit pairs
//     r[0] = b0 {k0,k0+1}, r[1] = b0 {k0+8,k0+9}, r[2] = b1 {k0,k0+1}, r[3] = b1 {k0+8,k0+9}
// which is exactly the per-thread fragment of four 8x8 b16 matrices (rows = n, cols = k): one
// stmatrix.x4 writes them as 16-byte K-contiguous chunks into the K-major operand tile.
//
// Values are exact integers (q - 8 / q - 128) in fp16 / bf16; multiplying by the group scale with
// one rounding reproduces the reference's w_ref = (q - zp) * s bit for bit
// (quant_utils.py:73; reference device code csrc/quantization/gptq_marlin/gptq_marlin.cu:153-259).
#pragma once

#include "common.cuh"

namespace b200 {
namespace marlin {

template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
  return r;
}
constexpr int kAndOr = (0xf0 & 0xcc) | 0xaa;  // (a & b) | c

__device__ __forceinline__ uint32_t h2_sub(uint32_t a, uint32_t b) {
  __half2 r = __hsub2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c) {
  __half2 r = __hfma2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b),
                      *reinterpret_cast<__half2*>(&c));
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t bf2_fma(uint32_t a, uint32_t b, uint32_t c) {
  __nv_bfloat162 r = __hfma2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b),
                             *reinterpret_cast<__nv_bfloat162*>(&c));
  return *reinterpret_cast<uint32_t*>(&r);
}

template <typename T>
__device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t s) {
  if constexpr (sizeof(T) == 2 && DT<T>::id == B200_F16) {
    __half2 r = __hmul2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  } else {
    __nv_bfloat162 r = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  }
}
// broadcast the 16-bit scale at half index `idx` (0/1) of `w` to both halves
__device__ __forceinline__ uint32_t dup16(uint32_t w, int hi) {
  return hi ? __byte_perm(w, w, 0x3232) : __byte_perm(w, w, 0x1010);
}

// 4-bit word -> r[0..3] = exact (q - 8) pairs
template <typename T>
__device__ __forceinline__ void unpack4(uint32_t q, uint32_t (&r)[4]) {
  if constexpr (DT<T>::id == B200_F16) {
    // 0x6400 | q = 1024 + q ; high nibble of each byte: 1024 + 16 q -> * 1/16 - 72 = q - 8
    constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400;
    constexpr uint32_t SUB = 0x64086408, MUL = 0x2c002c00, ADD = 0xd480d480;
    const uint32_t q8 = q >> 8;
    r[0] = h2_sub(lop3<kAndOr>(q, LO, EX), SUB);
    r[1] = h2_fma(lop3<kAndOr>(q, HI, EX), MUL, ADD);
    r[2] = h2_sub(lop3<kAndOr>(q8, LO, EX), SUB);
    r[3] = h2_fma(lop3<kAndOr>(q8, HI, EX), MUL, ADD);
  } else {
    // 0x4300 | q = 128 + q (bf16) ; - 136 = q - 8
    constexpr uint32_t MASK = 0x000f000f, EX = 0x43004300, ONE = 0x3f803f80, ADD = 0xc308c308;
    r[0] = bf2_fma(lop3<kAndOr>(q, MASK, EX), ONE, ADD);
    r[1] = bf2_fma(lop3<kAndOr>(q >> 4, MASK, EX), ONE, ADD);
    r[2] = bf2_fma(lop3<kAndOr>(q >> 8, MASK, EX), ONE, ADD);
    r[3] = bf2_fma(lop3<kAndOr>(q >> 12, MASK, EX), ONE, ADD);
  }
}

// 8-bit word (bytes k0, k0+8, k0+1, k0+9) -> lo = {k0,k0+1}, hi = {k0+8,k0+9}, exact (q - 128)
template <typename T>
__device__ __forceinline__ void unpack8(uint32_t q, uint32_t& lo, uint32_t& hi) {
  if constexpr (DT<T>::id == B200_F16) {
    constexpr uint32_t EX = 0x64646464, SUB = 0x64806480;  // 1024 + q - 1152
    lo = h2_sub(__byte_perm(q, EX, 0x5250), SUB);
    hi = h2_sub(__byte_perm(q, EX, 0x5351), SUB);
  } else {
    const float f0 = (float)((int)(q & 0xff) - 128), f1 = (float)((int)((q >> 8) & 0xff) - 128);
    const float f2 = (float)((int)((q >> 16) & 0xff) - 128), f3 = (float)((int)(q >> 24) - 128);
    lo = DT<T>::pack2(f0, f2);
    hi = DT<T>::pack2(f1, f3);
  }
}

// fp8 e4m3 weights in the same 8-bit word (fp8_marlin_gemm, csrc/quantization/fp8/fp8_marlin.cu:
// 120-192): sign and exponent/mantissa bits are moved to their fp16 / bf16 positions; the result is
// the fp8 value times 2^-8 (fp16) or 2^-120 (bf16), exactly.  The missing power of two is folded
// into the channel scale (fp8_scale_bias), so that ONE multiply rounds value x scale once.
template <typename T>
__device__ __forceinline__ void unpack8_fp8(uint32_t q, uint32_t& lo, uint32_t& hi) {
  constexpr int SH = DT<T>::id == B200_F16 ? 1 : 4;                         // exponent 5 | 8 bits vs 4
  constexpr uint32_t EM = DT<T>::id == B200_F16 ? 0x3F803F80u : 0x07F007F0u;  // (0x7F007F00 >> SH)
  const uint32_t ql = q << 8;  // bytes 0, 2 -> the high byte of each half
  hi = lop3<kAndOr>(q >> SH, EM, q & 0x80008000u);
  lo = lop3<kAndOr>(ql >> SH, EM, ql & 0x80008000u);
}
template <typename T>
__device__ __forceinline__ uint32_t fp8_scale_bias(uint32_t s2) {  // s2 x 2^8 | 2^120 (exact)
  if constexpr (DT<T>::id == B200_F16) return mul2<T>(s2, 0x5C005C00u);  // 256.0
  else return mul2<T>(s2, 0x7B807B80u);                                   // 2^120 in bf16
}

// position of column n inside a grouped-scale row (marlin_perms.py:41-43: 64-column chunks,
// position 8*a + c holds column a + 8*c)
__host__ __device__ __forceinline__ int scale_pos_grouped(int n) {
  const int c64 = n & 63;
  return (n & ~63) + 8 * (c64 & 7) + (c64 >> 3);
}
// single-group rows (marlin_perms.py:44-47: 32-column chunks, position 8*i + t holds column
// 2*i + {0,1,8,9,16,17,24,25}[t])
__host__ __device__ __forceinline__ int scale_pos_single(int n) {
  const int c32 = n & 31;
  const int i = (c32 & 7) >> 1, t = 2 * (c32 >> 3) + (c32 & 1);
  return (n & ~31) + 8 * i + t;
}

__device__ __forceinline__ void stmatrix_x4(uint32_t smem_addr, const uint32_t (&r)[4]) {
  asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1,%2,%3,%4};" ::"r"(smem_addr),
               "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3])
               : "memory");
}

}  // namespace marlin
}  // namespace b200
