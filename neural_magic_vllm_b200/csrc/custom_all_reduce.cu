// One-shot all-reduce over NVLink peer memory for the latency-bound decode all-reduces of the
// row-parallel linears.  Replaces csrc/custom_all_reduce.cuh:179-253
// / custom_all_reduce.cu:50-148 of the reference (one-shot path) with a single kernel that has no
// grid-wide barrier and no host-side epoch:
//
//   every rank owns one IPC-shared buffer  = [ ArState | staging[2][max_bytes] ]
//   block b of rank r:  copies slice b of its input into its staging[parity]      (16-byte stores)
//                       release-stores `epoch` into flags[b][r] of EVERY rank      (system scope)
//                       acquire-spins until its own flags[b][*] all reached `epoch`
//                       sums slice b over the ranks' staging buffers in rank order (fp32, so every
//                       rank produces the same bits) and writes the result.
//   Block b only ever touches slice b, on every rank, so per-block flags are a sufficient barrier.
//   epoch lives in the rank's own state, one counter per block; the grid is always kArGrid blocks
//   and EVERY block advances its counter on every call (blocks beyond the active ones do nothing
//   else), so all counters move in lockstep = one call counter without a grid-wide barrier, and the
//   kernel is CUDA-graph replayable.  Staging is double-buffered by epoch parity: a rank can be at
//   most one call ahead of the slowest peer (it cannot pass call N's wait before every peer
//   STARTED call N, i.e. finished call N-1), so call N+1's writes never race call N-1's reads.
//   A protocol failure traps after a bounded spin instead of hanging the GPU.
//
// Messages up to kLLMaxPayload (the decode sizes: 8 KiB at batch 1 ... 2 MiB at batch 256) take the
// flag-in-data ("LL") kernel instead: every rank PUSHES its input into each peer's receive area as
// 16-byte lines {data, epoch, data, epoch} (each 8-byte half is valid on its own), and every thread
// polls its own lines of its own receive area until both epochs match, then sums in rank order.
// No staging copy, no fence, no block barrier: one posted NVLink store and local polling.
#include "common.cuh"

#include <cstring>

namespace b200 {
namespace {

constexpr int kArMaxRanks = 8;
constexpr int kArMaxBlocks = 128;
constexpr int kArThreads = 512;
constexpr int kArGrid = 72;  // fixed grid: see the epoch note above

struct ArState {
  uint32_t flags[kArMaxBlocks][kArMaxRanks];  // written by peers
  uint32_t epoch[kArMaxBlocks];               // local call counter per block
};
constexpr int64_t kArDataOffset = 8192;
constexpr int64_t kLLMaxPayload = 2048 * 1024;          // bytes of payload per call on the LL path
constexpr int64_t kLLSrcBytes = 2 * kLLMaxPayload;      // receive area per (parity, source rank)
constexpr int64_t kLLBytes = 2 * kArMaxRanks * kLLSrcBytes;
static_assert(sizeof(ArState) <= kArDataOffset, "state must fit in front of the staging area");

struct ArPeers {
  uint8_t* base[kArMaxRanks];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_volatile16(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

template <typename T, int W>
__global__ void __launch_bounds__(kArThreads) one_shot_all_reduce_kernel(
    const ArPeers peers, const int rank, const T* in, T* out, const int64_t nvec, const int active,
    const int64_t max_bytes) {
  constexpr int N = 16 / (int)sizeof(T);
  const int b = blockIdx.x, tid = threadIdx.x;
  ArState* me = reinterpret_cast<ArState*>(peers.base[rank]);
  __shared__ uint32_t s_epoch;
  if (tid == 0) s_epoch = me->epoch[b] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (b >= active) {  // idle block: keep the call counter in lockstep, nothing else
    if (tid == 0) me->epoch[b] = epoch;
    return;
  }
  const int64_t stage_off = kArDataOffset + (int64_t)(epoch & 1) * max_bytes;
  const int64_t per = (nvec + active - 1) / active;
  const int64_t lo = (int64_t)b * per;
  const int64_t hi = lo + per < nvec ? lo + per : nvec;

  uint4* mine = reinterpret_cast<uint4*>(peers.base[rank] + stage_off);
  const uint4* src = reinterpret_cast<const uint4*>(in);
  for (int64_t i = lo + tid; i < hi; i += kArThreads) mine[i] = src[i];
  __threadfence_system();
  __syncthreads();
  if (tid < W) {
    st_release_sys(&reinterpret_cast<ArState*>(peers.base[tid])->flags[b][rank], epoch);
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(&me->flags[b][tid]) - epoch) < 0) {
      if (++spins > (1u << 26)) __trap();  // a peer never arrived: fail loudly, do not hang
    }
  }
  __syncthreads();

  for (int64_t i = lo + tid; i < hi; i += kArThreads) {
    float acc[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < W; ++r) {
      union {
        uint4 u;
        T e[N];
      } v;
      v.u = ld_volatile16(reinterpret_cast<const uint4*>(peers.base[r] + stage_off) + i);
#pragma unroll
      for (int j = 0; j < N; ++j) acc[j] += DT<T>::to_f(v.e[j]);
    }
    union {
      uint4 u;
      T e[N];
    } o;
#pragma unroll
    for (int j = 0; j < N; ++j) o.e[j] = DT<T>::from_f(acc[j]);
    reinterpret_cast<uint4*>(out)[i] = o.u;
  }
  if (tid == 0) me->epoch[b] = epoch;
}


__device__ __forceinline__ void st_volatile16(uint4* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// unit = 8 bytes of payload = one 16-byte line.  Thread-independent: push, poll, sum.
// TWO (two-shot, the reference's second algorithm: custom_all_reduce.cuh:255-300): rank o owns the
// units [o * units / W, (o + 1) * units / W).  Every rank pushes its input units to their owners,
// the owner sums them in rank order and pushes the result into every rank's gather area (the upper
// half of the same per-source receive block), and every rank reads all results from there: 2 (W - 1)
// / W of the message crosses NVLink per rank instead of W - 1 times the message, for two flag waits
// instead of one.  Same bits as one-shot (fp32 sum in rank order, rounded once, by one rank).
template <typename T, int W, bool TWO>
__global__ void __launch_bounds__(kArThreads) ll_all_reduce_kernel(
    const ArPeers peers, const int rank, const T* in, T* out, const int64_t units, const int active,
    const int64_t ll_off) {
  constexpr int N = 8 / (int)sizeof(T);
  const int b = blockIdx.x, tid = threadIdx.x;
  ArState* me = reinterpret_cast<ArState*>(peers.base[rank]);
  __shared__ uint32_t s_epoch;
  // PDL: resident behind the row-parallel GEMM's tail; its partial sums (and the previous call's
  // epoch) are visible after the wait.  The kernel behind this one (a GEMM: its weight stages do
  // not depend on us) may start streaming while we exchange.
  pdl_wait();
  pdl_launch_dependents();
  if (tid == 0) s_epoch = me->epoch[b] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (tid == 0) me->epoch[b] = epoch;  // only this block reads it, next call
  if (b >= active) return;
  const int64_t par_off = ll_off + (int64_t)(epoch & 1) * kArMaxRanks * kLLSrcBytes;
  const uint2* src = reinterpret_cast<const uint2*>(in);
  const int64_t first = (int64_t)b * kArThreads + tid, step = (int64_t)active * kArThreads;
  if constexpr (TWO) {
    const int64_t seg = units / W;  // the host guarantees units % W == 0
    auto poll = [&](const uint4* slot) {
      uint4 l = ld_volatile16(slot);
      uint32_t spins = 0;
      while (l.y != epoch || l.w != epoch) {
        if (++spins > (1u << 26)) __trap();
        l = ld_volatile16(slot);
      }
      return make_uint2(l.x, l.z);
    };
    // A: my units of every other owner's segment -> that owner
    for (int64_t i = first; i < units; i += step) {
      const int o = (int)(i / seg);
      if (o == rank) continue;
      const uint2 d = src[i];
      st_volatile16(reinterpret_cast<uint4*>(peers.base[o] + par_off + (int64_t)rank * kLLSrcBytes) + (i - o * seg),
                    make_uint4(d.x, epoch, d.y, epoch));
    }
    // B: my segment -- sum in rank order, result to every rank's gather area (mine included)
    for (int64_t li = first; li < seg; li += step) {
      float acc[N];
#pragma unroll
      for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
      for (int r = 0; r < W; ++r) {
        union {
          uint2 u;
          T e[N];
        } v;
        v.u = r == rank ? src[rank * seg + li]
                        : poll(reinterpret_cast<const uint4*>(peers.base[rank] + par_off + (int64_t)r * kLLSrcBytes) + li);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] += DT<T>::to_f(v.e[j]);
      }
      union {
        uint2 u;
        T e[N];
      } o;
#pragma unroll
      for (int j = 0; j < N; ++j) o.e[j] = DT<T>::from_f(acc[j]);
      const uint4 line = make_uint4(o.u.x, epoch, o.u.y, epoch);
#pragma unroll
      for (int r = 0; r < W; ++r)
        st_volatile16(reinterpret_cast<uint4*>(peers.base[r] + par_off + (int64_t)rank * kLLSrcBytes + kLLSrcBytes / 2) + li,
                      line);
    }
    // C: all results from the local gather area
    for (int64_t i = first; i < units; i += step) {
      const int o = (int)(i / seg);
      reinterpret_cast<uint2*>(out)[i] = poll(reinterpret_cast<const uint4*>(
          peers.base[rank] + par_off + (int64_t)o * kLLSrcBytes + kLLSrcBytes / 2) + (i - o * seg));
    }
    return;
  }
  // pass 1: push everything (posted stores: NVLink bandwidth, not round trips, bounds this)
  for (int64_t i = first; i < units; i += step) {
    const uint2 d = src[i];
    const uint4 line = make_uint4(d.x, epoch, d.y, epoch);
#pragma unroll
    for (int r = 0; r < W; ++r) {
      if (r == rank) continue;
      st_volatile16(reinterpret_cast<uint4*>(peers.base[r] + par_off + (int64_t)rank * kLLSrcBytes) + i, line);
    }
  }
  // pass 2: poll the local receive area, sum in rank order
  for (int64_t i = first; i < units; i += step) {
    float acc[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < W; ++r) {
      union {
        uint2 u;
        T e[N];
      } v;
      if (r == rank) {
        v.u = src[i];
      } else {
        const uint4* slot = reinterpret_cast<const uint4*>(peers.base[rank] + par_off + (int64_t)r * kLLSrcBytes) + i;
        uint4 l = ld_volatile16(slot);
        uint32_t spins = 0;
        while (l.y != epoch || l.w != epoch) {
          if (++spins > (1u << 26)) __trap();  // a peer never arrived: fail loudly, do not hang
          l = ld_volatile16(slot);
        }
        v.u = make_uint2(l.x, l.z);
      }
#pragma unroll
      for (int j = 0; j < N; ++j) acc[j] += DT<T>::to_f(v.e[j]);
    }
    union {
      uint2 u;
      T e[N];
    } o;
#pragma unroll
    for (int j = 0; j < N; ++j) o.e[j] = DT<T>::from_f(acc[j]);
    reinterpret_cast<uint2*>(out)[i] = o.u;
  }
}

// Two-shot pays once the W - 1 copies of the message that one-shot pulls per rank cost more than a
// second flag round trip: world >= 4 and >= 128 KiB (measured, profiles/r2c_all_reduce.md).
// B200_AR_TWO_SHOT = 0 / 1 forces it off / on (tests run it at world 2).
inline bool use_two_shot(int world, int64_t bytes, bool divisible) {
  if (!divisible) return false;
  static const int knob = [] {
    const char* e = getenv("B200_AR_TWO_SHOT");
    return e ? atoi(e) : -1;
  }();
  if (knob >= 0) return knob != 0;
  return world >= 4 && bytes >= 128 * 1024;
}

template <typename T>
int launch_ll(const ArPeers& peers, int rank, int world, const void* in, void* out, int64_t units,
              int64_t ll_off, cudaStream_t st) {
  int64_t active = (units + kArThreads - 1) / kArThreads;
  active = active < 1 ? 1 : (active > kArGrid ? kArGrid : active);
  const bool two = use_two_shot(world, units * 8, units % world == 0);
#define LL_CASE(W)                                                       \
  case W:                                                                \
    if (two)                                                             \
      launch_pdl(ll_all_reduce_kernel<T, W, true>, dim3(kArGrid), dim3(kArThreads), 0, st, peers, rank, \
                 (const T*)in, (T*)out, units, (int)active, ll_off);                                   \
    else                                                                 \
      launch_pdl(ll_all_reduce_kernel<T, W, false>, dim3(kArGrid), dim3(kArThreads), 0, st, peers, rank, \
                 (const T*)in, (T*)out, units, (int)active, ll_off);                                    \
    break;
  switch (world) {
    LL_CASE(2)
    LL_CASE(4)
    LL_CASE(6)
    LL_CASE(8)
    default:
      set_error("all_reduce: unsupported world size %d (2, 4, 6, 8)", world);
      return B200_ENOTSUP;
  }
#undef LL_CASE
  return check_launch("all_reduce (LL)");
}

// ---------------------------------------------------------------------------------------------
// Fused: all-reduce of the row-parallel partial sums + residual add + RMSNorm, one kernel.
// The pair the decode layer issues after o_proj / down_proj (linear.py:792-793 followed by
// layernorm.py: fused_add_rms_norm).  Same flag-in-data push as ll_all_reduce_kernel; the thread
// that has summed a unit of a token row keeps it in registers, adds the residual, and the CTA
// that owns the row reduces the sum of squares -- the reduced activations never go to HBM between
// the collective and the norm.  Rounding steps are those of the two separate ops:
//   o = T(sum_r partial_r)  (fp32, rank order);  z = T(o + residual);  residual = z;
//   x = T(T(z * rstd) * weight).
// One CTA per token row (rows strided over the fixed grid); units per thread bounded by kRowUnits.
// ---------------------------------------------------------------------------------------------
constexpr int kRowUnits = 8;  // 8-byte units per thread per row: hidden <= 16384 (fp16/bf16)

template <typename T, int W, bool TWO>
__global__ void __launch_bounds__(kArThreads) ll_all_reduce_add_rms_norm_kernel(
    const ArPeers peers, const int rank, T* x, T* residual, const T* __restrict__ weight,
    const float eps, const int tokens, const int hidden, const int64_t ll_off) {
  constexpr int N = 8 / (int)sizeof(T);
  const int b = blockIdx.x, tid = threadIdx.x;
  ArState* me = reinterpret_cast<ArState*>(peers.base[rank]);
  __shared__ uint32_t s_epoch;
  __shared__ float red[32];
  pdl_wait();
  pdl_launch_dependents();
  if (tid == 0) s_epoch = me->epoch[b] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (tid == 0) me->epoch[b] = epoch;
  if (b >= tokens) return;
  const int64_t par_off = ll_off + (int64_t)(epoch & 1) * kArMaxRanks * kLLSrcBytes;
  const int row_units = hidden / N;
  // TWO (two-shot): rank o owns the units [o * seg, (o + 1) * seg) of every row (seg = row_units / W,
  // a multiple of 32: a warp's units have one owner); see ll_all_reduce_kernel
  const int seg = TWO ? row_units / W : row_units;
  auto poll = [&](const uint4* slot) {
    uint4 l = ld_volatile16(slot);
    uint32_t spins = 0;
    while (l.y != epoch || l.w != epoch) {
      if (++spins > (1u << 26)) __trap();
      l = ld_volatile16(slot);
    }
    return make_uint2(l.x, l.z);
  };
  if constexpr (TWO) {
    // A: the units other ranks own -> their owners
    for (int t = b; t < tokens; t += gridDim.x) {
      const uint2* src = reinterpret_cast<const uint2*>(x + (int64_t)t * hidden);
      for (int u = tid; u < row_units; u += kArThreads) {
        const int o = u / seg;
        if (o == rank) continue;
        const uint2 d = src[u];
        st_volatile16(reinterpret_cast<uint4*>(peers.base[o] + par_off + (int64_t)rank * kLLSrcBytes) +
                          ((int64_t)t * seg + (u - o * seg)),
                      make_uint4(d.x, epoch, d.y, epoch));
      }
    }
    // B: the units this rank owns -- sum in rank order, result into every rank's gather area
    for (int t = b; t < tokens; t += gridDim.x) {
      const uint2* src = reinterpret_cast<const uint2*>(x + (int64_t)t * hidden);
      for (int lu = tid; lu < seg; lu += kArThreads) {
        const int64_t li = (int64_t)t * seg + lu;
        float acc[N];
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
        for (int r = 0; r < W; ++r) {
          union {
            uint2 u2;
            T e[N];
          } v;
          v.u2 = r == rank ? src[rank * seg + lu]
                           : poll(reinterpret_cast<const uint4*>(peers.base[rank] + par_off + (int64_t)r * kLLSrcBytes) + li);
#pragma unroll
          for (int j = 0; j < N; ++j) acc[j] += DT<T>::to_f(v.e[j]);
        }
        union {
          uint2 u2;
          T e[N];
        } o;
#pragma unroll
        for (int j = 0; j < N; ++j) o.e[j] = DT<T>::from_f(acc[j]);  // the all-reduce's output rounding
        const uint4 line = make_uint4(o.u2.x, epoch, o.u2.y, epoch);
#pragma unroll
        for (int r = 0; r < W; ++r)
          st_volatile16(reinterpret_cast<uint4*>(peers.base[r] + par_off + (int64_t)rank * kLLSrcBytes + kLLSrcBytes / 2) + li,
                        line);
      }
    }
  } else {
    // pass 1: push this CTA's rows to every peer
    for (int t = b; t < tokens; t += gridDim.x) {
      const uint2* src = reinterpret_cast<const uint2*>(x + (int64_t)t * hidden);
      for (int u = tid; u < row_units; u += kArThreads) {
        const uint2 d = src[u];
        const uint4 line = make_uint4(d.x, epoch, d.y, epoch);
        const int64_t i = (int64_t)t * row_units + u;
#pragma unroll
        for (int r = 0; r < W; ++r) {
          if (r == rank) continue;
          st_volatile16(reinterpret_cast<uint4*>(peers.base[r] + par_off + (int64_t)rank * kLLSrcBytes) + i, line);
        }
      }
    }
  }
  // pass 2: per row -- poll + sum, add residual, sum of squares, normalise
  for (int t = b; t < tokens; t += gridDim.x) {
    uint2* xrow = reinterpret_cast<uint2*>(x + (int64_t)t * hidden);
    uint2* rrow = reinterpret_cast<uint2*>(residual + (int64_t)t * hidden);
    float z[kRowUnits][N];
    float ss = 0.f;
    // residual and weight do not depend on the peers: fetch them before polling
    const uint2* wrow = reinterpret_cast<const uint2*>(weight);
    uint2 rpre[kRowUnits], wpre[kRowUnits];
#pragma unroll
    for (int k = 0; k < kRowUnits; ++k) {
      const int u = tid + k * kArThreads;
      if (u < row_units) rpre[k] = rrow[u], wpre[k] = __ldg(wrow + u);
    }
#pragma unroll
    for (int k = 0; k < kRowUnits; ++k) {
      const int u = tid + k * kArThreads;
      if (u < row_units) {
        const int64_t i = (int64_t)t * row_units + u;
        float acc[N];
        if constexpr (TWO) {  // the owner's rounded sum, from the local gather area
          const int o = u / seg;
          union {
            uint2 u2;
            T e[N];
          } v;
          v.u2 = poll(reinterpret_cast<const uint4*>(peers.base[rank] + par_off + (int64_t)o * kLLSrcBytes + kLLSrcBytes / 2) +
                      ((int64_t)t * seg + (u - o * seg)));
#pragma unroll
          for (int j = 0; j < N; ++j) acc[j] = DT<T>::to_f(v.e[j]);
        } else {
#pragma unroll
          for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
          for (int r = 0; r < W; ++r) {
            union {
              uint2 u2;
              T e[N];
            } v;
            v.u2 = r == rank ? xrow[u]
                             : poll(reinterpret_cast<const uint4*>(peers.base[rank] + par_off + (int64_t)r * kLLSrcBytes) + i);
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] += DT<T>::to_f(v.e[j]);
          }
        }
        union {
          uint2 u2;
          T e[N];
        } rv, zo;
        rv.u2 = rpre[k];
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const float o = DT<T>::to_f(DT<T>::from_f(acc[j]));            // the all-reduce's output rounding
          zo.e[j] = DT<T>::from_f(__fadd_rn(o, DT<T>::to_f(rv.e[j])));  // fused_add: z = T(o + residual)
          z[k][j] = DT<T>::to_f(zo.e[j]);
          ss = fmaf(z[k][j], z[k][j], ss);
        }
        rrow[u] = zo.u2;
      }
    }
    // block-wide sum of squares
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    float tot = (tid & 31) < (kArThreads / 32) ? red[tid & 31] : 0.f;
    tot = warp_sum(tot);
    __syncthreads();  // red is reused by the next row
    const float rstd = rsqrtf(tot / hidden + eps);
#pragma unroll
    for (int k = 0; k < kRowUnits; ++k) {
      const int u = tid + k * kArThreads;
      if (u < row_units) {
        union {
          uint2 u2;
          T e[N];
        } wv, ov;
        wv.u2 = wpre[k];
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const float n1 = DT<T>::to_f(DT<T>::from_f(__fmul_rn(z[k][j], rstd)));
          ov.e[j] = DT<T>::from_f(__fmul_rn(n1, DT<T>::to_f(wv.e[j])));
        }
        xrow[u] = ov.u2;
      }
    }
  }
}

template <typename T>
int launch_ll_norm(const ArPeers& peers, int rank, int world, void* x, void* residual, const void* weight,
                   float eps, int tokens, int hidden, int64_t ll_off, cudaStream_t st) {
  constexpr int N = 8 / (int)sizeof(T);
  const int row_units = hidden / N;
  const bool two = use_two_shot(world, (int64_t)tokens * hidden * (int64_t)sizeof(T),
                                row_units % world == 0 && (row_units / world) % 32 == 0);
#define LLN_CASE(W)                                                                          \
  case W:                                                                                    \
    if (two)                                                                                 \
      launch_pdl(ll_all_reduce_add_rms_norm_kernel<T, W, true>, dim3(kArGrid), dim3(kArThreads), 0, st,  \
                 peers, rank, (T*)x, (T*)residual, (const T*)weight, eps, tokens, hidden, ll_off);      \
    else                                                                                     \
      launch_pdl(ll_all_reduce_add_rms_norm_kernel<T, W, false>, dim3(kArGrid), dim3(kArThreads), 0, st, \
                 peers, rank, (T*)x, (T*)residual, (const T*)weight, eps, tokens, hidden, ll_off);      \
    break;
  switch (world) {
    LLN_CASE(2)
    LLN_CASE(4)
    LLN_CASE(6)
    LLN_CASE(8)
    default:
      set_error("all_reduce_add_rms_norm: unsupported world size %d (2, 4, 6, 8)", world);
      return B200_ENOTSUP;
  }
#undef LLN_CASE
  return check_launch("all_reduce_add_rms_norm");
}

template <typename T>
int launch_ar(const ArPeers& peers, int rank, int world, const void* in, void* out, int64_t nvec,
              int64_t max_bytes, cudaStream_t st) {
  int64_t active = (nvec + kArThreads * 4 - 1) / (kArThreads * 4);
  active = active < 1 ? 1 : (active > kArGrid ? kArGrid : active);  // same on every rank (f(nvec))
#define AR_CASE(W)                                                        \
  case W:                                                                 \
    one_shot_all_reduce_kernel<T, W><<<kArGrid, kArThreads, 0, st>>>(     \
        peers, rank, (const T*)in, (T*)out, nvec, (int)active, max_bytes); \
    break;
  switch (world) {
    AR_CASE(2)
    AR_CASE(4)
    AR_CASE(6)
    AR_CASE(8)
    default:
      set_error("all_reduce: unsupported world size %d (2, 4, 6, 8)", world);
      return B200_ENOTSUP;
  }
#undef AR_CASE
  return check_launch("all_reduce");
}

}  // namespace
}  // namespace b200

extern "C" {

int64_t b200_ar_state_bytes(void) { return b200::kArDataOffset; }

int b200_ar_alloc(int64_t max_bytes, void** ptr, unsigned char* ipc_handle) {
  using namespace b200;
  B200_REQUIRE(ptr && ipc_handle && max_bytes > 0 && max_bytes % 16 == 0, B200_EINVAL,
               "ar_alloc: bad argument");
  const size_t total = (size_t)kArDataOffset + 2 * (size_t)max_bytes + (size_t)kLLBytes;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, total);
  if (e == cudaSuccess) e = cudaMemset(p, 0, total);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    if (p) cudaFree(p);
    set_error("ar_alloc: %s", cudaGetErrorString(e));
    return B200_ECUDA;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  memcpy(ipc_handle, &h, sizeof(h));
  *ptr = p;
  return B200_OK;
}

int b200_ar_open(const unsigned char* ipc_handle, void** ptr) {
  using namespace b200;
  B200_REQUIRE(ptr && ipc_handle, B200_EINVAL, "ar_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("ar_open: %s", cudaGetErrorString(e));
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_ar_close(void* ptr) {
  if (ptr && cudaIpcCloseMemHandle(ptr) != cudaSuccess) {
    cudaGetLastError();
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_ar_free(void* ptr) {
  if (ptr && cudaFree(ptr) != cudaSuccess) {
    cudaGetLastError();
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_all_reduce(void* const* bases, int32_t rank, int32_t world, const void* input, void* output,
                    int64_t numel, int32_t dtype, int64_t max_bytes, void* stream) {
  using namespace b200;
  B200_REQUIRE(bases && world >= 2 && world <= kArMaxRanks && rank >= 0 && rank < world, B200_EINVAL,
               "all_reduce: bad rank/world (%d/%d)", rank, world);
  if (numel == 0) return B200_OK;
  B200_REQUIRE(input && output && numel > 0, B200_EINVAL, "all_reduce: null pointer");
  const int64_t esz = dtype == B200_F32 ? 4 : 2;
  const int64_t bytes = numel * esz;
  // custom_all_reduce.cu:37-48 of the reference: 16-byte multiples only, bounded by the buffer
  B200_REQUIRE(bytes % 16 == 0 && bytes <= max_bytes, B200_ENOTSUP,
               "all_reduce: %lld bytes is not a multiple of 16 within max_size %lld", (long long)bytes,
               (long long)max_bytes);
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(input) | reinterpret_cast<uintptr_t>(output)) & 15) == 0,
               B200_ENOTSUP, "all_reduce: input/output must be 16-byte aligned");
  ArPeers peers{};
  for (int r = 0; r < world; ++r) {
    B200_REQUIRE(bases[r] != nullptr, B200_EINVAL, "all_reduce: null peer buffer %d", r);
    peers.base[r] = static_cast<uint8_t*>(bases[r]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t ll_off = kArDataOffset + 2 * max_bytes;
  const bool ll = bytes <= kLLMaxPayload;  // same decision on every rank (function of the size)
  switch (dtype) {
    case B200_F16:
      return ll ? launch_ll<__half>(peers, rank, world, input, output, bytes / 8, ll_off, st)
                : launch_ar<__half>(peers, rank, world, input, output, bytes / 16, max_bytes, st);
    case B200_BF16:
      return ll ? launch_ll<__nv_bfloat16>(peers, rank, world, input, output, bytes / 8, ll_off, st)
                : launch_ar<__nv_bfloat16>(peers, rank, world, input, output, bytes / 16, max_bytes, st);
    case B200_F32:
      return ll ? launch_ll<float>(peers, rank, world, input, output, bytes / 8, ll_off, st)
                : launch_ar<float>(peers, rank, world, input, output, bytes / 16, max_bytes, st);
  }
  set_error("all_reduce: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

int b200_all_reduce_add_rms_norm(void* const* bases, int32_t rank, int32_t world, void* input,
                                 void* residual, const void* weight, float epsilon,
                                 int32_t num_tokens, int32_t hidden, int32_t dtype, int64_t max_bytes,
                                 void* stream) {
  using namespace b200;
  B200_REQUIRE(bases && world >= 2 && world <= kArMaxRanks && rank >= 0 && rank < world, B200_EINVAL,
               "all_reduce_add_rms_norm: bad rank/world (%d/%d)", rank, world);
  B200_REQUIRE(num_tokens >= 0 && hidden >= 0, B200_EINVAL, "all_reduce_add_rms_norm: negative size");
  if (num_tokens == 0 || hidden == 0) return B200_OK;
  B200_REQUIRE(input && residual && weight, B200_EINVAL, "all_reduce_add_rms_norm: null pointer");
  const int64_t esz = dtype == B200_F32 ? 4 : 2;
  const int64_t row_bytes = (int64_t)hidden * esz, bytes = row_bytes * num_tokens;
  B200_REQUIRE(row_bytes % 8 == 0 && row_bytes / 8 <= (int64_t)kRowUnits * kArThreads &&
                   bytes <= kLLMaxPayload && bytes <= max_bytes,
               B200_ENOTSUP,
               "all_reduce_add_rms_norm: %d x %d does not fit the fused kernel (row <= %d bytes, total "
               "<= %lld); call all_reduce + fused_add_rms_norm",
               num_tokens, hidden, kRowUnits * kArThreads * 8, (long long)kLLMaxPayload);
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(input) | reinterpret_cast<uintptr_t>(residual) |
                 reinterpret_cast<uintptr_t>(weight)) & 7) == 0,
               B200_ENOTSUP, "all_reduce_add_rms_norm: pointers must be 8-byte aligned");
  ArPeers peers{};
  for (int r = 0; r < world; ++r) {
    B200_REQUIRE(bases[r] != nullptr, B200_EINVAL, "all_reduce_add_rms_norm: null peer buffer %d", r);
    peers.base[r] = static_cast<uint8_t*>(bases[r]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t ll_off = kArDataOffset + 2 * max_bytes;
  switch (dtype) {
    case B200_F16: return launch_ll_norm<__half>(peers, rank, world, input, residual, weight, epsilon, num_tokens, hidden, ll_off, st);
    case B200_BF16: return launch_ll_norm<__nv_bfloat16>(peers, rank, world, input, residual, weight, epsilon, num_tokens, hidden, ll_off, st);
    case B200_F32: return launch_ll_norm<float>(peers, rank, world, input, residual, weight, epsilon, num_tokens, hidden, ll_off, st);
  }
  set_error("all_reduce_add_rms_norm: unsupported dtype %d", dtype);
  return B200_EINVAL;
}

}  // extern "C"
