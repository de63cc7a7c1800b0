// gptq_marlin_gemm decode kernel, second generation: persistent, stream-K, one CTA per SM.
// (included by marlin_gemm.cu; same operator, same layouts, same bit-exact dequant as the kernels
// there -- replaces the stripe loop of csrc/quantization/gptq_marlin/gptq_marlin.cu:933-988.)
//
// Decode (M <= 64) is weight streaming: the only thing that matters is that every SM pulls its
// 1/148th of the packed weights at the HBM rate and that the dequant warps keep up.
//
//   work      the (256-column n tile, 64-k block) units of the layer are laid out tile-major and cut
//             into `grid` equal contiguous ranges, one per CTA (stream-K): every SM streams the same
//             number of bytes whatever the layer shape.  A range covers the tail of one tile, whole
//             tiles, and the head of another ("pieces").
//   warp 0    TMA producer: per unit one activation tile [BM x 64] (128-byte swizzle), the packed
//             weights (1 KiB rows of the Marlin tensor) and one group-scale row into a ring of
//             ~20 stages (~210 KiB in flight per SM -- an epilogue never starves HBM)
//   warps 1, 22  MMA issuers, one per 128-row sub-tile (each accumulator has ONE issuer: fixed
//             summation order): tcgen05.mma kind::f16, A = dequantised weights in TMEM, B =
//             activations in shared memory, D = fp32 in TMEM, double-buffered per piece
//   warps 2-17  four dequant teams of four warps (one per TMEM quadrant).  A job is one 128 x 64
//             sub-tile of a stage; jobs go round-robin to the teams, each team owns two 32-column
//             A slots in TMEM, so four jobs are in flight while the MMAs consume in order.
//             LDS.64 packed words (fetched one job ahead) -> lop3 / hsub2 / hfma2 exact (q - 8) ->
//             x group scale (one rounding: bit-identical to the reference's w_ref) -> tcgen05.st
//             16x128b.x8 straight to TMEM.
//   warps 18-21 epilogue: tcgen05.ld the finished piece; a whole tile is rounded once and stored.
//             A tile split over several CTAs is fixed up by the CTA that owns its lowest k range
//             (that piece is the last of its range; the others' pieces come first in theirs):
//             publishers store fp32 pieces to their slot and bump the tile's counter with one
//             red.release.gpu; the fixer waits for parts - 1 arrivals, sums its own piece (from
//             TMEM) and the slots in CTA order (deterministic), rounds once, stores.  The counters
//             are the caller's `workspace` (zero on entry, zero on exit, as the reference
//             requires: gptq_marlin.cu:329-332).
// (no include guard / namespaces: this file is spliced into marlin_gemm.cu's b200::{anonymous})

constexpr int kSNT = 2;                       // 128-row sub-tiles per n tile
constexpr int kSCtaN = kSNT * kTileN;         // 256
constexpr int kSTeams = 4;
constexpr int kSDequantWarps = kSTeams * 4;
constexpr int kSEpiWarps = 4;
constexpr int kSMma1Warp = 2 + kSDequantWarps + kSEpiWarps;       // 22: MMA issuer of sub-tile 1
constexpr int kSThreads = (kSMma1Warp + 1) * 32;                  // 736
constexpr int kSASlots = 2 * kSTeams;         // TMEM A slots of 32 columns (64 k)
constexpr int kSDCol = kSASlots * kTmemAStageCols;  // 256: first accumulator column
constexpr int kSMaxStages = 24;

template <int BITS, int BM>
struct PlanS {
  static constexpr int kPackedBytes = kStageK * kSCtaN * BITS / 8;  // 8 KiB / 16 KiB
  static constexpr int kScaleBytes = kSCtaN * 2;                    // one group-scale row
  static constexpr int kXBytes = BM * 128;
  static constexpr int kStageBytes = (kXBytes + kPackedBytes + kScaleBytes + 1023) / 1024 * 1024;
  static constexpr int kBudget = 224 * 1024;
  static constexpr int kStages = kBudget / kStageBytes > kSMaxStages ? kSMaxStages : kBudget / kStageBytes;
  static constexpr int kBarBytes = (2 * kStages + 2 * kSASlots + 4) * 8 + 64;
  static constexpr int kBytes = kStages * kStageBytes + 1024 + kBarBytes;
  static_assert(kStages >= 4, "ring too shallow");
  static_assert(kSDCol + 2 * kSNT * BM <= 512, "TMEM columns");
};

// Optional cycle accounting per role (build with -DB200_STREAM_PROF; read back through
// b200_debug_stream_prof): where each warp role of one CTA spends its time.
#ifdef B200_STREAM_PROF
__device__ unsigned long long g_stream_prof[160 * 32];
#define SPROF_DECL long long sp_t = clock64(), sp_n; unsigned long long sp_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define SPROF(i) (sp_n = clock64(), sp_acc[i] += (unsigned long long)(sp_n - sp_t), sp_t = sp_n)
#define SPROF_ENTRY const long long sp_entry = clock64()
#define SPROF_MARK(slot) g_stream_prof[blockIdx.x * 32 + (slot)] = (unsigned long long)(clock64() - sp_entry)
#define SPROF_FLUSH(base)                                                              \
  do {                                                                                 \
    for (int sp_i = 0; sp_i < 8; ++sp_i) g_stream_prof[blockIdx.x * 32 + (base) + sp_i] = sp_acc[sp_i]; \
  } while (0)
#else
#define SPROF_DECL
#define SPROF(i)
#define SPROF_FLUSH(base)
#define SPROF_ENTRY
#define SPROF_MARK(slot)
#endif

struct StreamArgs {
  int n_tiles;      // ceil(n / 256)
  int grid;         // CTAs = ranges
  int units;        // n_tiles * num_kb
};

// range start of CTA c; CTA that owns unit u
__device__ __forceinline__ uint32_t s_range_start(const StreamArgs& s, int c) {
  return (uint32_t)c * (uint32_t)s.units / (uint32_t)s.grid;
}
// (the host guarantees (units + 1) * grid < 2^31: 32-bit divisions)
__device__ __forceinline__ int s_cta_of_unit(const StreamArgs& s, int u) {
  return (int)(((uint32_t)(u + 1) * (uint32_t)s.grid - 1u) / (uint32_t)s.units);
}

__device__ __forceinline__ void red_release_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int BITS, int BM, typename T, bool GROUPED>
__global__ void __launch_bounds__(kSThreads, 1)
marlin_gemm_stream_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                          const __grid_constant__ CUtensorMap tm_s, const MarlinArgs args, const StreamArgs sa) {
  using P = PlanS<BITS, BM>;
  constexpr int kStages = P::kStages;
  constexpr int KIND = DT<T>::id == B200_F16 ? KIND_F16 : KIND_BF16;
  constexpr int kDCols = kSNT * BM;  // accumulator columns of one piece
  SPROF_ENTRY;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stages = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kStages * P::kStageBytes);
  uint64_t* full_p = bars;
  uint64_t* empty_p = full_p + kStages;
  uint64_t* full_a = empty_p + kStages;
  uint64_t* empty_a = full_a + kSASlots;
  uint64_t* tmem_full = empty_a + kSASlots;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  // warp-uniform by construction: say so, so that role code keeps its operands in uniform registers
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int num_kb = args.num_kb;
  const int u0 = (int)s_range_start(sa, blockIdx.x), u1 = (int)s_range_start(sa, blockIdx.x + 1);
  const int total = u1 - u0;
  const int tile0 = u0 / num_kb, kb0 = u0 - tile0 * num_kb;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_w);
    if (args.s_rows > 0) tma_prefetch_desc(&tm_s);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_p[s], 1);
      mbar_init(&empty_p[s], kSNT * 4 + kSNT);
    }
    for (int s = 0; s < kSASlots; ++s) {
      mbar_init(&full_a[s], 4);
      mbar_init(&empty_a[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], kSNT);
      mbar_init(&tmem_empty[s], kSEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  if (threadIdx.x == 0) SPROF_MARK(6);

  if (warp == 0) {
    // ======================= TMA producer =======================
    // The whole warp runs the loops (uniform control flow: coordinates, shared-memory addresses and
    // barrier addresses stay in uniform registers); one elected lane issues.
    const bool leader = elect_one();
    const uint64_t pol_w = l2_policy_evict_first();  // streamed exactly once
    const uint64_t pol_x = l2_policy_evict_last();   // every CTA re-reads the activations
    constexpr int kWordsPerN = 16 * BITS / 32;
    constexpr int kLoadsW = kSCtaN * kWordsPerN / 256;
    const uint32_t tx_bytes = P::kXBytes + P::kPackedBytes + (uint32_t)args.s_rows * (kSCtaN * 2);
    const bool staged_scales = args.s_rows > 0;
    // Programmatic dependent launch: weights and scales do not depend on the predecessor grid, so
    // the first pass over the ring streams them in before griddepcontrol.wait; activations after.
    const int pre = min(total, kStages);
    {
      int t = tile0, k = kb0;
      for (int i = 0; i < pre; ++i) {
        uint8_t* w_dst = stages + i * P::kStageBytes + P::kXBytes;
        if (leader) {
          mbar_arrive_expect_tx(&full_p[i], tx_bytes);
#pragma unroll
          for (int l = 0; l < kLoadsW; ++l)
            tma_load_2d(w_dst + l * 4096, &tm_w, t * (kSCtaN * kWordsPerN) + l * 256, k * 4, &full_p[i], pol_w);
          if (staged_scales)
            tma_load_2d(w_dst + P::kPackedBytes, &tm_s, t * kSCtaN, (k * kStageK) >> args.group_shift, &full_p[i],
                        pol_x);
        }
        if (++k == num_kb) k = 0, ++t;
      }
    }
    pdl_wait();
    int tile = tile0, kb = kb0;
    for (int i = 0; i < pre; ++i) {
      if (leader) tma_load_2d(stages + i * P::kStageBytes, &tm_x, kb * kStageK, 0, &full_p[i], pol_x);
      if (++kb == num_kb) kb = 0, ++tile;
    }
    int stage = pre == kStages ? 0 : pre;
    uint32_t phase = pre == kStages ? 1 : 0;
    SPROF_DECL;
    for (int i = pre; i < total; ++i) {
      SPROF(1);
      mbar_wait_bounded(&empty_p[stage], phase ^ 1);
      SPROF(0);
      uint8_t* x_dst = stages + stage * P::kStageBytes;
      uint8_t* w_dst = x_dst + P::kXBytes;
      if (leader) {
        mbar_arrive_expect_tx(&full_p[stage], tx_bytes);
        tma_load_2d(x_dst, &tm_x, kb * kStageK, 0, &full_p[stage], pol_x);
#pragma unroll
        for (int l = 0; l < kLoadsW; ++l)
          tma_load_2d(w_dst + l * 4096, &tm_w, tile * (kSCtaN * kWordsPerN) + l * 256, kb * 4, &full_p[stage], pol_w);
        if (staged_scales)
          tma_load_2d(w_dst + P::kPackedBytes, &tm_s, tile * kSCtaN, (kb * kStageK) >> args.group_shift,
                      &full_p[stage], pol_x);
      }
      if (++kb == num_kb) kb = 0, ++tile;
      if (++stage == kStages) stage = 0, phase ^= 1;
    }
    SPROF(1);
    if (lane == 0) SPROF_FLUSH(16);
  } else if (warp == 1 || warp == kSMma1Warp) {
    // ======================= MMA issuers (one warp per 128-row sub-tile) =======================
    // Uniform control flow again: all lanes wait on the barrier, one elected lane issues the MMAs
    // with descriptors / TMEM addresses computed in uniform registers.  A full_a arrival implies
    // the stage's full_p (the dequant warps waited on it first), so the activation tile is visible
    // without a second barrier wait.  One issuer per accumulator keeps the summation order fixed.
    const int h = warp == 1 ? 0 : 1;
    const bool leader = elect_one();
    constexpr uint32_t idesc = instr_desc<KIND>(kTileN, BM);
    int stage = 0;
    uint32_t job = h, piece = 0;
    int done = 0, kb = kb0;
    SPROF_DECL;
    while (done < total) {
      const int cnt = min(num_kb - kb, total - done);
      const uint32_t buf = piece & 1;
      SPROF(2);
      if (piece >= 2) mbar_wait_bounded(&tmem_empty[buf], ((piece >> 1) - 1) & 1);
      SPROF(3);
      const uint32_t d_col = tmem_base + kSDCol + buf * kDCols + h * BM;
      for (int i = 0; i < cnt; ++i) {
        const uint32_t aslot = job & (kSASlots - 1);
        const uint64_t dx = smem_desc_sw128(smem_u32(stages + stage * P::kStageBytes));
        const uint32_t a_col = tmem_base + aslot * kTmemAStageCols;
        SPROF(2);
        mbar_wait_bounded(&full_a[aslot], (job >> 3) & 1);
        SPROF(1);
        fence_after_sync();
        if (leader) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_ts_f16(d_col, a_col + k * 8, desc_advance(dx, k * 32), idesc, (i > 0 || k > 0) ? 1u : 0u);
          mma_commit(&empty_a[aslot]);
          mma_commit(&empty_p[stage]);
        }
        job += kSNT;
        if (++stage == kStages) stage = 0;
      }
      if (leader) mma_commit(&tmem_full[buf]);
      ++piece;
      done += cnt;
      kb = 0;
    }
    pdl_launch_dependents();
    SPROF(2);
    if (warp == 1 && lane == 0) {
      SPROF_FLUSH(0);
      SPROF_MARK(7);
    }
  } else if (warp < 2 + kSDequantWarps) {
    // ======================= dequant teams =======================
    // job (unit i, sub-tile h) has index 2 i + h and goes to team (2 i + h) % 4: team t takes
    // sub-tile t & 1 of the units i = t >> 1 (mod 2)
    const int team = (warp - 2) >> 2;
    const int h = team & 1;
    const int qd = warp & 3;          // TMEM quadrant = rows 32 qd .. 32 qd + 31 of the 128-row sub-tile
    const int nt = qd >> 1;           // 64-column Marlin tile inside the sub-tile
    const int j0 = (qd & 1) * 2;      // first of its two 16-row fragments
    const int a_row = lane >> 2;
    uint32_t sc[4];  // sc[2*jj + b]: scale of n = 64 nt + 16 (j0 + jj) + a_row + 8 b, in both halves
#pragma unroll
    for (int i = 0; i < 4; ++i) sc[i] = DT<T>::pack2(1.f, 1.f);
    int sc_tile = -1;  // tile the channel-wise scales in sc[] belong to
    const uint32_t lane_src = smem_u32(stages) + P::kXBytes + chunk_off<BITS>(0, 2 * h + nt) + lane * (BITS * 4) +
                              (qd & 1) * (BITS * 2);
    const uint32_t scale_src = smem_u32(stages) + P::kXBytes + P::kPackedBytes + (h * kTileN + nt * 64 + 8 * a_row) * 2 +
                               (qd & 1) * 8;
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
    int i = team >> 1;
    int stage = i % kStages;
    uint32_t phase = (i / kStages) & 1;
    int tile = tile0, kb = kb0 + i;
    while (kb >= num_kb) kb -= num_kb, ++tile;
    SPROF_DECL;
    constexpr int kW = BITS == 4 ? 2 : 4;
    // software pipeline over two register sets: the packed words (and scales) of the team's NEXT
    // job are fetched from shared memory before the current job is unpacked
    auto fetch = [&](uint32_t (&w)[4][kW], uint2& sv, int st) {
      const uint32_t w_src = lane_src + st * P::kStageBytes;
#pragma unroll
      for (int kt = 0; kt < 4; ++kt) {
        if constexpr (BITS == 4) {
          const uint2 v = lds64(w_src + kt * 1024);
          w[kt][0] = v.x, w[kt][1] = v.y;
        } else {
          const uint4 v = lds128(w_src + kt * 1024);
          w[kt][0] = v.x, w[kt][1] = v.y, w[kt][2] = v.z, w[kt][3] = v.w;
        }
      }
      if constexpr (GROUPED) sv = lds64(scale_src + st * P::kStageBytes);
    };
    auto job_step = [&](uint32_t (&cur)[4][kW], uint2& cur_sv, uint32_t (&nxt)[4][kW], uint2& nxt_sv) {
      const uint32_t job = 2 * i + h;
      const uint32_t aslot = job & (kSASlots - 1);
      SPROF(5);
      int nstage = stage + 2;
      uint32_t nphase = phase;
      if (nstage >= kStages) nstage -= kStages, nphase ^= 1;
      if constexpr (GROUPED) {
        sc[0] = dup16(cur_sv.x, 0), sc[1] = dup16(cur_sv.x, 1), sc[2] = dup16(cur_sv.y, 0), sc[3] = dup16(cur_sv.y, 1);
      } else if (sc_tile != tile) {
        sc_tile = tile;
        const uint16_t* s = reinterpret_cast<const uint16_t*>(args.scales);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int n = tile * kSCtaN + h * kTileN + nt * 64 + 16 * (j0 + (e >> 1)) + a_row + 8 * (e & 1);
          const uint32_t v = n < args.n ? s[scale_pos_single(n)] : 0u;
          sc[e] = v | (v << 16);
          if (BITS == 8 && args.fp8) sc[e] = fp8_scale_bias<T>(sc[e]);
        }
      }
      mbar_wait_bounded(&empty_a[aslot], ((job >> 3) & 1) ^ 1);
      SPROF(2);
      fence_after_sync();
      const uint32_t t_dst = t_lane + aslot * kTmemAStageCols;
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        // 16 rows x 64 k of the job = 32 TMEM columns = one tcgen05.st 16x128b.x8:
        // unit 2 kt = {b0, b1} of the k-lo quad, unit 2 kt + 1 = {b0, b1} of the k-hi quad
        uint32_t st[16];
#pragma unroll
        for (int kt = 0; kt < 4; ++kt) {
          uint32_t rg[4];
          if constexpr (BITS == 4) {
            unpack4<T>(cur[kt][jj], rg);
          } else {
            if (args.fp8) {
              unpack8_fp8<T>(cur[kt][2 * jj], rg[0], rg[1]);
              unpack8_fp8<T>(cur[kt][2 * jj + 1], rg[2], rg[3]);
            } else {
              unpack8<T>(cur[kt][2 * jj], rg[0], rg[1]);
              unpack8<T>(cur[kt][2 * jj + 1], rg[2], rg[3]);
            }
          }
          st[4 * kt + 0] = mul2<T>(rg[0], sc[2 * jj]);
          st[4 * kt + 2] = mul2<T>(rg[1], sc[2 * jj]);
          st[4 * kt + 1] = mul2<T>(rg[2], sc[2 * jj + 1]);
          st[4 * kt + 3] = mul2<T>(rg[3], sc[2 * jj + 1]);
        }
        tmem_st_16x128b_x8(t_dst + ((uint32_t)(16 * jj) << 16), st);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_p[stage]);  // every lane has consumed its packed words + scales
      SPROF(3);
      // the next job's operands are fetched while the tcgen05.st above drain into TMEM
      if (i + 2 < total) {
        mbar_wait_bounded(&full_p[nstage], nphase);
        fetch(nxt, nxt_sv, nstage);
      }
      SPROF(0);
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_a[aslot]);
      SPROF(4);
      kb += 2;
      while (kb >= num_kb) kb -= num_kb, ++tile;
      stage = nstage, phase = nphase;
      i += 2;
    };
    uint32_t wa[4][kW], wb[4][kW];
    uint2 sva = make_uint2(0u, 0u), svb = make_uint2(0u, 0u);
    if (i < total) {
      mbar_wait_bounded(&full_p[stage], phase);
      fetch(wa, sva, stage);
    }
    while (i < total) {
      job_step(wa, sva, wb, svb);
      if (i >= total) break;
      job_step(wb, svb, wa, sva);
    }
    SPROF(5);
    if (warp == 2 && lane == 0) SPROF_FLUSH(8);
  } else {
    // ======================= epilogue + split-K fix-up =======================
    // A whole tile is rounded once and stored.  A tile that is split over several CTAs is fixed up
    // by the CTA that owns its LOWEST k range: the unit order makes that piece the last one of its
    // CTA's range, while the pieces of the higher k ranges come first in theirs -- the other
    // contributors have usually published long before the fixer gets there.
    //   publisher (piece starts at kb > 0): fp32 piece -> its slot ([token][column]: one 128-byte
    //     line per warp access) -> barrier -> one gpu-scope release increment of the tile's counter
    //   fixer (piece starts at kb = 0): waits for parts - 1 arrivals (acquire), then sums in CTA
    //     order -- its own piece straight from TMEM first, then the slots -- rounds once, stores,
    //     and hands the counter back as zero.  Fixed order: deterministic.
    const int qd = warp & 3;
    const int et = (warp - 2 - kSDequantWarps) * 32 + lane;  // 0..127
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16) + kSDCol;
    constexpr int kSlotFloats = BM * kSCtaN;
    float* my_slot = args.partial + (int64_t)blockIdx.x * kSlotFloats;
    T* out = reinterpret_cast<T*>(args.c);
    uint32_t piece = 0;
    int done = 0, tile = tile0, kb = kb0;
    SPROF_DECL;
    while (done < total) {
      const int cnt = min(num_kb - kb, total - done);
      const bool whole = cnt == num_kb;
      const bool fixer = !whole && kb == 0;
      const uint32_t buf = piece & 1;
      int c_hi = 0;
      if (fixer) {
        // contributors: this CTA (lowest k) .. c_hi; they publish parts - 1 pieces
        c_hi = s_cta_of_unit(sa, (tile + 1) * num_kb - 1);
        if (et == 0) {
          const int need = c_hi - (int)blockIdx.x;
          uint32_t spins = 0;
          uint64_t t0 = 0;
          while (ld_acquire_gpu(&args.locks[tile]) < need) {
            if ((++spins & 0x3ffu) == 0) {
              const uint64_t t = global_timer_ns();
              if (t0 == 0) t0 = t;
              else if (t - t0 > kMbarTimeoutNs) __trap();
            }
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // thread 0's acquire covers the role
      }
      SPROF(1);
      mbar_wait_bounded(&tmem_full[buf], (piece >> 1) & 1);
      SPROF(0);
      fence_after_sync();
      const uint32_t t_src = t_lane + buf * kDCols;
      if (!whole && !fixer) {
#pragma unroll 1
        for (int h = 0; h < kSNT; ++h) {
          float* dst = my_slot + h * kTileN + qd * 32 + lane;
#pragma unroll 1
          for (int c = 0; c < BM / 16; ++c) {
            if (c * 16 >= args.m) break;
            uint32_t rr[16];
            tmem_ld16(t_src + h * BM + c * 16, rr);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c * 16 + j < args.m) dst[(c * 16 + j) * kSCtaN] = __uint_as_float(rr[j]);
          }
        }
        // the barrier orders the 128 threads' stores before thread 0's release (cumulative)
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) red_release_add(&args.locks[tile], 1);
      } else {
        // chunks of 16 tokens x this thread's column, flattened over (sub-tile, token chunk); the
        // fixer prefetches the first other contributor's piece of chunk i + 1 while it sums chunk i
        const int me = (int)blockIdx.x;
        const int n_chunks = kSNT * ((min(args.m, BM) + 15) / 16);
        const int per_h = n_chunks / kSNT;
        const float* slot1 = args.partial + (int64_t)(me + 1) * kSlotFloats;
        auto chunk_src = [&](int i) {
          return slot1 + (i % per_h) * 16 * kSCtaN + (i / per_h) * kTileN + qd * 32 + lane;
        };
        auto load16 = [&](const float* sp, int i, float (&dstv)[16]) {
          const int rows = min(args.m - (i % per_h) * 16, 16);
          const bool ok = tile * kSCtaN + (i / per_h) * kTileN + qd * 32 + lane < args.n;
#pragma unroll
          for (int j = 0; j < 16; ++j) dstv[j] = (ok && j < rows) ? __ldcg(sp + j * kSCtaN) : 0.f;
        };
        float nxt[16];
        if (fixer) load16(chunk_src(0), 0, nxt);
#pragma unroll 1
        for (int i = 0; i < n_chunks; ++i) {
          const int h = i / per_h, c = i % per_h;
          const int n = tile * kSCtaN + h * kTileN + qd * 32 + lane;
          uint32_t own[16];
          tmem_ld16(t_src + h * BM + c * 16, own);
          float cur[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) cur[j] = nxt[j];
          if (fixer && i + 1 < n_chunks) load16(chunk_src(i + 1), i + 1, nxt);
          tmem_ld_wait();
          float acc[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = __uint_as_float(own[j]);
          if (fixer) {
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] += cur[j];
            for (int cc = me + 2; cc <= c_hi; ++cc) {  // further contributors, fixed order
              float v[16];
              load16(chunk_src(i) + (int64_t)(cc - me - 1) * kSlotFloats, i, v);
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] += v[j];
            }
          }
          if (n < args.n) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int mm = c * 16 + j;
              if (mm < args.m) out[(int64_t)mm * args.n + n] = DT<T>::from_f(acc[j]);
            }
          }
        }
        if (fixer && et == 0) args.locks[tile] = 0;
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[buf]);
      ++piece;
      done += cnt;
      kb = 0;
      ++tile;
    }
    SPROF(1);
    if (et == 0) {
      SPROF_FLUSH(24);
      SPROF_MARK(30);
    }
  }
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
  if (threadIdx.x == 0) SPROF_MARK(31);
}

inline int64_t stream_ws_bytes(int m) {
  const int bm = m <= 16 ? 16 : m <= 32 ? 32 : 64;
  return (int64_t)num_sms() * bm * kSCtaN * 4;  // one fp32 slot per CTA (a range publishes at most one piece)
}

template <int BITS, int BM, typename T>
int launch_stream(const b200_gptq_marlin_gemm_params& p, cudaStream_t st) {
  using P = PlanS<BITS, BM>;
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
  rc = make_weight_maps(p, p.num_groups > 1, kSCtaN, &tm_w, &tm_s, &a);
  if (rc != B200_OK) return rc;
  StreamArgs sa{};
  sa.n_tiles = (p.size_n + kSCtaN - 1) / kSCtaN;
  sa.units = sa.n_tiles * a.num_kb;
  sa.grid = (int)(sa.units < num_sms() ? sa.units : num_sms());
  a.partial = reinterpret_cast<float*>(p.splitk_ws);
  a.locks = p.workspace;
  auto kern = p.num_groups > 1 ? marlin_gemm_stream_kernel<BITS, BM, T, true>
                               : marlin_gemm_stream_kernel<BITS, BM, T, false>;
  static size_t smem_configured[2][kMaxDevices] = {};
  if (ensure_dyn_smem(kern, P::kBytes, smem_configured[p.num_groups > 1 ? 1 : 0]) != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_gemm: cannot reserve %d B of shared memory", (int)P::kBytes);
    return B200_ECUDA;
  }
  cudaError_t le = launch_pdl(kern, dim3(sa.grid), dim3(kSThreads), (size_t)P::kBytes, st, tm_x, tm_w, tm_s, a, sa);
  if (le != cudaSuccess) {
    cudaGetLastError();
    set_error("gptq_marlin_gemm: launch failed: %s", cudaGetErrorString(le));
    return B200_ECUDA;
  }
  return B200_OK;
}
