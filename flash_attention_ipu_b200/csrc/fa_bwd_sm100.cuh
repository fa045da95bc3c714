// fa_bwd_sm100.cuh -- backward attention kernel for B200 (sm_100a).
//
// Replaces the backward graph of the reference, flashAttentionQKVPackedGrad
// (flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:348-512).  Same loop order as the
// reference -- KV tile outer / stationary, Q tiles inner (:410-497) -- and the same math:
//   P = exp(S - LSE) (:457-459), dV += P^T dO (:462), dP = dO V^T (:465),
//   dS = P o (dP - D) (:466-467), dQ += dS K (:470), dK += dS^T Q (:472),
// but O/LSE come from the forward instead of a recompute (:357-359) and D = rowsum(O o dO)
// (:374-377) from a small pre-kernel.  Everything is computed TRANSPOSED (keys on the TMEM
// lanes) so that P^T and dS^T are directly the A operands of the dV and dK GEMMs:
//
//   S^T  = K_j Q_i^T      SS   A = K (K-major)      B = Q_i  (K-major)      -> TMEM S
//   dP^T = V_j dO_i^T     SS   A = V (K-major)      B = dO_i (K-major)      -> TMEM dP
//   dV  += P^T dO_i       TS   A = P^T (TMEM, bf16) B = dO_i (MN-major)     -> TMEM dV
//   dQ_i = dS K_j         SS   A = dS^T smem (MN-major) B = K (MN-major)    -> TMEM dQ (aliases dP)
//   dK  += dS^T Q_i       SS   A = dS^T smem (K-major)  B = Q_i (MN-major)  -> TMEM dK
//
// TMEM (512 cols, D=128): dK [0,128) dV [128,256) S/P^T [256,384) dP^T/dQ [384,512).
// Issue order per Q tile:  dP(i) dV(i) S(i+1) dQ(i) dK(i)  -- the exp of tile i+1 and the drain of
// dQ(i) both happen in the shadow of other GEMMs.  dQ tiles are accumulated across KV tiles in
// an fp32 scratch by TMA reduce-add and cast afterwards.
//
//   warps 0-7  : compute (2 warpgroups; thread <-> key row; WG w owns query columns [64w,64w+64))
//   warps 8-11 : dQ drain (TMEM -> regs -> smem -> TMA reduce-add)
//   warp 12 TMA producer | warp 13 MMA issuer | warp 14 TMEM alloc | warp 15 idle
//
// Build switches (all default off; the default build is the measured v6/v7 kernel.  DESIGN.md 4.6):
//   FA_BWD_DK_TS         dS^T also in TMEM, dK as a TS product           correct, +4.9 % cycles
//   FA_BWD_DRAIN_SINGLE  one 4 KB staging slab per drain warp            correct, +6.9 %
//   FA_BWD_STATS_MMA     -lse / -delta added by the tensor core          correct, no gain (implies DRAIN_SINGLE)
//   FA_BWD_DRAIN_PIPE    one slab per bulk group, wait_group.read 1      parity cases green, not yet timed
//   FA_BWD_DRAIN_REGS    all dQ columns in registers, TMEM released first  compiled only (written without a GPU)
//   FA_EXPERIMENT_*      timing experiments that remove work (wrong results), FA_BWD_TRACE: wait timers
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "fa_fwd_sm100.cuh"

namespace fa {

constexpr int kBwdThreads = 512;
#ifndef FA_BWD_POLY_PAIRS
#define FA_BWD_POLY_PAIRS 12
#endif
constexpr int kBwdPolyPairs = FA_BWD_POLY_PAIRS;
// FA_BWD_STATS_MMA: the per-query statistics (-lse, -delta) are added to S^T and dP^T by the TENSOR
// CORE as one extra K=16 step each (A = a constant [1 1 1 0 ...] slab, B = a [128 q x 16] slab holding
// the statistic split into three 16-bit terms, both in the no-swizzle K-major layout) instead of
// being read by every compute thread with broadcast LDS.128: those loads are 1016 of the 4270
// shared-memory wavefronts per Q tile that bound the kernel (profiles/r01_v6_ncu_bwd.txt); the
// extra MMA steps cost 128.  The slabs take 16 KB, which the dQ staging ring gives up (one 4 KB
// slab per drain warp instead of two).
#ifndef FA_BWD_STATS_MMA
#define FA_BWD_STATS_MMA 0
#endif
constexpr bool kStatsMma = FA_BWD_STATS_MMA != 0;
#ifndef FA_BWD_DRAIN_SINGLE
#define FA_BWD_DRAIN_SINGLE FA_BWD_STATS_MMA
#endif
constexpr bool kDrainSingle = FA_BWD_DRAIN_SINGLE != 0;
// FA_BWD_DRAIN_PIPE: the two 4 KB staging slabs of a drain warp are re-filled one at a time, each
// reduce-add in its own bulk group and `wait_group.read 1` before a slab is written again, so the
// reduce engine always has the other slab queued while this one is staged.  (The default stages
// two slabs, issues two reduces and then waits for BOTH to be read: the trace build shows that
// wait at 1 300 clk per iteration and the drain warp busy 93 % of the time -- it, not the tensor
// pipe, paces the loop.)
#ifndef FA_BWD_DRAIN_PIPE
#define FA_BWD_DRAIN_PIPE 0
#endif
constexpr bool kDrainPipe = FA_BWD_DRAIN_PIPE != 0;
// FA_BWD_DRAIN_REGS: a drain warp reads ALL of its dQ columns into registers at once and releases
// TMEM (dq_empty) before it touches the staging slabs, so dP^T of the next Q tile never waits for
// the reduce engine.  Costs 152 registers per drain thread instead of 104; the compute warps give
// them up (168 -> 144), which they can afford once the row statistics no longer live in registers
// (FA_BWD_STATS_MMA).  Untimed: written after the round's GPU budget ended.
#ifndef FA_BWD_DRAIN_REGS
#define FA_BWD_DRAIN_REGS 0
#endif
constexpr bool kDrainRegs = FA_BWD_DRAIN_REGS != 0;
#ifndef FA_BWD_RED_COLS
#define FA_BWD_RED_COLS 0
#endif
constexpr int kBwdRegCompute = kDrainRegs ? 144 : 168, kBwdRegDrain = kDrainRegs ? 152 : 104;  // one 4 KB staging slab per drain warp instead of two  // of a thread's 32 pairs of exponentials, how many run on the FMA pipe

struct BwdParams {
  const float *lse2;   // [G][Lq_pad]  -lse * log2(e); -inf for padding / rows without keys
  const float *delta;  // [G][Lq_pad]  -rowsum(O o dO); 0 for padding
  // FA_BWD_STATS_MMA: the same two statistics as ready-made shared-memory images, one 4 KB slab per
  // 128-row Q tile ([G][Lq_pad / 128][4096] bytes each; see fa_bwd_prep_kernel)
  const uint8_t *slab_lse, *slab_delta;
  float *dq_acc;       // [G][Lq][D] fp32 dQ accumulator (FA_BWD_DRAIN_RED addresses it directly)
  int Lq, Lk, H, Lq_pad;
  int Hkv, kv_ratio;   // K/V heads and query heads per K/V head (H = Hkv * kv_ratio; 1 = multi-head)
  int causal, shift;
  float scale, scale_log2;
  int num_kv_tiles;    // ceil(Lk / 128)
#ifdef FA_BWD_TRACE
  unsigned long long *trace;  // CTA 0: cycles the MMA warp spent in each of its waits
#endif
};

template <int D>
struct BwdSmem {
  static constexpr int kTileBytes = 128 * D * 2;
  static constexpr int kHalfBytes = 128 * 128;
  static constexpr int kKOff = 0;
  static constexpr int kVOff = kKOff + kTileBytes;
  static constexpr int kQOff = kVOff + kTileBytes;            // 2 stages (epilogue: dK | dV staging)
  static constexpr int kDOOff = kQOff + 2 * kTileBytes;       // 1 stage
  static constexpr int kDSOff = kDOOff + kTileBytes;          // dS^T [128 keys][128 q] bf16
  static constexpr int kDQOff = kDSOff + 2 * kHalfBytes;      // fp32 staging: 4 drain warps x kDrainSlabs x 4 KB
  // FA_BWD_SLAB2K (with FA_BWD_DRAIN_REGS): the staging ring of a drain warp is two slabs of 32 rows x
  // 16 fp32 columns (2 KB, 64-byte rows, SWIZZLE_64B) instead of 32 columns -- 16 KB for the four
  // warps, which leaves room for the statistic operand slabs of FA_BWD_STATS_MMA
#ifdef FA_BWD_SLAB2K
  static constexpr int kSlabCols = 16, kDrainSlabs = 2;
#else
  static constexpr int kSlabCols = 32, kDrainSlabs = kDrainSingle ? 1 : 2;
#endif
  static constexpr int kSlabBytes = 32 * kSlabCols * 4;
  static constexpr int kStatOff = kDQOff + 4 * kDrainSlabs * kSlabBytes;
  // !kStatsMma: 2 stages x (lse2[128] | delta[128]) fp32.   kStatsMma: four 4 KB no-swizzle K-major
  // [128 x 16] operand slabs: lse (stage 0, 1; travel with Q), delta (travels with dO), ones
  static constexpr int kSlabLseOff = kStatOff, kSlabDeltaOff = kStatOff + 2 * 4096, kSlabOnesOff = kStatOff + 3 * 4096;
  static constexpr int kBarOff = kStatOff + (kStatsMma ? 4 * 4096 : 2 * 1024);
  static constexpr int kNumBars = 16;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kBytes = kTmemPtrOff + 16;
};

// first Q tile that sees any key of KV tile j
__device__ __forceinline__ int bwd_first_q_tile(const BwdParams &p, int j) {
  if (!p.causal) return 0;
  const int a = j * 128 - p.shift;
  return a <= 0 ? 0 : a / 128;
}

#ifdef FA_BWD_TRACE
#define FA_BT_WAIT(slot, ...) do { const long long t0__ = clock64(); __VA_ARGS__; tr[slot] += clock64() - t0__; } while (0)
#else
#define FA_BT_WAIT(slot, ...) do { __VA_ARGS__; } while (0)
#endif

template <bool kGqa>
__device__ __forceinline__ void bwd_next_tile(int &h, int &i, int i0, int num_q_tiles) {
  ++i;
  if (kGqa && i == num_q_tiles) { i = i0; ++h; }
}

// kGqa = several query heads share the CTA's K/V head (the Q-tile loop then walks head by head);
// the multi-head instantiation keeps the plain induction variable i = i0 + it, which the compiler
// holds in uniform registers (the loop-carried (h, i) pair cost 6 % of the backward).
template <int D, bool kBf16, bool kGqa>
__global__ void __launch_bounds__(kBwdThreads, 1)
fa_bwd_sm100_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                    const __grid_constant__ CUtensorMap tmDQ, const __grid_constant__ CUtensorMap tmDK,
                    const __grid_constant__ CUtensorMap tmDV, const BwdParams p) {
  using L = BwdSmem<D>;
  constexpr int kHalves = D / 64;
  constexpr uint32_t kColDK = 0, kColDV = D, kColS = 2 * D, kColDP = 2 * D + 128;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kBarOff);
  uint64_t *kv_full = bars + 0;
  uint64_t *q_full = bars + 1;    // [2]
  uint64_t *q_empty = bars + 3;   // [2]
  uint64_t *do_full = bars + 5;
  uint64_t *do_empty = bars + 6;
  uint64_t *s_full = bars + 7;    // MMA -> compute : S^T(i) ready
  uint64_t *p_full = bars + 8;    // compute -> MMA : P^T(i) in TMEM        (8 warp arrivals)
  uint64_t *dp_full = bars + 9;   // MMA -> compute : dP^T(i) ready
  uint64_t *ds_full = bars + 10;  // compute -> MMA : dS^T(i) in smem       (8 warp arrivals)
  uint64_t *dq_full = bars + 11;  // MMA -> drain   : dQ(i) ready
  uint64_t *dq_empty = bars + 12; // drain -> MMA   : dQ(i) left TMEM       (4 warp arrivals)
  uint64_t *dkv_full = bars + 13; // MMA -> compute : final dK, dV ready
  uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(smem + L::kTmemPtrOff);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
#ifdef FA_BWD_TRACE
  const long long t_entry = clock64();
#endif
  const int gk = blockIdx.x / p.num_kv_tiles;  // (batch, K/V head)
  const int j = blockIdx.x % p.num_kv_tiles;   // ascending: the KV tiles with most Q tiles first
  const int b = gk / p.Hkv, hk = gk % p.Hkv;
  const int num_q_tiles = (p.Lq + 127) / 128;
  const int i0 = bwd_first_q_tile(p, j);
  // The CTA walks the Q tiles of every query head that shares this K/V head (grouped-query
  // attention: kv_ratio heads; dK, dV simply keep accumulating in TMEM across them).
  const int n1 = max(num_q_tiles - i0, 0);     // Q tiles per query head
  const int n = kGqa ? n1 * p.kv_ratio : n1;   // loop iterations: heads outer, Q tiles i0.. inner
  const int h0 = kGqa ? hk * p.kv_ratio : hk;

  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) {
    printf("fa_b200: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  if (warp == 12 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ); tma_prefetch_desc(&tmDK); tma_prefetch_desc(&tmDV);
  }
  if (warp == 13 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
    mbar_init(do_full, 1); mbar_init(do_empty, 1);
    mbar_init(s_full, 1); mbar_init(p_full, 8); mbar_init(dp_full, 1); mbar_init(ds_full, 8);  // per-warp arrivals
    mbar_init(dq_full, 1); mbar_init(dq_empty, 4); mbar_init(dkv_full, 1);
    fence_barrier_init();
  }
  if (kStatsMma && threadIdx.x < 128) {
    // constant A operand of the two statistic steps: row r = [1 1 1 0 0 0 0 0 | 0 x 8] (16-bit floats)
    constexpr uint32_t kOne = kBf16 ? 0x3F80u : 0x3C00u;
    uint8_t *row = smem + L::kSlabOnesOff + (threadIdx.x >> 3) * 256 + (threadIdx.x & 7) * 16;
    *reinterpret_cast<uint4 *>(row) = make_uint4(kOne | (kOne << 16), kOne, 0u, 0u);
    *reinterpret_cast<uint4 *>(row + 128) = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  if (warp == 14) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 12) {
    // ============================== TMA producer ==============================
    reg_dealloc<72>();
    if (lane == 0 && n > 0) {
      mbar_arrive_expect_tx(kv_full, 2 * L::kTileBytes);
      for (int hh = 0; hh < kHalves; ++hh) {
        tma_load_4d(&tmK, kv_full, smem + L::kKOff + hh * L::kHalfBytes, hh * 64, j * 128, hk, b);
        tma_load_4d(&tmV, kv_full, smem + L::kVOff + hh * L::kHalfBytes, hh * 64, j * 128, hk, b);
      }
      for (int it = 0, h = h0, i = i0; it < n; ++it, bwd_next_tile<kGqa>(h, i, i0, num_q_tiles)) {
        const int s = it & 1;
        const float *lse_g = p.lse2 + static_cast<size_t>(b * p.H + h) * p.Lq_pad;
        const float *delta_g = p.delta + static_cast<size_t>(b * p.H + h) * p.Lq_pad;
        mbar_wait(&q_empty[s], ((it >> 1) & 1) ^ 1);
#ifdef FA_EXPERIMENT_HALF_LOADS  // timing experiment only (wrong results): skip every other tile load
        if ((it & 3) >= 2) {
          mbar_arrive(&q_full[s]);
          mbar_wait(do_empty, (it & 1) ^ 1);
          mbar_arrive(do_full);
          continue;
        }
#endif
        const size_t slab = (static_cast<size_t>(b * p.H + h) * (p.Lq_pad >> 7) + i) * 4096;
        mbar_arrive_expect_tx(&q_full[s], L::kTileBytes + (kStatsMma ? 4096 : 1024));
        for (int hh = 0; hh < kHalves; ++hh)
          tma_load_4d(&tmQ, &q_full[s], smem + L::kQOff + s * L::kTileBytes + hh * L::kHalfBytes,
                      hh * 64, i * 128, h, b);
        if (kStatsMma) {
          bulk_load_1d(smem + L::kSlabLseOff + s * 4096, p.slab_lse + slab, 4096, &q_full[s]);
        } else {
          bulk_load_1d(smem + L::kStatOff + s * 1024, lse_g + i * 128, 512, &q_full[s]);
          bulk_load_1d(smem + L::kStatOff + s * 1024 + 512, delta_g + i * 128, 512, &q_full[s]);
        }
        mbar_wait(do_empty, (it & 1) ^ 1);
        mbar_arrive_expect_tx(do_full, L::kTileBytes + (kStatsMma ? 4096 : 0));
        for (int hh = 0; hh < kHalves; ++hh)
          tma_load_4d(&tmDO, do_full, smem + L::kDOOff + hh * L::kHalfBytes, hh * 64, i * 128, h, b);
        if (kStatsMma) bulk_load_1d(smem + L::kSlabDeltaOff, p.slab_delta + slab, 4096, do_full);
      }
    }
  } else if (warp == 13) {
    // ============================== MMA issuer ==============================
    // Warp-uniform loop, one elected lane issues; (lo, hi) descriptors with constant offsets.
    reg_dealloc<72>();
    if (n > 0) {
      constexpr int kFmt = kBf16 ? 1 : 0;
      constexpr uint32_t idesc_kk = make_idesc(128, 128, kFmt, 0, 0);  // S^T, dP^T
      constexpr uint32_t idesc_dv = make_idesc(128, D, kFmt, 0, 1);    // dV (A in TMEM), dK (A K-major)
      constexpr uint32_t idesc_dq = make_idesc(128, D, kFmt, 1, 1);    // dQ (A and B MN-major)
      constexpr uint32_t hi = smem_desc_hi(1024);
      constexpr uint32_t kTile16 = L::kTileBytes >> 4, kHalf16 = L::kHalfBytes >> 4;
      // K-major views (LBO unused) and MN-major views (LBO = distance between 64-column slabs)
      const uint32_t k_kmaj = smem_desc_lo(smem_u32(smem + L::kKOff), 16);
      const uint32_t v_kmaj = smem_desc_lo(smem_u32(smem + L::kVOff), 16);
      const uint32_t q_kmaj = smem_desc_lo(smem_u32(smem + L::kQOff), 16);
      const uint32_t do_kmaj = smem_desc_lo(smem_u32(smem + L::kDOOff), 16);
      const uint32_t ds_kmaj = smem_desc_lo(smem_u32(smem + L::kDSOff), 16);
      const uint32_t k_mn = smem_desc_lo(smem_u32(smem + L::kKOff), L::kHalfBytes);
      const uint32_t q_mn = smem_desc_lo(smem_u32(smem + L::kQOff), L::kHalfBytes);
      const uint32_t do_mn = smem_desc_lo(smem_u32(smem + L::kDOOff), L::kHalfBytes);
      const uint32_t ds_mn = smem_desc_lo(smem_u32(smem + L::kDSOff), L::kHalfBytes);
      const uint32_t tS = tmem_base + kColS, tDP = tmem_base + kColDP;
      const uint32_t tDK = tmem_base + kColDK, tDV = tmem_base + kColDV;
      const bool leader = elect_one();

      // no-swizzle K-major [128 x 16] slabs: LBO = 128 (next 8 K elements), SBO = 256 (next 8 rows)
      // (tests/cuda/umma_probe.cu mode 5)
      constexpr uint32_t hi_ns = ((256u >> 4) & 0x3FFFu) | (1u << 14);
      const uint32_t ones_ns = smem_desc_lo(smem_u32(smem + L::kSlabOnesOff), 128);
      const uint32_t lse_ns = smem_desc_lo(smem_u32(smem + L::kSlabLseOff), 128);
      const uint32_t delta_ns = smem_desc_lo(smem_u32(smem + L::kSlabDeltaOff), 128);
      // d = a b^T (+ 1 stat^T: the per-query statistic added to every key row by one more K step)
      auto issue_kk = [&](uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t stat_lo) {
        #pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = ((k / 4) * L::kHalfBytes + (k % 4) * 32) >> 4;
          umma_ss2(d_tmem, a_lo + off, hi, b_lo + off, hi, idesc_kk, k > 0);
        }
        if (kStatsMma) umma_ss2(d_tmem, ones_ns, hi_ns, stat_lo, hi_ns, idesc_kk, 1);
      };

      mbar_wait(kv_full, 0);
      mbar_wait(&q_full[0], 0);
      tc_fence_after();
      if (leader) {
        issue_kk(tS, k_kmaj, q_kmaj, lse_ns);  // S^T(0)
        umma_commit(s_full);
      }
#ifdef FA_BWD_TRACE
      long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      const long long t_loop = clock64();
#endif
#ifdef FA_BWD_DK_TS
      // Variant: dS^T is ALSO written to TMEM (packed, over the dP^T columns its warpgroup has
      // already read) and dK += dS^T Q takes it from there (TS form: no shared-memory A reads).
      // dQ aliases dP^T, so dK must be issued before dQ; the order becomes
      //   dV(it) dP(it) S(it+1) dK(it) dQ(it)
      // and dV(it+1) is what runs while the drain empties dQ(it).
      for (int it = 0; it < n; ++it) {
        const uint32_t q_stage = (it & 1) * kTile16;
#ifdef FA_BWD_ORDER2  // dP(it) dV(it) S(it+1) dK(it) dQ(it): dS^T gets dV + S as cover, dP(it+1) waits for the (short) drain
        FA_BT_WAIT(0, mbar_wait(do_full, it & 1));
        if (it > 0) FA_BT_WAIT(1, mbar_wait(dq_empty, (it - 1) & 1));
        tc_fence_after();
        if (leader) {
          issue_kk(tDP, v_kmaj, do_kmaj, delta_ns);
          umma_commit(dp_full);
        }
        FA_BT_WAIT(2, mbar_wait(p_full, it & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ts2(tDV, tS + (k < 4 ? 8 * k : 64 + 8 * (k - 4)), do_mn + k * (2048 >> 4), hi, idesc_dv,
                     (it > 0 || k > 0));
          umma_commit(do_empty);
        }
#else
        FA_BT_WAIT(2, mbar_wait(p_full, it & 1));
        FA_BT_WAIT(0, mbar_wait(do_full, it & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ts2(tDV, tS + (k < 4 ? 8 * k : 64 + 8 * (k - 4)), do_mn + k * (2048 >> 4), hi, idesc_dv,
                     (it > 0 || k > 0));
        }
        if (it > 0) FA_BT_WAIT(1, mbar_wait(dq_empty, (it - 1) & 1));
        tc_fence_after();
        if (leader) {
          issue_kk(tDP, v_kmaj, do_kmaj, delta_ns);
          umma_commit(dp_full);
          umma_commit(do_empty);
        }
#endif
        if (it + 1 < n) {
          FA_BT_WAIT(3, mbar_wait(&q_full[(it + 1) & 1], ((it + 1) >> 1) & 1));
          tc_fence_after();
          if (leader) {
            issue_kk(tS, k_kmaj, q_kmaj + ((it + 1) & 1) * kTile16, lse_ns + ((it + 1) & 1) * (4096 >> 4));
            umma_commit(s_full);
          }
        }
        FA_BT_WAIT(4, mbar_wait(ds_full, it & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ts2(tDK, tDP + (k < 4 ? 8 * k : 64 + 8 * (k - 4)), q_mn + q_stage + k * (2048 >> 4), hi,
                     idesc_dv, (it > 0 || k > 0));
          umma_commit(&q_empty[it & 1]);
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ss2(tDP, ds_mn + k * (2048 >> 4), hi, k_mn + k * (2048 >> 4), hi, idesc_dq, k > 0);
          umma_commit(dq_full);
        }
        __syncwarp();
      }
      (void)ds_kmaj;
#else
      for (int it = 0; it < n; ++it) {
        const uint32_t q_stage = (it & 1) * kTile16;
        // dP^T(it) = V dO^T   (overwrites dQ(it-1): the drain must have emptied it)
        FA_BT_WAIT(0, mbar_wait(do_full, it & 1));
        if (it > 0) FA_BT_WAIT(1, mbar_wait(dq_empty, (it - 1) & 1));
        tc_fence_after();
        if (leader) {
          issue_kk(tDP, v_kmaj, do_kmaj, delta_ns);
          umma_commit(dp_full);
        }
        // dV += P^T dO
        FA_BT_WAIT(2, mbar_wait(p_full, it & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ts2(tDV, tS + (k < 4 ? 8 * k : 64 + 8 * (k - 4)), do_mn + k * (2048 >> 4), hi, idesc_dv,
                     (it > 0 || k > 0));
          umma_commit(do_empty);
        }
        // S^T(it+1) = K Q^T  (overwrites P^T(it), which dV(it) has consumed by then: in-order pipe)
        if (it + 1 < n) {
          FA_BT_WAIT(3, mbar_wait(&q_full[(it + 1) & 1], ((it + 1) >> 1) & 1));
          tc_fence_after();
          if (leader) {
            issue_kk(tS, k_kmaj, q_kmaj + ((it + 1) & 1) * kTile16, lse_ns + ((it + 1) & 1) * (4096 >> 4));
            umma_commit(s_full);
          }
        }
        // dQ(it) = dS K ; dK += dS^T Q
        FA_BT_WAIT(4, mbar_wait(ds_full, it & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ss2(tDP, ds_mn + k * (2048 >> 4), hi, k_mn + k * (2048 >> 4), hi, idesc_dq, k > 0);
          umma_commit(dq_full);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t off = ((k / 4) * L::kHalfBytes + (k % 4) * 32) >> 4;
            umma_ss2(tDK, ds_kmaj + off, hi, q_mn + q_stage + k * (2048 >> 4), hi, idesc_dv,
                     (it > 0 || k > 0));
          }
          umma_commit(&q_empty[it & 1]);
        }
        __syncwarp();
      }
#endif
      if (leader) umma_commit(dkv_full);
#ifdef FA_BWD_TRACE
      if (blockIdx.x == 0 && lane == 0) {
        for (int t = 0; t < 5; ++t) p.trace[t] = tr[t];
        p.trace[5] = clock64() - t_loop;
        p.trace[6] = n;
        p.trace[12] = t_loop - t_entry;  // CTA entry -> first S^T issued (prologue)
        p.trace[13] = clock64();         // loop end (absolute; the CTA-exit stamp subtracts it)
      }
#endif
      (void)kHalf16;
    }
  } else if (warp >= 8 && warp < 12) {
    // ============================== dQ drain ==============================
    // Each warp drains its own 32 query rows independently: TMEM -> registers -> a private
    // 2 x 4 KB staging ring -> TMA reduce-add (32-row boxes).  No cross-warp barriers, and one
    // reduce is in flight while the next slab is being staged.
    if (kDrainRegs) reg_alloc<kBwdRegDrain>(); else reg_dealloc<kBwdRegDrain>();
    const int w = warp & 3;
    const uint32_t tDQ = tmem_base + kColDP + (static_cast<uint32_t>(w * 32) << 16);
    uint8_t *wstage = smem + L::kDQOff + w * (L::kDrainSlabs * L::kSlabBytes);
    for (int it = 0, h = h0, i = i0; it < n; ++it, bwd_next_tile<kGqa>(h, i, i0, num_q_tiles)) {
#ifdef FA_BWD_TRACE
      const long long tf0 = clock64();
#endif
      mbar_wait(dq_full, it & 1);
#ifdef FA_BWD_TRACE
      if (blockIdx.x == 0 && w == 0 && lane == 0) p.trace[10] += clock64() - tf0;  // drain warp idle: waiting for dQ
#endif
      tc_fence_after();
      if (kDrainSingle && !kDrainRegs) {
        // One 4 KB slab (32 fp32 columns) per warp: a round = wait until the previous reduce has
        // read the slab, store 32 columns, reduce.  Two rounds of columns are held in registers, so
        // dQ has left TMEM (dq_empty) once the slabs of rounds 0 .. D/32-3 are on their way.
        constexpr int kRounds = D / 32;
        uint32_t r[2][32];
        tmem_ld32(tDQ, r[0]);
        tmem_ld32(tDQ + 32, r[1]);
        #pragma unroll
        for (int q = 0; q < kRounds; ++q) {
          if (q + 2 <= kRounds) {  // a load is outstanding: initial pair, or the one issued last round
            tmem_ld_wait();
            if (q + 2 == kRounds) { tc_fence_before(); mbar_arrive_warp(dq_empty); }
          }
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
#ifndef FA_EXPERIMENT_NO_DRAIN_STS
          #pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v4 = make_uint4(r[q & 1][c * 4], r[q & 1][c * 4 + 1], r[q & 1][c * 4 + 2], r[q & 1][c * 4 + 3]);
            *reinterpret_cast<uint4 *>(wstage + lane * 128 + ((c ^ (lane & 7)) * 16)) = v4;
          }
#else
          asm volatile("" ::"r"(r[q & 1][0]), "r"(r[q & 1][31]));
#endif
          if (q + 2 < kRounds) tmem_ld32(tDQ + (q + 2) * 32, r[q & 1]);
          fence_proxy_async_smem();
          __syncwarp();
#ifndef FA_EXPERIMENT_NO_DRAIN_STS
          if (lane == 0) {
            tma_reduce_add_4d(&tmDQ, wstage, q * 32, i * 128 + w * 32, h, b);
            tma_store_commit();
          }
#endif
        }
        continue;
      }
      if (kDrainRegs) {
        constexpr int kSlabs = D / 32;
#ifdef FA_BWD_TRACE
        const bool tr_on = blockIdx.x == 0 && w == 0 && lane == 0;
        const long long tq0 = clock64();
#endif
#ifdef FA_BWD_DRAIN_RED
        // dQ leaves the SM as red.global.add.v2.f32 straight from registers: the 16x256b TMEM load
        // shape gives lanes 4r .. 4r+3 eight consecutive fp32 of row r (and of row r + 8), so one
        // warp instruction adds 8 whole 32-byte sectors.  No staging, no TMA.
        uint32_t r[2][64];
        tmem_ld_16x256b_x16(tDQ, r[0]);
        tmem_ld_16x256b_x16(tDQ + (16u << 16), r[1]);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive_warp(dq_empty);
#ifdef FA_BWD_TRACE
        if (tr_on) p.trace[11] += clock64() - tq0;
#endif
        {
          const int q_row0 = i * 128 + w * 32;
          float *base = p.dq_acc + (static_cast<size_t>(b * p.H + h) * p.Lq + q_row0) * D + 2 * (lane & 3);
          #pragma unroll
          for (int hf = 0; hf < 2; ++hf)
            #pragma unroll
            for (int sub = 0; sub < 2; ++sub) {
              const int row = hf * 16 + sub * 8 + (lane >> 2);
              if (q_row0 + row < p.Lq) {
                float *rp = base + static_cast<size_t>(row) * D;
                #pragma unroll
                for (int c = 0; c < 16; ++c)
                  red_add_v2(rp + c * 8, r[hf][c * 4 + sub * 2], r[hf][c * 4 + sub * 2 + 1]);
              }
            }
        }
        continue;
#else
        uint32_t r[kSlabs][32];
        #pragma unroll
        for (int k = 0; k < kSlabs; ++k) tmem_ld32(tDQ + k * 32, r[k]);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive_warp(dq_empty);  // dQ has left TMEM: nothing below is on the tensor pipe's path
#ifdef FA_BWD_TRACE
        if (tr_on) p.trace[11] += clock64() - tq0;
#endif
        // FA_BWD_RED_COLS: the last kRedCols columns of the tile leave as red.global.add.v4.f32 straight
        // from the registers (LSU path) while the rest goes through staging + TMA reduce-add: the TMA
        // unit moves reduce-adds at 25.4 B/clk/SM (tests/cuda/smem_contention.cu) and is the unit that
        // paces the loop, so part of the volume takes the other road into L2.
        constexpr int kRedCols = FA_BWD_RED_COLS;
        if (kRedCols > 0) {
          const int qrow = i * 128 + w * 32 + lane;
          if (qrow < p.Lq) {
            float *rp = p.dq_acc + (static_cast<size_t>(b * p.H + h) * p.Lq + qrow) * D + (D - kRedCols);
            #pragma unroll
            for (int c = 0; c < kRedCols; c += 4) {
              const int col = D - kRedCols + c;
              red_add_v4(rp + c, r[col >> 5][col & 31], r[col >> 5][(col & 31) + 1], r[col >> 5][(col & 31) + 2],
                         r[col >> 5][(col & 31) + 3]);
            }
          }
        }
        constexpr int kOps = (D - kRedCols) / L::kSlabCols;  // TMA reduce-adds per warp per Q tile
        constexpr int kChunks = L::kSlabCols / 4;        // 16-byte chunks per staged row
        #pragma unroll
        for (int k = 0; k < kOps; ++k) {
          uint8_t *slab = wstage + (k % L::kDrainSlabs) * L::kSlabBytes;
#ifdef FA_BWD_TRACE
          const long long tw0 = clock64();
#endif
          if (lane == 0) tma_store_wait_read<L::kDrainSlabs - 1>();  // the reduce that last used this slab has read it
#ifdef FA_BWD_TRACE
          if (tr_on) p.trace[8] += clock64() - tw0;
#endif
          __syncwarp();
          #pragma unroll
          for (int c = 0; c < kChunks; ++c) {
            const int col = k * L::kSlabCols + c * 4;
            uint4 v4 = make_uint4(r[col >> 5][col & 31], r[col >> 5][(col & 31) + 1], r[col >> 5][(col & 31) + 2],
                                  r[col >> 5][(col & 31) + 3]);
            // SWIZZLE_128B: chunk ^= row % 8 (128-byte rows); SWIZZLE_64B: chunk ^= (row / 2) % 4 (64-byte rows)
            const int pc = kChunks == 8 ? (c ^ (lane & 7)) : (c ^ ((lane >> 1) & 3));
            *reinterpret_cast<uint4 *>(slab + lane * (L::kSlabCols * 4) + pc * 16) = v4;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
#if defined(FA_EXPERIMENT_DRAIN_STORE)     // timing experiment only (wrong results): plain TMA store, no L2 reduction
            tma_store_4d(&tmDQ, slab, k * L::kSlabCols, i * 128 + w * 32, h, b);
#elif defined(FA_EXPERIMENT_DRAIN_SAMETILE)  // timing experiment only (wrong results): every Q tile reduces into tile i0
            tma_reduce_add_4d(&tmDQ, slab, k * L::kSlabCols, i0 * 128 + w * 32, h, b);
#else
            tma_reduce_add_4d(&tmDQ, slab, k * L::kSlabCols, i * 128 + w * 32, h, b);
#endif
            tma_store_commit();
          }
        }
        continue;
#endif
      }
      if (kDrainPipe) {
        constexpr int kSlabs = D / 32;
        uint32_t r[2][32];
        tmem_ld32(tDQ, r[0]);
        tmem_ld32(tDQ + 32, r[1]);
        #pragma unroll
        for (int k = 0; k < kSlabs; ++k) {
          if ((k & 1) == 0) {  // the loads of slabs k and k+1 were issued together
            tmem_ld_wait();
            if (k + 2 == kSlabs) { tc_fence_before(); mbar_arrive_warp(dq_empty); }
          }
          // this buffer was last used two bulk groups ago; the previous group (other buffer) may still be pending
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          uint8_t *slab = wstage + (k & 1) * 4096;
          #pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v4 = make_uint4(r[k & 1][c * 4], r[k & 1][c * 4 + 1], r[k & 1][c * 4 + 2], r[k & 1][c * 4 + 3]);
            *reinterpret_cast<uint4 *>(slab + lane * 128 + ((c ^ (lane & 7)) * 16)) = v4;
          }
          if (k + 2 < kSlabs) tmem_ld32(tDQ + (k + 2) * 32, r[k & 1]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_reduce_add_4d(&tmDQ, slab, k * 32, i * 128 + w * 32, h, b);
            tma_store_commit();
          }
        }
        continue;
      }
#ifdef FA_BWD_TRACE
      const bool tr_on = blockIdx.x == 0 && w == 0 && lane == 0;
      long long tq0 = clock64();
#endif
      // 64 fp32 columns (two 32-column slabs, one per staging buffer) at a time.  The TMEM reads of
      // the next half are issued as soon as this half's registers have been stored, so dQ leaves
      // TMEM (dq_empty -> the MMA warp may overwrite it with dP^T(it+1)) after two loads, one
      // round of stores and no wait on the reduce engine.
      uint32_t r[2][32];
      tmem_ld32(tDQ, r[0]);
      tmem_ld32(tDQ + 32, r[1]);
      #pragma unroll
      for (int hh = 0; hh < kHalves; ++hh) {
        tmem_ld_wait();
        if (hh == kHalves - 1) {
          tc_fence_before(); mbar_arrive_warp(dq_empty);
#ifdef FA_BWD_TRACE
          if (tr_on) p.trace[11] += clock64() - tq0;  // dq_full seen -> dq_empty arrived
#endif
        }
#ifdef FA_BWD_TRACE
        const long long tw0 = clock64();
#endif
        if (lane == 0) tma_store_wait_read<0>();  // the reduces that last used the two buffers have read them
#ifdef FA_BWD_TRACE
        if (tr_on) p.trace[8 + hh] += clock64() - tw0;   // [8]: wait_read before half 0, [9]: before half 1
#endif
        __syncwarp();
#ifndef FA_EXPERIMENT_NO_DRAIN_STS  // (timing experiment only when defined: wrong results)
        #pragma unroll
        for (int sl = 0; sl < 2; ++sl) {
          #pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v4 = make_uint4(r[sl][c * 4], r[sl][c * 4 + 1], r[sl][c * 4 + 2], r[sl][c * 4 + 3]);
            *reinterpret_cast<uint4 *>(wstage + sl * 4096 + lane * 128 + ((c ^ (lane & 7)) * 16)) = v4;
          }
        }
#else
        asm volatile("" ::"r"(r[0][0]), "r"(r[1][31]));
#endif
        if (hh + 1 < kHalves) {
          tmem_ld32(tDQ + (hh + 1) * 64, r[0]);
          tmem_ld32(tDQ + (hh + 1) * 64 + 32, r[1]);
        }
        fence_proxy_async_smem();
        __syncwarp();
#if !defined(FA_EXPERIMENT_NO_REDUCE) && !defined(FA_EXPERIMENT_NO_DRAIN_STS)
        if (lane == 0) {
          tma_reduce_add_4d(&tmDQ, wstage, hh * 64, i * 128 + w * 32, h, b);
#ifndef FA_EXPERIMENT_HALF_REDUCE  // (timing experiment only when defined: wrong results)
          tma_reduce_add_4d(&tmDQ, wstage + 4096, hh * 64 + 32, i * 128 + w * 32, h, b);
#endif
          tma_store_commit();
        }
#endif
      }
    }
    // the staging slabs must stay valid until the last reduce has READ them; completion of the
    // writes themselves is guaranteed by the kernel boundary (the convert pass is a later launch)
    if (lane == 0) tma_store_wait_read<0>();
  } else if (warp < 8) {
    // ============================== compute: P^T, dS^T ; epilogue dK, dV ==============================
    reg_alloc<kBwdRegCompute>();
    const int wg = warp >> 2;                      // which 64 query columns of the tile
    const int row = (warp & 3) * 32 + lane;        // key row within the KV tile (= TMEM lane)
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + kColS + lane_addr + 64 * wg;
    const uint32_t tDP = tmem_base + kColDP + lane_addr + 64 * wg;
    uint8_t *ds_row = smem + L::kDSOff + wg * L::kHalfBytes + row * 128;
    const int key = j * 128 + row;

    for (int it = 0, hx = 0, i = i0; it < n; ++it, bwd_next_tile<kGqa>(hx, i, i0, num_q_tiles)) {
      const float4 *stat = reinterpret_cast<const float4 *>(smem + L::kStatOff + (it & 1) * 1024) + wg * 16;
      // -lse2 of this warpgroup's 64 query columns: loaded (broadcast LDS.128) BEFORE the S^T wait so
      // that the shared-memory latency hides behind it.  (kStatsMma: the tensor core has already
      // added -lse / softmax_scale to S^T and -delta to dP^T; nothing to load.)
      float nl[kStatsMma ? 2 : 64];
      if (!kStatsMma) {
        mbar_wait(&q_full[it & 1], (it >> 1) & 1);   // makes the lse2 / delta stage visible to this thread
        #pragma unroll
        for (int c4 = 0; c4 < (kStatsMma ? 0 : 16); ++c4) {
#ifdef FA_EXPERIMENT_NO_STATS  // timing experiment only (wrong results): no broadcast loads of the row statistics
          const float4 l4 = make_float4(-3.f, -3.f, -3.f, -3.f); (void)stat;
#else
          const float4 l4 = stat[c4];
#endif
          nl[4 * c4] = l4.x; nl[4 * c4 + 1] = l4.y; nl[4 * c4 + 2] = l4.z; nl[4 * c4 + 3] = l4.w;
        }
      }
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      float pr[64];
      {
        uint32_t s[2][32];
        tmem_ld32(tS, s[0]);
        tmem_ld32(tS + 32, s[1]);
        tmem_ld_wait();
        const uint64_t c2 = f2_pack(p.scale_log2, p.scale_log2);
        // kBwdPolyPairs of the 32 pairs run as a polynomial on the FMA pipe, the rest on MUFU; the
        // two warpgroups (one warp of each per SM sub-partition) take the two kinds in opposite
        // order so that one keeps the FMA pipe busy while the other occupies MUFU.
        auto exp_all = [&](auto poly_first) {
          #pragma unroll
          for (int c = 0; c < 64; c += 2) {
            float x0, x1;
            if (kStatsMma)
              f2_unpack(f2_mul(f2_pack(s[c >> 5][c & 31], s[c >> 5][(c & 31) + 1]), c2), x0, x1);
            else
              f2_unpack(f2_fma(f2_pack(s[c >> 5][c & 31], s[c >> 5][(c & 31) + 1]), c2,
                               f2_pack(nl[kStatsMma ? 0 : c], nl[kStatsMma ? 1 : c + 1])), x0, x1);
            const int pair = c >> 1;
            if (decltype(poly_first)::value ? pair < kBwdPolyPairs : pair >= 32 - kBwdPolyPairs) {
              poly_exp2_pair(x0, x1, pr[c], pr[c + 1]);
            } else {
              pr[c] = fast_exp2(x0);
              pr[c + 1] = fast_exp2(x1);
            }
          }
        };
        if (wg) exp_all(BoolC<true>{}); else exp_all(BoolC<false>{});
      }
      // causal mask on tiles that cross the diagonal: query column visible iff key <= q + shift
      if (p.causal && j * 128 + 127 > i * 128 + p.shift) {
        const int cmin = key - p.shift - i * 128 - 64 * wg;  // columns below cmin are masked
        #pragma unroll
        for (int c = 0; c < 64; ++c)
          if (c < cmin) pr[c] = 0.f;
      }
      // key rows past the end of K (last KV tile of a ragged Lk): K is zero-filled there, so
      // S = 0 and exp2(-lse2) is unbounded for rows with a strongly negative LSE.  Select, not
      // multiply: the value may be inf / polynomial garbage.  (The forward masks the same keys
      // through `limit = Lk - 1 - col0`.)
      if (j * 128 + 128 > p.Lk && key >= p.Lk) {
        #pragma unroll
        for (int c = 0; c < 64; ++c) pr[c] = 0.f;
      }
      {
        uint32_t pk[32];
        #pragma unroll
        for (int c = 0; c < 32; ++c) pk[c] = pack2<kBf16>(pr[2 * c], pr[2 * c + 1]);
        tmem_st32(tS, pk);  // P^T over the first half of this warpgroup's own S^T columns
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive_warp(p_full);
      }
      // dS^T = P^T o (dP^T - delta); -delta of each 32-column half is fetched ahead of the
      // barrier / TMEM load it is used after
      float nd[kStatsMma ? 2 : 32];
      #pragma unroll
      for (int c4 = 0; c4 < (kStatsMma ? 0 : 8); ++c4) {
#ifdef FA_EXPERIMENT_NO_STATS
        const float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
#else
        const float4 d4 = stat[32 + c4];
#endif
        nd[4 * c4] = d4.x; nd[4 * c4 + 1] = d4.y; nd[4 * c4 + 2] = d4.z; nd[4 * c4 + 3] = d4.w;
      }
      mbar_wait(dp_full, it & 1);
      tc_fence_after();
      #pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t dp[32];
        tmem_ld32(tDP + hf * 32, dp);
        tmem_ld_wait();
        uint32_t w[16];
        #pragma unroll
        for (int c = 0; c < 32; c += 2) {
          float a0, a1;
          if (kStatsMma)
            f2_unpack(f2_mul(f2_pack(pr[hf * 32 + c], pr[hf * 32 + c + 1]), f2_pack(dp[c], dp[c + 1])), a0, a1);
          else
            f2_unpack(f2_mul(f2_pack(pr[hf * 32 + c], pr[hf * 32 + c + 1]),
                             f2_add(f2_pack(dp[c], dp[c + 1]),
                                    f2_pack(nd[kStatsMma ? 0 : c], nd[kStatsMma ? 1 : c + 1]))), a0, a1);
          w[c >> 1] = pack2<kBf16>(a0, a1);
        }
        if (hf == 0 && !kStatsMma) {
          #pragma unroll
          for (int c4 = 0; c4 < (kStatsMma ? 0 : 8); ++c4) {
#ifdef FA_EXPERIMENT_NO_STATS
            const float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
#else
            const float4 d4 = stat[40 + c4];
#endif
            nd[4 * c4] = d4.x; nd[4 * c4 + 1] = d4.y; nd[4 * c4 + 2] = d4.z; nd[4 * c4 + 3] = d4.w;
          }
        }
#ifdef FA_BWD_DK_TS
        tmem_st16(tDP + hf * 16, w);  // the TMEM copy (A operand of dK): columns this thread has already read
#endif
        #pragma unroll
        for (int c8 = 0; c8 < 4; ++c8) {  // 8 columns -> one 16-byte chunk
          const int chunk = (hf * 4 + c8) ^ (row & 7);
          *reinterpret_cast<uint4 *>(ds_row + chunk * 16) =
              make_uint4(w[4 * c8], w[4 * c8 + 1], w[4 * c8 + 2], w[4 * c8 + 3]);
        }
      }
      fence_proxy_async_smem();
#ifdef FA_BWD_DK_TS
      tmem_st_wait();
#endif
      tc_fence_before();
      mbar_arrive_warp(ds_full);
    }

    // ---------------- epilogue: dK * scale, dV -> smem (SW128) -> TMA store ----------------
    if (n > 0) {
      mbar_wait(dkv_full, 0);
      tc_fence_after();
    }
    constexpr int kItems = (D == 128) ? 2 : 1;
    #pragma unroll
    for (int item = 0; item < kItems; ++item) {
      const int which = (D == 128) ? item : wg;   // 0 = dK, 1 = dV
      const int half = (D == 128) ? wg : 0;       // which 64-column slab
      const uint32_t tsrc = tmem_base + (which == 0 ? kColDK : kColDV) + lane_addr + 64 * half;
      const float mul = which == 0 ? p.scale : 1.f;
      uint8_t *dst_row = smem + L::kQOff + which * L::kTileBytes + half * L::kHalfBytes + row * 128;
      #pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t r[32];
        if (n > 0) {
          tmem_ld32(tsrc + hf * 32, r);
          tmem_ld_wait();
        } else {
          #pragma unroll
          for (int e = 0; e < 32; ++e) r[e] = 0u;
        }
        #pragma unroll
        for (int c8 = 0; c8 < 4; ++c8) {
          uint32_t w[4];
          #pragma unroll
          for (int e = 0; e < 4; ++e)
            w[e] = pack2<kBf16>(__uint_as_float(r[c8 * 8 + 2 * e]) * mul, __uint_as_float(r[c8 * 8 + 2 * e + 1]) * mul);
          const int chunk = (hf * 4 + c8) ^ (row & 7);
          *reinterpret_cast<uint4 *>(dst_row + chunk * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
    fence_proxy_async_smem();
    named_bar_sync(2, 256);
    if (threadIdx.x == 0) {
      for (int hh = 0; hh < kHalves; ++hh) {
        tma_store_4d(&tmDK, smem + L::kQOff + hh * L::kHalfBytes, hh * 64, j * 128, hk, b);
        tma_store_4d(&tmDV, smem + L::kQOff + L::kTileBytes + hh * L::kHalfBytes, hh * 64, j * 128, hk, b);
      }
      tma_store_commit();
      tma_store_wait_read<0>();
    }
  } else {
    reg_dealloc<72>();
  }

  tc_fence_before();
  __syncthreads();
#ifdef FA_BWD_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0) p.trace[14] = clock64() - p.trace[13];  // MMA loop end -> CTA exit (epilogue)
#endif
  if (warp == 14) tmem_dealloc<512>(tmem_base);
}

// Pre-kernel: delta = -rowsum(O o dO) (cpp/flash_attention_qkv_packed.cpp:374-377) and
// lse2 = -lse * log2(e) (both stored NEGATED so the main kernel folds them in with packed
// fma/add and no negate modifier; -inf for padded rows and rows without keys => P = 0), both padded to a multiple of 128 rows per (batch, head) so that the main
// kernel can bulk-copy 128-row slices without bounds checks; also zeroes the fp32 dQ accumulator
// row (saves a separate memset pass).  D/8 threads per row, 16-byte loads.
template <typename T, int D>
__global__ void __launch_bounds__(256)
fa_bwd_prep_kernel(const T *o, int64_t o_sb, int64_t o_sh, int64_t o_sl, const T *dout, int64_t d_sb,
                   int64_t d_sh, int64_t d_sl, const float *lse, float *delta, float *lse2,
                   float *dq_acc, int B, int H, int Lq, int Lq_pad, uint8_t *slab_lse = nullptr,
                   uint8_t *slab_delta = nullptr, float inv_scale = 1.f) {
  constexpr int kTpr = D / 8;  // threads per row
  const int sub = threadIdx.x % kTpr;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * (256 / kTpr) + threadIdx.x / kTpr;
  const bool in_range = rid < static_cast<int64_t>(B) * H * Lq_pad;
  const int r = in_range ? static_cast<int>(rid % Lq_pad) : Lq;
  const int g = in_range ? static_cast<int>(rid / Lq_pad) : 0;
  float s = 0.f, l2 = -INFINITY;
  if (r < Lq) {
    const T *orow = o + (g / H) * o_sb + (g % H) * o_sh + static_cast<int64_t>(r) * o_sl + sub * 8;
    const T *drow = dout + (g / H) * d_sb + (g % H) * d_sh + static_cast<int64_t>(r) * d_sl + sub * 8;
    const uint4 ov = *reinterpret_cast<const uint4 *>(orow);
    const uint4 dv = *reinterpret_cast<const uint4 *>(drow);
    const T *oe = reinterpret_cast<const T *>(&ov);
    const T *de = reinterpret_cast<const T *>(&dv);
    #pragma unroll
    for (int e = 0; e < 8; ++e) s = fmaf(static_cast<float>(oe[e]), static_cast<float>(de[e]), s);
    float4 *acc = reinterpret_cast<float4 *>(dq_acc + (static_cast<int64_t>(g) * Lq + r) * D + sub * 8);
    acc[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    acc[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float l = lse[static_cast<int64_t>(g) * Lq + r];
    l2 = (l == -INFINITY) ? -INFINITY : -l * kLog2e;
  }
  #pragma unroll
  for (int off = kTpr / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if (kStatsMma) {
    // the statistics as [128 q x 16] no-swizzle K-major operand slabs (one 4 KB image per Q tile):
    // row q = [hi mid lo 0 0 0 0 0 | 0 x 8] with hi + mid + lo = -lse / softmax_scale (resp. -delta)
    // to 3 x the mantissa of T; rows that must give P = 0 carry a huge negative value instead of -inf
    // (inf x 0 inside the MMA would be NaN)
    if (in_range && sub == 0) {
      const float big = std::is_same<T, __half>::value ? -60000.f : -1e30f;
      const float vals[2] = {l2 == -INFINITY ? big : l2 * (inv_scale / kLog2e), -s};  // l2 = -lse * log2(e)
      uint8_t *dst[2] = {slab_lse, slab_delta};
      const size_t off = static_cast<size_t>(rid >> 7) * 4096 + ((rid & 127) >> 3) * 256 + (rid & 7) * 16;
      #pragma unroll
      for (int w = 0; w < 2; ++w) {
        const T hi = static_cast<T>(vals[w]);
        const float r1 = vals[w] - static_cast<float>(hi);
        const T mid = static_cast<T>(r1);
        const T lo = static_cast<T>(r1 - static_cast<float>(mid));
        const uint32_t w0 = static_cast<uint32_t>(*reinterpret_cast<const unsigned short *>(&hi)) |
                            (static_cast<uint32_t>(*reinterpret_cast<const unsigned short *>(&mid)) << 16);
        const uint32_t w1 = *reinterpret_cast<const unsigned short *>(&lo);
        *reinterpret_cast<uint4 *>(dst[w] + off) = make_uint4(w0, w1, 0u, 0u);
        *reinterpret_cast<uint4 *>(dst[w] + off + 128) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    return;
  }
  if (in_range && sub == 0) { delta[rid] = -s; lse2[rid] = l2; }
}

}  // namespace fa
