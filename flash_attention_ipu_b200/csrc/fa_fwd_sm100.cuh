// fa_fwd_sm100.cuh -- forward attention kernel for B200 (sm_100a).
//
// Replaces the chunked forward graph of the reference,
// flashAttentionQKVPackedWithStash (flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:175-334):
// S = Q K^T (:254), block-causal mask with block skipping (:244, :268-278), online softmax
// (:281-296), O = O*c + P V (:299-300), O /= l and LSE = m + log l (:314-316) -- all fused in
// one kernel.  Nothing of the Poplar structure is kept: tiles arrive by TMA into 128B-swizzled
// shared memory, both GEMMs run on tcgen05 with fp32 accumulators in TMEM, P never leaves the
// SM (it overwrites S in TMEM as a 16-bit A operand), and warps are specialised:
//
//   warps 0-3  : softmax warpgroup for Q tile 0 (thread t <-> query row t, TMEM lane t)
//   warps 4-7  : softmax warpgroup for Q tile 1
//   warp  8    : TMA producer (Q once, K/V rings)
//   warp  9    : MMA issuer (one elected lane)
//   warps 10-11: TMEM allocation / idle
//
// One CTA owns 2 x 128 query rows of one (batch, head) and walks the KV tiles 0..n-1.  The
// issuer interleaves the two Q tiles as  S0(j), PV1(j-1), S1(j), PV0(j)  so that the tensor
// pipe works on one tile while the other tile's softmax runs.
//
// TMEM (512 columns): S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D);
// P_i aliases the first 64 columns of S_i.
#pragma once
#include "sm100_ptx.cuh"

namespace fa {

constexpr int kBlockM = 128;     // query rows per Q tile (= TMEM lanes)
constexpr int kBlockN = 128;     // keys per KV tile
constexpr int kFwdThreads = 384;
#ifndef FA_FWD_K_STAGES
#define FA_FWD_K_STAGES 3
#endif
constexpr int kKStages = FA_FWD_K_STAGES;  // K ring depth (the TMA round trip is latency-, not bandwidth-bound)
#ifndef FA_FWD_V_STAGES
#define FA_FWD_V_STAGES 2
#endif
constexpr int kVStages = FA_FWD_V_STAGES;
#ifndef FA_FWD_P_SPLIT
#define FA_FWD_P_SPLIT 2
#endif
// P_i is handed to the MMA warp in kPSplit column groups, so that O_i += P_i V_j starts on the first
// keys while the exponentials of the last ones are still being computed (shortens the
// S -> softmax -> PV dependency chain that bounds the kernel)
constexpr int kPSplit = FA_FWD_P_SPLIT;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
#ifndef FA_FWD_POLY_EVERY
#define FA_FWD_POLY_EVERY 4
#endif
constexpr int kFwdPolyEvery = FA_FWD_POLY_EVERY;  // every n-th pair of exponentials on the FMA pipe (0: none)
template <bool B> struct BoolC { static constexpr bool value = B; };
constexpr float kRescaleThreshold = 8.0f;  // lazy O rescale: only when the max grows by > 2^8

struct FwdParams {
  float *lse;         // [G][Lq] fp32
  int Lq, Lk, H;      // H: heads per batch (group g -> (b = g / H, h = g % H))
  int kv_ratio;       // query heads per K/V head (1 = multi-head; >1 = grouped-query attention)
  int causal;
  int shift;          // q_pos0 - kv_pos0: key c visible to query r iff c <= r + shift
  float scale_log2;   // softmax_scale * log2(e)
  int num_pairs;      // ceil(Lq / 256)
#ifdef FA_FWD_TRACE  // timeline instrumentation build (never the product)
  unsigned long long *trace;
#endif
};
#ifdef FA_FWD_TRACE
#define FA_TR(cond, j, slot) do { if (blockIdx.x == 16 && (cond) && (j) >= 10 && (j) < 14) p.trace[((j) - 10) * 32 + (slot)] = clock64(); } while (0)
#else
#define FA_TR(cond, j, slot) do { } while (0)
#endif

template <int D>
struct FwdSmem {
  static constexpr int kTileBytes = kBlockM * D * 2;  // one 128-row tile of 16-bit elements
  static constexpr int kHalfBytes = kBlockM * 128;    // one 64-element (128 B) column slab
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + 2 * kTileBytes;
  static constexpr int kVOff = kKOff + kKStages * kTileBytes;
  static constexpr int kBarOff = kVOff + kVStages * kTileBytes;
  static constexpr int kNumBars = 2 + 2 * kKStages + 2 * kVStages + 2 + 2 * kPSplit + 2;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kBytes = kTmemPtrOff + 16;
  static constexpr int kAlloc = kBytes;  // the dynamic shared memory base is 1024-byte aligned (checked)
};

// number of KV tiles Q tile `t` must visit
__device__ __forceinline__ int fwd_num_kv_tiles(const FwdParams &p, int t) {
  if (t * kBlockM >= p.Lq) return 0;
  int hi = p.Lk;  // exclusive bound on visible key index
  if (p.causal) {
    int last_row = min(t * kBlockM + kBlockM, p.Lq) - 1;
    hi = min(hi, last_row + p.shift + 1);
  }
  return hi <= 0 ? 0 : (hi + kBlockN - 1) / kBlockN;
}

template <int D, bool kBf16>
__global__ void __launch_bounds__(kFwdThreads, 1)
fa_fwd_sm100_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                    const FwdParams p) {
  using L = FwdSmem<D>;
  constexpr int kHalves = D / 64;
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) {
    printf("fa_b200: dynamic shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kBarOff);
  uint64_t *q_full = bars;                       // [2]
  uint64_t *k_full = q_full + 2;                 // [kKStages]
  uint64_t *k_empty = k_full + kKStages;         // [kKStages]
  uint64_t *v_full = k_empty + kKStages;         // [kVStages]
  uint64_t *v_empty = v_full + kVStages;         // [kVStages]
  uint64_t *s_full = v_empty + kVStages;         // [2]  MMA -> softmax : S_i ready
  uint64_t *p_full = s_full + 2;                 // [2][kPSplit]  softmax -> MMA : part of P_i ready (4 warp arrivals)
  uint64_t *o_full = p_full + 2 * kPSplit;               // [2]  MMA -> softmax : final O_i ready
  uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(smem + L::kTmemPtrOff);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;

  // work decode: consecutive CTAs share a (batch, head) so K/V stay in L2; heavy tiles first
  const int g = blockIdx.x / p.num_pairs;
  const int pair = p.num_pairs - 1 - (blockIdx.x % p.num_pairs);
  const int b = g / p.H, h = g % p.H;
  const int t0 = 2 * pair;
  const int n_kv[2] = {fwd_num_kv_tiles(p, t0), fwd_num_kv_tiles(p, t0 + 1)};
  const int n_max = max(n_kv[0], n_kv[1]);

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1); mbar_init(&s_full[i], 1);
      for (int q = 0; q < kPSplit; ++q) mbar_init(&p_full[i * kPSplit + q], 4);  // one arrival per warp
      mbar_init(&o_full[i], 1);
    }
    for (int i = 0; i < kKStages; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < kVStages; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    fence_barrier_init();
  }
  if (warp == 10) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 8) {
    // ============================== TMA producer ==============================
    reg_dealloc<88>();
    if (lane == 0) {
      for (int i = 0; i < 2; ++i) {
        mbar_arrive_expect_tx(&q_full[i], L::kTileBytes);
        for (int hh = 0; hh < kHalves; ++hh)
          tma_load_4d(&tmQ, &q_full[i], smem + L::kQOff + i * L::kTileBytes + hh * L::kHalfBytes,
                      hh * 64, (t0 + i) * kBlockM, h, b);
      }
      for (int j = 0; j < n_max; ++j) {
        const int ks = j % kKStages, vs = j % kVStages;
        mbar_wait(&k_empty[ks], ((j / kKStages) & 1) ^ 1);
#ifdef FA_EXPERIMENT_FWD_NO_KV_LOADS  // timing experiment only (wrong results): ring handshake without the copies
        if (j >= 6) {
          mbar_arrive(&k_full[ks]);
          mbar_wait(&v_empty[vs], ((j / kVStages) & 1) ^ 1);
          mbar_arrive(&v_full[vs]);
          continue;
        }
#endif
        mbar_arrive_expect_tx(&k_full[ks], L::kTileBytes);
        for (int hh = 0; hh < kHalves; ++hh)
          tma_load_4d(&tmK, &k_full[ks], smem + L::kKOff + ks * L::kTileBytes + hh * L::kHalfBytes,
                      hh * 64, j * kBlockN, h / p.kv_ratio, b);
        mbar_wait(&v_empty[vs], ((j / kVStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&v_full[vs], L::kTileBytes);
        for (int hh = 0; hh < kHalves; ++hh)
          tma_load_4d(&tmV, &v_full[vs], smem + L::kVOff + vs * L::kTileBytes + hh * L::kHalfBytes,
                      hh * 64, j * kBlockN, h / p.kv_ratio, b);
      }
    }
  } else if (warp == 9) {
    // ============================== MMA issuer ==============================
    // The whole warp walks the loop (warp-uniform control flow, waits included); one elected
    // lane issues the MMAs and their commits.  Descriptors are (lo, hi) words whose per-MMA
    // offsets are compile-time constants, so an MMA costs a couple of integer instructions.
    reg_dealloc<88>();
    {
      constexpr uint32_t idesc_s = make_idesc(kBlockM, kBlockN, kBf16 ? 1 : 0, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc(kBlockM, D, kBf16 ? 1 : 0, 0, 1);
      constexpr uint32_t hi = smem_desc_hi(1024);
      const uint32_t q_lo = smem_desc_lo(smem_u32(smem + L::kQOff), 16);
      const uint32_t k_lo = smem_desc_lo(smem_u32(smem + L::kKOff), 16);
      const uint32_t v_lo = smem_desc_lo(smem_u32(smem + L::kVOff), L::kHalfBytes);
      const bool leader = elect_one();
      const int n0 = n_kv[0], n1 = n_kv[1];

      auto issue_s = [&](int i, int j) {  // S_i = Q_i K_j^T   (both operands K-major, SW128)
        const uint32_t a = q_lo + i * (L::kTileBytes >> 4);
        const uint32_t bb = k_lo + (j % kKStages) * (L::kTileBytes >> 4);
        const uint32_t d = tmem_base + i * 128;
#ifndef FA_EXPERIMENT_FWD_NO_S  // timing experiments only (wrong results)
        #pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = ((k / 4) * L::kHalfBytes + (k % 4) * 32) >> 4;
          umma_ss2(d, a + off, hi, bb + off, hi, idesc_s, k > 0);
        }
#endif
        umma_commit(&s_full[i]);
      };
      // O_i += P_i V_j (A = P in TMEM, B = V MN-major), one group of key columns at a time as the
      // softmax warps release them
      auto issue_pv = [&](int i, int j, uint32_t parity) {
        const uint32_t bb = v_lo + (j % kVStages) * (L::kTileBytes >> 4);
        const uint32_t d = tmem_base + 256 + i * D, a = tmem_base + i * 128;
        constexpr int kSteps = kBlockN / 16 / kPSplit;
        #pragma unroll
        for (int q = 0; q < kPSplit; ++q) {
          mbar_wait(&p_full[i * kPSplit + q], parity);
          tc_fence_after();
          FA_TR(leader && i == 0 && q < 2, j, 1 + q);
#ifndef FA_EXPERIMENT_FWD_NO_PV
          if (leader) {
            #pragma unroll
            for (int k = q * kSteps; k < (q + 1) * kSteps; ++k)
              umma_ts2(d, a + k * 8, bb + k * (2048 >> 4), hi, idesc_pv, (j > 0 || k > 0));
          }
#endif
        }
      };

      mbar_wait(&q_full[0], 0);
      mbar_wait(&q_full[1], 0);
      for (int j = 0; j <= n_max; ++j) {
        FA_TR(leader, j, 16);
        if (j < n_max) {
          mbar_wait(&k_full[j % kKStages], (j / kKStages) & 1);
          tc_fence_after();
        }
        if (j < n0 && leader) {
          FA_TR(true, j, 0);
          issue_s(0, j);
          if (!(j < n1)) umma_commit(&k_empty[j % kKStages]);
          FA_TR(true, j, 17);
        }
        if (j >= 1 && j - 1 < n1) {
          mbar_wait(&v_full[(j - 1) % kVStages], ((j - 1) / kVStages) & 1);
          FA_TR(leader, j, 18);
          issue_pv(1, j - 1, (j - 1) & 1);
          FA_TR(leader, j, 19);
          if (leader) {
            umma_commit(&v_empty[(j - 1) % kVStages]);
            if (j - 1 == n1 - 1) umma_commit(&o_full[1]);
          }
        }
        if (j < n1 && leader) {
          issue_s(1, j);
          umma_commit(&k_empty[j % kKStages]);
          FA_TR(true, j, 20);
        }
        if (j < n0) {
          mbar_wait(&v_full[j % kVStages], (j / kVStages) & 1);
          FA_TR(leader, j, 21);
          issue_pv(0, j, j & 1);
          FA_TR(leader, j, 22);
          if (leader) {
            if (!(j < n1)) umma_commit(&v_empty[j % kVStages]);
            if (j == n0 - 1) umma_commit(&o_full[0]);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp < 8) {
    // ============================== softmax / correction / epilogue ==============================
    reg_alloc<208>();
    const int i = warp >> 2;                       // which Q tile
    const int row_in_tile = (warp & 3) * 32 + lane;
    const int row = (t0 + i) * kBlockM + row_in_tile;
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + i * 128 + lane_addr;
    const uint32_t tO = tmem_base + 256 + i * D + lane_addr;
    const int n = i ? n_kv[1] : n_kv[0];

    float m_used = -INFINITY;  // reference max (log2 domain) the exponentials are taken against
    float l = 0.f;
    for (int j = 0; j < n; ++j) {
      FA_TR(threadIdx.x == 0, j, 8);
      mbar_wait(&s_full[i], j & 1);
      tc_fence_after();
      FA_TR(threadIdx.x == 0, j, 9);
#ifdef FA_EXPERIMENT_FWD_NO_SOFTMAX
      tc_fence_before();
      for (int q = 0; q < kPSplit; ++q) mbar_arrive_warp(&p_full[i * kPSplit + q]);
      FA_TR(threadIdx.x == 0, j, 12);
      continue;
#endif
      uint32_t sr[4][32];
      #pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, sr[c]);
      tmem_ld_wait();
      FA_TR(threadIdx.x == 0, j, 10);

      // mask (only tiles that cross the causal diagonal or the end of the keys)
      const int col0 = j * kBlockN;
      const bool tile_masked = (p.causal && col0 + kBlockN - 1 > (t0 + i) * kBlockM + p.shift) ||
                               (col0 + kBlockN > p.Lk);
      if (tile_masked) {
        int limit = p.Lk - 1 - col0;
        if (p.causal) limit = min(limit, row + p.shift - col0);
        #pragma unroll
        for (int c = 0; c < 4; ++c)
          #pragma unroll
          for (int e = 0; e < 32; ++e)
            if (c * 32 + e > limit) sr[c][e] = 0xff800000u;  // -inf
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // 4 independent 3-input chains
      #pragma unroll
      for (int c = 0; c < 4; ++c)
        #pragma unroll
        for (int e = 0; e < 32; e += 2)
#ifndef FA_EXPERIMENT_FWD_NO_MAX
          mx4[(e >> 1) & 3] = fmax3(mx4[(e >> 1) & 3], __uint_as_float(sr[c][e]), __uint_as_float(sr[c][e + 1]));
#else
          if (e == 0) mx4[c] = __uint_as_float(sr[c][0]);
#endif
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      const float m_new = mx * p.scale_log2;
      FA_TR(threadIdx.x == 0, j, 11);

      // lazy rescale of the running accumulator
      const bool grow = m_new > m_used + kRescaleThreshold;  // also true when m_used == -inf
      if (j > 0 && __any_sync(0xffffffffu, grow)) {
        const float alpha = grow ? fast_exp2(m_used - m_new) : 1.f;  // exp2(-inf) = 0
        l *= alpha;
        #pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t o[32];
          tmem_ld32(tO + c * 32, o);
          tmem_ld_wait();
          #pragma unroll
          for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
          tmem_st32(tO + c * 32, o);
        }
      }
      if (grow) m_used = m_new;
      const float m_eff = (m_used == -INFINITY) ? 0.f : m_used;

      // P = exp2(S * scale_log2 - m): packed FFMA2 for the argument; every kFwdPolyEvery-th pair of
      // exponentials runs as a polynomial on the FMA pipe instead of MUFU (never on masked tiles:
      // the polynomial clamps at 2^-126 instead of returning an exact 0 for -inf)
      const uint64_t sc2 = f2_pack(p.scale_log2, p.scale_log2), nm2 = f2_pack(-m_eff, -m_eff);
      uint64_t sum2[4] = {0ull, 0ull, 0ull, 0ull};
      auto exp_tile = [&](auto use_poly) {
        #pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pk[16];
          #pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float x0, x1, p0, p1;
            f2_unpack(f2_fma(f2_pack(sr[c][e], sr[c][e + 1]), sc2, nm2), x0, x1);
            constexpr int kEvery = kFwdPolyEvery > 0 ? kFwdPolyEvery : 1;
            if (decltype(use_poly)::value && kFwdPolyEvery > 0 && (e >> 1) % kEvery == kEvery - 1) {
              poly_exp2_pair(x0, x1, p0, p1);
            } else {
#ifndef FA_EXPERIMENT_FWD_NO_EXP
              p0 = fast_exp2(x0);
              p1 = fast_exp2(x1);
#else
              p0 = x0; p1 = x1;
#endif
            }
            sum2[(e >> 1) & 3] = f2_add(sum2[(e >> 1) & 3], f2_pack(p0, p1));
#ifndef FA_EXPERIMENT_FWD_NO_PACK
            pk[e >> 1] = pack2<kBf16>(p0, p1);
#else
            if ((e >> 1) < 2) pk[e >> 1] = __float_as_uint(p0 + p1);
#endif
          }
#ifndef FA_EXPERIMENT_FWD_NO_PACK
          tmem_st16(tS + c * 16, pk);
#else
          asm volatile("" ::"r"(pk[0]), "r"(pk[1]));
#endif
          if ((c + 1) % (4 / kPSplit) == 0) {  // this group of P columns is complete: release it
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive_warp(&p_full[i * kPSplit + (c + 1) / (4 / kPSplit) - 1]);
            FA_TR(threadIdx.x == 0, j, 12 + (c + 1) / (4 / kPSplit) - 1);
          }
        }
      };
      if (tile_masked) exp_tile(BoolC<false>{}); else exp_tile(BoolC<true>{});
      {
        float a0, a1, b0, b1;
        f2_unpack(f2_add(f2_add(sum2[0], sum2[1]), f2_add(sum2[2], sum2[3])), a0, a1);
        (void)b0; (void)b1;
        l += a0 + a1;
      }
    }

    // ---------------- epilogue: O_i / l -> smem (SW128) -> TMA store ; LSE ----------------
    uint8_t *o_smem = smem + L::kQOff + i * L::kTileBytes;  // Q_i is dead once o_full fires
    const float inv_l = l > 0.f ? 1.f / l : 0.f;
    if (n > 0) {
      mbar_wait(&o_full[i], 0);
      tc_fence_after();
    } else {
      mbar_wait(&q_full[i], 0);  // the Q load must have landed before its buffer is reused
    }
    #pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t o[32];
      if (n > 0) {
        tmem_ld32(tO + c * 32, o);
        tmem_ld_wait();
      } else {
        #pragma unroll
        for (int e = 0; e < 32; ++e) o[e] = 0u;
      }
      uint8_t *row_base = o_smem + (c >> 1) * L::kHalfBytes + row_in_tile * 128;
      #pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        uint4 w;
        w.x = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 0]) * inv_l, __uint_as_float(o[q4 * 8 + 1]) * inv_l);
        w.y = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 2]) * inv_l, __uint_as_float(o[q4 * 8 + 3]) * inv_l);
        w.z = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 4]) * inv_l, __uint_as_float(o[q4 * 8 + 5]) * inv_l);
        w.w = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 6]) * inv_l, __uint_as_float(o[q4 * 8 + 7]) * inv_l);
        const int chunk = ((c & 1) * 4 + q4) ^ (row_in_tile & 7);
        *reinterpret_cast<uint4 *>(row_base + chunk * 16) = w;
      }
    }
    if (row < p.Lq)
      p.lse[static_cast<size_t>(g) * p.Lq + row] = l > 0.f ? (m_used + log2f(l)) * kLn2 : -INFINITY;
    fence_proxy_async_smem();
    named_bar_sync(1 + i, 128);
    if ((warp & 3) == 0 && lane == 0 && (t0 + i) * kBlockM < p.Lq) {
      for (int hh = 0; hh < kHalves; ++hh)
        tma_store_4d(&tmO, o_smem + hh * L::kHalfBytes, hh * 64, (t0 + i) * kBlockM, h, b);
      tma_store_commit();
      tma_store_wait<0>();
    }
  } else {
    reg_dealloc<88>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 10) tmem_dealloc<512>(tmem_base);
}

}  // namespace fa
