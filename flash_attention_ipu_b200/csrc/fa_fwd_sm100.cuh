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
//   warp  8    : TMA producer + scheduler (Q per item, K/V rings)
//   warp  9    : MMA issuer (one elected lane)
//   warps 10-11: TMEM allocation / idle
//
// One work item = 2 x 128 query rows of one (batch, head) against the KV tiles 0..n-1.  The
// issuer interleaves the two Q tiles as  S0(j), PV1(j-1), S1(j), PV0(j)  so that the tensor
// pipe works on one tile while the other tile's softmax runs.
//
// PERSISTENT: one CTA per SM walks items handed out by an atomic counter in the order the
// hardware scheduler used to launch them (group-major so K/V stay in L2, the pairs with most KV
// tiles first).  Rings and hand-off barriers run on counters that continue across items: the next
// item's Q tiles load as soon as this item's last S products have completed, its first K/V tiles
// as soon as ring stages free up, and its first S products are issued while the softmax warps
// are still busy with this item's output.  O_i is staged in the dead Q_i buffer and leaves by TMA
// store; only the storing thread waits for the store to have read the buffer, then releases it to
// the producer.  Measured alternatives (profiles/r02_fwd_*_cycles.log, config 3): this 2.787 M
// cycles; a staging buffer of its own (16 KB per warpgroup, taken from a third K stage that buys
// nothing) so that Q_i reloads earlier 3.086 M -- the next item's SS products then run while the
// store reads shared memory, and TMA reads yield to the tensor core's operand reads; stores
// straight from registers (32 rows x 16 bytes per warp instruction) 3.33 M; non-persistent 2.84 M.
//
// TMEM (512 columns): S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [256+D,256+2D);
// P_i aliases the first 64 columns of S_i.
#pragma once
#include "sm100_ptx.cuh"

namespace fa {

constexpr int kBlockM = 128;     // query rows per Q tile (= TMEM lanes)
constexpr int kBlockN = 128;     // keys per KV tile
constexpr int kFwdThreads = 384;
#ifndef FA_FWD_HINT_Q
#define FA_FWD_HINT_Q 0    // L2 eviction hints (sm100_ptx.cuh): Q tiles are read once,
#endif
#ifndef FA_FWD_HINT_OUT
#define FA_FWD_HINT_OUT 0  // O is written once
#endif
#ifndef FA_FWD_K_STAGES
#define FA_FWD_K_STAGES 3
#endif
#ifndef FA_FWD_OWN_OSTAGE
#define FA_FWD_OWN_OSTAGE 1
#endif
// D = 64 leaves 80 KB of shared memory unused: the output tiles get a staging buffer of their own there, so
// that the next item's Q tiles load as soon as this item's last S products have completed instead of after
// its output has left (short items -- BERT / GPT-2 shapes walk 2-8 KV tiles -- pay that wait once per
// item).  At D = 128 there is no room, and a buffer taken from the K ring measured slower (header).
template <int D> struct FwdOwnOStage { static constexpr bool value = D == 64 && FA_FWD_OWN_OSTAGE; };
constexpr int kKStages = FA_FWD_K_STAGES;  // K ring depth (2 measures the same once the kernel is persistent)
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
constexpr int kFwdConsumerWarps = 9;       // 8 softmax warps + the MMA warp read every scheduled item

struct FwdParams {
  float *lse;         // [G][Lq] fp32
  const int *kv_lens; // [B*H] keys of each (batch, head) that exist (<= Lk; the rest is padding), or nullptr
  int *sched_counter; // [2], both zero at launch: [0] next unclaimed item - gridDim.x, [1] CTAs that are done claiming;
                      // the last of them zeroes both again (no memset between launches).  nullptr: grid = items
  int Lq, Lk, H;      // H: heads per batch (group g -> (b = g / H, h = g % H))
  int kv_ratio;       // query heads per K/V head (1 = multi-head; >1 = grouped-query attention)
  int causal;
  int shift;          // q_pos0 - kv_pos0: key c visible to query r iff c <= r + shift
  float scale_log2;   // softmax_scale * log2(e)
  int num_pairs;      // ceil(Lq / 256)
  int num_items;      // B * H * num_pairs
  int order_chunk;    // item order: the groups are walked in chunks of this many (K + V of a chunk stay in L2), and
                      // inside a chunk the heaviest pairs of ALL its groups come first, so that a chunk -- and the
                      // whole launch -- ends with its lightest items.  1 = group-major, >= groups = one global order
  // dropout on P (kDrop instantiation only; sm100_ptx.cuh: drop_hash): keep key iff its hash byte >= thresh
  uint32_t drop_thresh4;  // thresh * 0x01010101
  uint32_t drop_seed;     // 32-bit fold of fa_b200_problem.dropout_seed
  float drop_scale;       // 256 / (256 - thresh)
  int drop_q0, drop_k40;  // q_pos0 ; kv_pos0 / 4 (global positions feed the hash)
  // additive score bias (kBias instantiation only): S' = scale * q.k + bias[b][h][q][k], bias in the kernel's
  // 16-bit dtype, element strides (0 = broadcast over that dim), unit stride on k
  const uint16_t *bias;
  long long bias_sb, bias_sh, bias_sl;
};

template <int D>
struct FwdSmem {
  static constexpr int kTileBytes = kBlockM * D * 2;  // one 128-row tile of 16-bit elements
  static constexpr int kHalfBytes = kBlockM * 128;    // one 64-element (128 B) column slab
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kQOff + 2 * kTileBytes;
  static constexpr int kVOff = kKOff + kKStages * kTileBytes;
  static constexpr int kOOff = kVOff + kVStages * kTileBytes;  // 2 output staging tiles (FwdOwnOStage only)
  static constexpr int kBarOff = kOOff + (FwdOwnOStage<D>::value ? 2 * kTileBytes : 0);
  static constexpr int kNumBars = 4 + 2 * kKStages + 2 * kVStages + 2 + 2 * kPSplit + 2 + 4;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kSchedOff = kTmemPtrOff + 8;   // int[2]: scheduled item per slot (-1 = no more work)
  static constexpr int kBytes = kSchedOff + 8;
  static constexpr int kAlloc = kBytes;  // the dynamic shared memory base is 1024-byte aligned (checked)
};

// number of KV tiles Q tile `t` must visit; `lk` = keys of this group that exist
__device__ __forceinline__ int fwd_num_kv_tiles(const FwdParams &p, int t, int lk) {
  if (t * kBlockM >= p.Lq) return 0;
  int hi = lk;  // exclusive bound on visible key index
  if (p.causal) {
    int last_row = min(t * kBlockM + kBlockM, p.Lq) - 1;
    hi = min(hi, last_row + p.shift + 1);
  }
  return hi <= 0 ? 0 : (hi + kBlockN - 1) / kBlockN;
}

// one work item: Q tiles t0, t0 + 1 of group g; consecutive items share a (batch, head) so K/V stay
// in L2, and within a group the pairs with most KV tiles come first
struct FwdItem {
  int g, b, h, t0, n0, n1, n_max, lk;
};
__device__ __forceinline__ FwdItem fwd_decode(const FwdParams &p, int w) {
  FwdItem t;
  const int groups = p.num_items / p.num_pairs;
  const int chunk = w / (p.order_chunk * p.num_pairs);
  const int g0 = chunk * p.order_chunk, gc = min(p.order_chunk, groups - g0);  // the last chunk may be short
  const int r = w - g0 * p.num_pairs;
  t.g = g0 + r % gc;
  const int pair = p.num_pairs - 1 - r / gc;
  t.b = t.g / p.H;
  t.h = t.g % p.H;
  t.t0 = 2 * pair;
  t.lk = p.kv_lens ? min(p.Lk, max(__ldg(p.kv_lens + t.g), 0)) : p.Lk;  // key-padding: trailing keys do not exist
  t.n0 = fwd_num_kv_tiles(p, t.t0, t.lk);
  t.n1 = fwd_num_kv_tiles(p, t.t0 + 1, t.lk);
  t.n_max = max(t.n0, t.n1);
  return t;
}

template <int D, bool kBf16, bool kDrop = false, bool kBias = false>
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
  uint64_t *q_empty = q_full + 2;                // [2]  softmax -> producer : O_i has left the Q_i buffer it was staged in
  uint64_t *k_full = q_empty + 2;                // [kKStages]
  uint64_t *k_empty = k_full + kKStages;         // [kKStages]
  uint64_t *v_full = k_empty + kKStages;         // [kVStages]
  uint64_t *v_empty = v_full + kVStages;         // [kVStages]
  uint64_t *s_full = v_empty + kVStages;         // [2]  MMA -> softmax : S_i ready
  uint64_t *p_full = s_full + 2;                 // [2][kPSplit]  softmax -> MMA : part of P_i ready (4 warp arrivals)
  uint64_t *o_full = p_full + 2 * kPSplit;       // [2]  MMA -> softmax : final O_i ready
  uint64_t *sched_full = o_full + 2;             // [2]  producer -> consumers : slot holds the next item
  uint64_t *sched_empty = sched_full + 2;        // [2]  consumers -> producer : every consumer warp has read it
  uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(smem + L::kTmemPtrOff);
  volatile int *sched_item = reinterpret_cast<volatile int *>(smem + L::kSchedOff);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); mbar_init(&s_full[i], 1);
      for (int q = 0; q < kPSplit; ++q) mbar_init(&p_full[i * kPSplit + q], 4);  // one arrival per warp
      mbar_init(&o_full[i], 1);
      mbar_init(&sched_full[i], 1); mbar_init(&sched_empty[i], kFwdConsumerWarps);
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
  griddep_launch_dependents();
  griddep_wait();  // nothing above touched global memory (sm100_ptx.cuh: programmatic dependent launch)

  // every consumer warp fetches item number `k` of this CTA the same way
  auto next_item = [&](int k) -> int {
    const int s = k & 1;
    mbar_wait(&sched_full[s], (k >> 1) & 1);
    const int w = sched_item[s];
    mbar_arrive_warp(&sched_empty[s]);
    return w;
  };

  if (warp == 8) {
    // ============================== scheduler + TMA producer ==============================
    reg_dealloc<88>();
    if (lane == 0) {
      int w = blockIdx.x;
      uint32_t kj = 0, kq = 0;  // KV tiles loaded so far (ring position); items with loads so far
      for (int k = 0;; ++k) {
        const int s = k & 1;
        mbar_wait(&sched_empty[s], ((k >> 1) & 1) ^ 1);
        sched_item[s] = w < p.num_items ? w : -1;
        mbar_arrive(&sched_full[s]);
        if (w >= p.num_items) {
          // this CTA will not touch the work counter again; the last CTA to get here leaves it zeroed
          // for the next launch that is handed this slot (stream order separates the two launches)
          if (p.sched_counter) {
            __threadfence();
            if (atomicAdd(p.sched_counter + 1, 1) == static_cast<int>(gridDim.x) - 1) {
              p.sched_counter[0] = 0;
              p.sched_counter[1] = 0;
            }
          }
          break;
        }
        const FwdItem t = fwd_decode(p, w);
        w = p.sched_counter ? gridDim.x + atomicAdd(p.sched_counter, 1) : p.num_items;
        if (t.n_max == 0) continue;
        for (int i = 0; i < 2; ++i) {
          mbar_wait(&q_empty[i], (kq & 1) ^ 1);  // the previous item's S_i products have read the buffer
          mbar_arrive_expect_tx(&q_full[i], L::kTileBytes);
          for (int hh = 0; hh < kHalves; ++hh)
            tma_load_4d_h<FA_FWD_HINT_Q>(&tmQ, &q_full[i], smem + L::kQOff + i * L::kTileBytes + hh * L::kHalfBytes,
                                         hh * 64, (t.t0 + i) * kBlockM, t.h, t.b);
        }
        ++kq;
        for (int j = 0; j < t.n_max; ++j, ++kj) {
          const uint32_t ks = kj % kKStages, vs = kj % kVStages;
          mbar_wait(&k_empty[ks], ((kj / kKStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&k_full[ks], L::kTileBytes);
          for (int hh = 0; hh < kHalves; ++hh)
            tma_load_4d(&tmK, &k_full[ks], smem + L::kKOff + ks * L::kTileBytes + hh * L::kHalfBytes,
                        hh * 64, j * kBlockN, t.h / p.kv_ratio, t.b);
          mbar_wait(&v_empty[vs], ((kj / kVStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&v_full[vs], L::kTileBytes);
          for (int hh = 0; hh < kHalves; ++hh)
            tma_load_4d(&tmV, &v_full[vs], smem + L::kVOff + vs * L::kTileBytes + hh * L::kHalfBytes,
                        hh * 64, j * kBlockN, t.h / p.kv_ratio, t.b);
        }
      }
    }
  } else if (warp == 9) {
    // ============================== MMA issuer ==============================
    // The whole warp walks the loop (warp-uniform control flow, waits included); one elected
    // lane issues the MMAs and their commits.  Descriptors are (lo, hi) words whose per-MMA
    // offsets are compile-time constants, so an MMA costs a couple of integer instructions.
    reg_dealloc<88>();
    constexpr uint32_t idesc_s = make_idesc(kBlockM, kBlockN, kBf16 ? 1 : 0, 0, 0);
    constexpr uint32_t idesc_pv = make_idesc(kBlockM, D, kBf16 ? 1 : 0, 0, 1);
    constexpr uint32_t hi = smem_desc_hi(1024);
    const uint32_t q_lo = smem_desc_lo(smem_u32(smem + L::kQOff), 16);
    const uint32_t k_lo = smem_desc_lo(smem_u32(smem + L::kKOff), 16);
    const uint32_t v_lo = smem_desc_lo(smem_u32(smem + L::kVOff), L::kHalfBytes);
    const bool leader = elect_one();
    uint32_t kbase = 0, kq = 0;  // ring position of the item's first KV tile; items with loads so far
    uint32_t cp[2] = {0, 0};     // softmax steps consumed so far per Q tile (parity of p_full)

    auto issue_s = [&](int i, uint32_t kj) {  // S_i = Q_i K_j^T   (both operands K-major, SW128)
      const uint32_t a = q_lo + i * (L::kTileBytes >> 4);
      const uint32_t bb = k_lo + (kj % kKStages) * (L::kTileBytes >> 4);
      const uint32_t d = tmem_base + i * 128;
      #pragma unroll
      for (int k = 0; k < D / 16; ++k) {
        const uint32_t off = ((k / 4) * L::kHalfBytes + (k % 4) * 32) >> 4;
        umma_ss2(d, a + off, hi, bb + off, hi, idesc_s, k > 0);
      }
      umma_commit(&s_full[i]);
    };
    // O_i += P_i V_j (A = P in TMEM, B = V MN-major), one group of key columns at a time as the
    // softmax warps release them; `first` = the item's first KV tile (the accumulator is overwritten)
    auto issue_pv = [&](int i, uint32_t kj, bool first) {
      const uint32_t bb = v_lo + (kj % kVStages) * (L::kTileBytes >> 4);
      const uint32_t d = tmem_base + 256 + i * D, a = tmem_base + i * 128;
      constexpr int kSteps = kBlockN / 16 / kPSplit;
      #pragma unroll
      for (int q = 0; q < kPSplit; ++q) {
        mbar_wait(&p_full[i * kPSplit + q], cp[i] & 1);
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int k = q * kSteps; k < (q + 1) * kSteps; ++k)
            umma_ts2(d, a + k * 8, bb + k * (2048 >> 4), hi, idesc_pv, (!first || k > 0));
        }
      }
      ++cp[i];
    };

    for (int k = 0;; ++k) {
      const int w = next_item(k);
      if (w < 0) break;
      const FwdItem t = fwd_decode(p, w);
      if (t.n_max == 0) continue;
      const int n0 = t.n0, n1 = t.n1;
      mbar_wait(&q_full[0], kq & 1);
      mbar_wait(&q_full[1], kq & 1);
      for (int j = 0; j <= t.n_max; ++j) {
        const uint32_t kj = kbase + j;
        if (j < t.n_max) {
          mbar_wait(&k_full[kj % kKStages], (kj / kKStages) & 1);
          tc_fence_after();
        }
        if (j < n0 && leader) {
          issue_s(0, kj);
          if (!(j < n1)) umma_commit(&k_empty[kj % kKStages]);
          if (FwdOwnOStage<D>::value && j == n0 - 1) umma_commit(&q_empty[0]);  // Q_0 has been read for the last time
        }
        if (j >= 1 && j - 1 < n1) {
          mbar_wait(&v_full[(kj - 1) % kVStages], ((kj - 1) / kVStages) & 1);
          issue_pv(1, kj - 1, j == 1);
          if (leader) {
            umma_commit(&v_empty[(kj - 1) % kVStages]);
            if (j - 1 == n1 - 1) umma_commit(&o_full[1]);
          }
        }
        if (j < n1 && leader) {
          issue_s(1, kj);
          umma_commit(&k_empty[kj % kKStages]);
          if (FwdOwnOStage<D>::value && j == n1 - 1) umma_commit(&q_empty[1]);
        }
        if (j < n0) {
          mbar_wait(&v_full[kj % kVStages], (kj / kVStages) & 1);
          issue_pv(0, kj, j == 0);
          if (leader) {
            if (!(j < n1)) umma_commit(&v_empty[kj % kVStages]);
            if (j == n0 - 1) umma_commit(&o_full[0]);
          }
        }
        __syncwarp();
      }
      kbase += t.n_max;
      ++kq;
    }
  } else if (warp < 8) {
    // ============================== softmax / correction / output ==============================
    reg_alloc<208>();
    const int i = warp >> 2;                       // which Q tile
    const int row_in_tile = (warp & 3) * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + i * 128 + lane_addr;
    const uint32_t tO = tmem_base + 256 + i * D + lane_addr;
    uint32_t cs = 0, co = 0;  // S_i tiles / final O_i tiles consumed so far (barrier parities)
    uint32_t kq = 0;          // items with loads so far (parity of q_full)

    for (int k = 0;; ++k) {
      const int w = next_item(k);
      if (w < 0) break;
      const FwdItem t = fwd_decode(p, w);
      const int row = (t.t0 + i) * kBlockM + row_in_tile;
      const int n = i ? t.n1 : t.n0;
      // dropout: the query's half of the hash input (sm100_ptx.cuh: drop_hash)
      const uint32_t drop_qh = kDrop ? drop_seed_of_group(p.drop_seed, static_cast<uint32_t>(t.g)) ^
                                       (static_cast<uint32_t>(p.drop_q0 + row) * 0x9E3779B1u) : 0u;

      const uint16_t *bias_row = kBias ? p.bias + t.b * p.bias_sb + t.h * p.bias_sh + static_cast<long long>(row) * p.bias_sl
                                       : nullptr;

      float m_used = -INFINITY;  // reference max (log2 domain) the exponentials are taken against
      float l = 0.f;
      for (int j = 0; j < n; ++j, ++cs) {
        // bias of this thread's row of the tile: 16 independent 16-byte loads (8 keys each), issued BEFORE the
        // wait for S so that their latency runs under it (Lk % 8 == 0 is checked by the shim, so a vector is
        // either entirely inside the row or entirely beyond it); the next tile's two lines are pulled into L2
        uint4 bv[kBias ? 16 : 1];
        if (kBias) {
          const bool row_ok = row < p.Lq;
          #pragma unroll
          for (int v = 0; v < 16; ++v) {
            const int col = j * kBlockN + v * 8;
            bv[v] = make_uint4(0u, 0u, 0u, 0u);
            if (row_ok && col + 8 <= p.Lk) bv[v] = __ldg(reinterpret_cast<const uint4 *>(bias_row + col));
          }
          if (row_ok && j + 1 < n) {
            #pragma unroll
            for (int ln = 0; ln < 2; ++ln) {
              const int col = (j + 1) * kBlockN + ln * 64;
              if (col < p.Lk) asm volatile("prefetch.global.L2 [%0];" ::"l"(bias_row + col));
            }
          }
        }
        mbar_wait(&s_full[i], cs & 1);
        tc_fence_after();
        uint32_t sr[4][32];
        #pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, sr[c]);
        tmem_ld_wait();

        const int col0 = j * kBlockN;
        if (kBias) {
          // S' = S * scale + bias, taken to the log2 domain here (the exponentials below then use sr as it is)
          #pragma unroll
          for (int c = 0; c < 4; ++c) {
            #pragma unroll
            for (int v8 = 0; v8 < 4; ++v8) {
              const uint4 b4 = bv[c * 4 + v8];
              const uint32_t w4[4] = {b4.x, b4.y, b4.z, b4.w};
              #pragma unroll
              for (int e2 = 0; e2 < 4; ++e2) {
                float b0, b1;
                unpack2<kBf16>(w4[e2], b0, b1);
                const int e = v8 * 8 + e2 * 2;
                sr[c][e] = __float_as_uint(fmaf(__uint_as_float(sr[c][e]), p.scale_log2, b0 * kLog2e));
                sr[c][e + 1] = __float_as_uint(fmaf(__uint_as_float(sr[c][e + 1]), p.scale_log2, b1 * kLog2e));
              }
            }
          }
        }
        // mask (only tiles that cross the causal diagonal or the end of the keys)
        const bool tile_masked = (p.causal && col0 + kBlockN - 1 > (t.t0 + i) * kBlockM + p.shift) ||
                                 (col0 + kBlockN > t.lk);
        if (tile_masked) {
          int limit = t.lk - 1 - col0;
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
            mx4[(e >> 1) & 3] = fmax3(mx4[(e >> 1) & 3], __uint_as_float(sr[c][e]), __uint_as_float(sr[c][e + 1]));
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_new = kBias ? mx : mx * p.scale_log2;  // with a bias the scores are already scaled

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
        const float sc1 = kBias ? 1.f : p.scale_log2;
        const uint64_t sc2 = f2_pack(sc1, sc1), nm2 = f2_pack(-m_eff, -m_eff);
        uint64_t sum2[4] = {0ull, 0ull, 0ull, 0ull};
        auto exp_tile = [&](auto use_poly) {
          #pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t pk[16];
            uint32_t keep[8];  // dropout: per-byte keep masks of this chunk's 8 groups of 4 keys
            if (kDrop) {
              const uint32_t k4 = static_cast<uint32_t>(p.drop_k40 + (col0 >> 2) + c * 8);
              #pragma unroll
              for (int g4 = 0; g4 < 8; ++g4)
                keep[g4] = drop_keep_bytes(drop_mix(drop_qh ^ ((k4 + g4) * 0x85EBCA77u)), p.drop_thresh4);
            }
            #pragma unroll
            for (int e = 0; e < 32; e += 2) {
              float x0, x1, p0, p1;
              f2_unpack(f2_fma(f2_pack(sr[c][e], sr[c][e + 1]), sc2, nm2), x0, x1);
              constexpr int kEvery = kFwdPolyEvery > 0 ? kFwdPolyEvery : 1;
              if (decltype(use_poly)::value && kFwdPolyEvery > 0 && (e >> 1) % kEvery == kEvery - 1) {
                poly_exp2_pair(x0, x1, p0, p1);
              } else {
                p0 = fast_exp2(x0);
                p1 = fast_exp2(x1);
              }
              sum2[(e >> 1) & 3] = f2_add(sum2[(e >> 1) & 3], f2_pack(p0, p1));
              pk[e >> 1] = pack2<kBf16>(p0, p1);
              // the row sum above is of the UNdropped P (softmax denominator); only the PV operand is thinned
              if (kDrop) pk[e >> 1] &= __byte_perm(keep[e >> 2], 0u, ((e >> 1) & 1) ? 0x3322u : 0x1100u);
            }
            tmem_st16(tS + c * 16, pk);
            if ((c + 1) % (4 / kPSplit) == 0) {  // this group of P columns is complete: release it
              tmem_st_wait();
              tc_fence_before();
              mbar_arrive_warp(&p_full[i * kPSplit + (c + 1) / (4 / kPSplit) - 1]);
            }
          }
        };
        // a bias may hold -inf anywhere: like masked tiles, never through the polynomial
        if (tile_masked || kBias) exp_tile(BoolC<false>{}); else exp_tile(BoolC<true>{});
        {
          float a0, a1;
          f2_unpack(f2_add(f2_add(sum2[0], sum2[1]), f2_add(sum2[2], sum2[3])), a0, a1);
          l += a0 + a1;
        }
      }

      // ---------------- item end: O_i / l -> global (straight from registers) ; LSE ----------------
      // Each thread owns one output row (D contiguous 16-bit values): 16-byte stores, two per 32-byte
      // sector.  No shared-memory staging: the Q buffers already belong to the next item.
      const float inv_l = (l > 0.f ? 1.f / l : 0.f) * (kDrop ? p.drop_scale : 1.f);
      if (n > 0) {
        mbar_wait(&o_full[i], co & 1);
        tc_fence_after();
        ++co;
      }
      if (t.n_max > 0) {
        constexpr bool kOwn = FwdOwnOStage<D>::value;
        // staged in the dead Q_i buffer (Q_i is dead once o_full fires), or in a buffer of its own
        uint8_t *o_smem = kOwn ? smem + L::kOOff + i * L::kTileBytes : smem + L::kQOff + i * L::kTileBytes;
        if (n == 0) mbar_wait(&q_full[i], kq & 1);              // its load must have landed before the buffer is reused
        ++kq;
        if (kOwn) {  // the previous item's store must have read the staging buffer (long done by now)
          if ((warp & 3) == 0 && lane == 0) tma_store_wait_read<0>();
          named_bar_sync(1 + i, 128);
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
            uint4 wv;
            wv.x = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 0]) * inv_l, __uint_as_float(o[q4 * 8 + 1]) * inv_l);
            wv.y = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 2]) * inv_l, __uint_as_float(o[q4 * 8 + 3]) * inv_l);
            wv.z = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 4]) * inv_l, __uint_as_float(o[q4 * 8 + 5]) * inv_l);
            wv.w = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 6]) * inv_l, __uint_as_float(o[q4 * 8 + 7]) * inv_l);
            const int chunk = ((c & 1) * 4 + q4) ^ (row_in_tile & 7);
            *reinterpret_cast<uint4 *>(row_base + chunk * 16) = wv;
          }
        }
        fence_proxy_async_smem();
        named_bar_sync(1 + i, 128);
        if ((warp & 3) == 0 && lane == 0) {
          if ((t.t0 + i) * kBlockM < p.Lq) {
            for (int hh = 0; hh < kHalves; ++hh)
              tma_store_4d_h<FA_FWD_HINT_OUT>(&tmO, o_smem + hh * L::kHalfBytes, hh * 64, (t.t0 + i) * kBlockM, t.h, t.b);
            tma_store_commit();
            if (!kOwn) tma_store_wait_read<0>();
          }
          // the next item's Q_i may be loaded (own staging: the MMA warp says so after the last S_i product,
          // unless this tile had none)
          if (!kOwn || n == 0) mbar_arrive(&q_empty[i]);
        }
      }
      if (row < p.Lq)
        p.lse[static_cast<size_t>(t.g) * p.Lq + row] = l > 0.f ? (m_used + log2f(l)) * kLn2 : -INFINITY;
      // the next item's first PV_i overwrites O_i: it waits for this warpgroup's next P, which is
      // issued after the loads above in program order
      tc_fence_before();
    }
    if (FwdOwnOStage<D>::value && (warp & 3) == 0 && lane == 0) tma_store_wait_read<0>();  // before the CTA gives up its smem
  } else {
    reg_dealloc<88>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 10) tmem_dealloc<512>(tmem_base);
}

}  // namespace fa
