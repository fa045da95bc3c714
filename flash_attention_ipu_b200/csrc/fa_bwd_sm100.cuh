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
//   S^T  = K_j Q_i^T      SS   A = K (K-major)          B = Q_i  (K-major)   -> TMEM S
//   dP^T = V_j dO_i^T     SS   A = V (K-major)          B = dO_i (K-major)   -> TMEM dP
//   dV  += P^T dO_i       TS   A = P^T  (TMEM, 16-bit)  B = dO_i (MN-major)  -> TMEM dV
//   dK  += dS^T Q_i       TS   A = dS^T (TMEM, 16-bit)  B = Q_i  (MN-major)  -> TMEM dK
//   dQ_i = dS K_j         SS   A = dS^T smem (MN-major) B = K (MN-major)     -> TMEM dQ (aliases dP)
//
// TMEM (512 cols, D=128): dK [0,128) dV [128,256) S/P^T [256,384) dP^T/dS^T/dQ [384,512).
// Issue order per Q tile:  dP(i) dV(i) S(i+1) dK(i) dQ(i).  dQ tiles are accumulated across KV
// tiles in an fp32 scratch by TMA reduce-add and cast afterwards.
//
// PERSISTENT: one CTA per SM walks work items (batch, K/V head, KV tile) handed out by an atomic
// counter in the order the hardware scheduler used to launch them (group-major, the KV tiles with
// most Q tiles first).  Every ring / hand-off barrier runs on counters that continue across items,
// so the loads of the next item's first Q / dO tiles, the L2 prefetch of its K / V and the store
// of this item's dK / dV overlap instead of costing a CTA launch + prologue + epilogue each
// (measured before: 6 000-7 000 clk entry -> first MMA, 5 800 clk last MMA -> exit, per KV tile).
//
//   warps 0-7  : compute (2 warpgroups; thread <-> key row; WG w owns query columns [64w,64w+64))
//   warps 8-11 : dQ drain (TMEM -> registers, TMEM released, -> smem slabs -> TMA reduce-add)
//   warp 12 TMA producer + scheduler | warp 13 MMA issuer | warp 14 TMEM alloc | warp 15 idle
//
// What bounds it: the TMA unit moves reduce-adds at 25 B/clk/SM
// and its shared-memory reads yield to the tensor core's operand reads, so every byte of dQ
// that leaves the SM costs tensor-pipe time.  Hence dK as a TS product (fewer operand reads),
// the drain releasing TMEM before it touches the staging slabs, and 4 KB reduce operations.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "fa_fwd_sm100.cuh"

#if defined(FA_EXPERIMENT_NO_REDUCE) && !defined(FA_ALLOW_EXPERIMENTS)
#error "FA_EXPERIMENT_* builds remove work and give wrong results: define FA_ALLOW_EXPERIMENTS to build one on purpose"
#endif

namespace fa {

constexpr int kBwdThreads = 512;
// L2 eviction hints of the backward's TMA traffic (0 none, 1 evict_first, 2 evict_last), probed at the power cap
#ifndef FA_BWD_HINT_KV
#define FA_BWD_HINT_KV 0   // K / V tiles: read once per item
#endif
#ifndef FA_BWD_HINT_Q
#define FA_BWD_HINT_Q 0    // Q / dO tiles: re-read by every KV tile of the group
#endif
#ifndef FA_BWD_HINT_DQ
#define FA_BWD_HINT_DQ 0   // fp32 dQ partial sums: re-touched by every KV tile of the group
#endif
#ifndef FA_BWD_HINT_OUT
#define FA_BWD_HINT_OUT 0  // dK / dV: written once
#endif
#ifndef FA_BWD_POLY_PAIRS
#define FA_BWD_POLY_PAIRS 4
#endif
// of a thread's 32 pairs of exponentials, how many run on the FMA pipe.  12 gives the fewest cycles (7.31 M on
// config 3), but the pool's B200s run this kernel AT THEIR 1000 W CAP, where the polynomial's ~10 FMA-pipe
// instructions per pair cost more energy than one MUFU op: 0 / 4 pairs measure 968-971 TFLOP/s, 8 / 12 pairs
// 940-946 (profiles/r02_power_probe_poly.log; sustained >= 1.5 s legs, alternating builds)
constexpr int kBwdPolyPairs = FA_BWD_POLY_PAIRS;
constexpr int kBwdRegCompute = 144, kBwdRegDrain = 152, kBwdRegOther = 72;  // 8x32x144 + 4x32x152 + 4x32x72 = 65536
constexpr int kBwdConsumerWarps = 13;             // compute 8 + drain 4 + MMA 1 read every scheduled item

struct BwdParams {
  const float *lse2;   // [G][Lq_pad]  -lse * log2(e); -inf for padding / rows without keys
  const float *delta;  // [G][Lq_pad]  -rowsum(O o dO); 0 for padding
  const int *kv_lens;  // [B*H] keys of each (batch, query head) that exist (<= Lk), or nullptr
  int *sched_counter;  // zeroed by fa_bwd_prep_kernel; next unclaimed item = gridDim.x + counter
  unsigned long long *clock_out;  // [2]: SM cycles and nanoseconds CTA 0 spent in the kernel (effective SM clock under load)
  // deterministic dQ (fa_b200_set_deterministic): dq_sem[(group * num_q_tiles + i) * 4 + drain warp] counts the KV
  // tiles whose partial of Q tile i has been reduce-added; zeroed by fa_bwd_prep_kernel.  nullptr: arrival order.
  int *dq_sem;
  int Lq, Lk, H, Lq_pad;
  int Hkv, kv_ratio;   // K/V heads and query heads per K/V head (H = Hkv * kv_ratio; 1 = multi-head)
  int causal, shift;
  float scale, scale_log2;
  int num_kv_tiles;    // ceil(Lk / 128)
  int num_items;       // B * Hkv * num_kv_tiles
  int order_chunk;     // groups per scheduling chunk (FwdParams::order_chunk): inside a chunk the KV tiles with most
                       // Q tiles of all its groups come first.  1 = group-major
  // dropout (kDrop instantiation only), as in FwdParams: the same hash regenerates the forward's mask
  uint32_t drop_thresh4, drop_seed;
  float drop_scale;
  int drop_q0, drop_k40;
  // additive score bias (kBias instantiation only), as in FwdParams.  It only enters P = exp(S' - LSE); dS is
  // the gradient with respect to S' as well, so dQ / dK / dV need nothing else (no gradient for the bias itself)
  const uint16_t *bias;
  long long bias_sb, bias_sh, bias_sl;
#ifdef FA_BWD_TRACE
  unsigned long long *trace;  // CTA 0: cycles the MMA warp spent in each of its waits
#endif
};

#ifndef FA_BWD_KV_STAGES2
#define FA_BWD_KV_STAGES2 1
#endif
// D = 64 leaves 80 KB of shared memory unused: K / V get a second stage there, so that the next item's K / V
// load while this item still runs instead of after its last MMAs (short items -- 4-8 Q tiles at the BERT /
// GPT-2 shapes -- pay that latency once per item).  No room at D = 128, and nothing to gain with 64-tile items.
template <int D> struct BwdKvStages { static constexpr int value = (D == 64 && FA_BWD_KV_STAGES2) ? 2 : 1; };

template <int D>
struct BwdSmem {
  static constexpr int kTileBytes = 128 * D * 2;
  static constexpr int kHalfBytes = 128 * 128;
  static constexpr int kKvStages = BwdKvStages<D>::value;
  static constexpr int kKvStageBytes = 2 * kTileBytes;        // K | V of one stage
  static constexpr int kKOff = 0;
  static constexpr int kVOff = kKOff + kTileBytes;
  static constexpr int kQOff = kKOff + kKvStages * kKvStageBytes;  // 2 stages
  static constexpr int kDOOff = kQOff + 2 * kTileBytes;       // 1 stage
  static constexpr int kDSOff = kDOOff + kTileBytes;          // dS^T [128 keys][128 q] 16-bit (item end: dK / dV staging)
  static constexpr int kDQOff = kDSOff + 2 * kHalfBytes;      // fp32 staging: 4 drain warps x 2 slabs x 4 KB
  static constexpr int kSlabBytes = 4096, kDrainSlabs = 2;
  static constexpr int kStatOff = kDQOff + 4 * kDrainSlabs * kSlabBytes;  // 2 stages x (lse2[128] | delta[128]) fp32
  static constexpr int kBarOff = kStatOff + 2 * 1024;
  static constexpr int kNumBars = 22;
  static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
  static constexpr int kSchedOff = kTmemPtrOff + 8;           // int[2]: scheduled item per slot (-1 = no more work)
  static constexpr int kBytes = kSchedOff + 8;
};

// first Q tile that sees any key of KV tile j
__device__ __forceinline__ int bwd_first_q_tile(const BwdParams &p, int j) {
  if (!p.causal) return 0;
  const int a = j * 128 - p.shift;
  return a <= 0 ? 0 : a / 128;
}

// one work item: KV tile j of (batch b, K/V head hk); the CTA walks the Q tiles i0 .. of every
// query head that shares the K/V head (grouped-query attention: kv_ratio heads, dK / dV simply keep
// accumulating in TMEM across them)
struct BwdItem {
  int b, hk, j, i0, n, h0, lk;
};
template <bool kGqa>
__device__ __forceinline__ BwdItem bwd_decode(const BwdParams &p, int w, int num_q_tiles) {
  BwdItem t;
  const int gks = p.num_items / p.num_kv_tiles;
  const int chunk = w / (p.order_chunk * p.num_kv_tiles);
  const int g0 = chunk * p.order_chunk, gc = min(p.order_chunk, gks - g0);  // the last chunk may be short
  const int r = w - g0 * p.num_kv_tiles;
  const int gk = g0 + r % gc;
  const int jw = r / gc;
  // ascending: the KV tiles with most Q tiles first.  Deterministic dQ: DESCENDING, because the
  // partials of a Q tile are then added in descending KV-tile order and the tile a CTA waits for
  // must have been claimed before its own (it is then running or done: no deadlock, and it reaches
  // every shared Q tile one iteration earlier, so the wait is normally already satisfied)
  t.j = p.dq_sem ? p.num_kv_tiles - 1 - jw : jw;
  t.b = gk / p.Hkv;
  t.hk = gk % p.Hkv;
  t.i0 = bwd_first_q_tile(p, t.j);
  t.h0 = kGqa ? t.hk * p.kv_ratio : t.hk;
  // key padding (one length per batch element; stored per (batch, query head)): a KV tile made of
  // padding only has no Q tile to visit and writes dK = dV = 0
  t.lk = p.kv_lens ? min(p.Lk, max(__ldg(p.kv_lens + t.b * p.H + t.h0), 0)) : p.Lk;
  const int n1 = t.j * 128 < t.lk ? max(num_q_tiles - t.i0, 0) : 0;
  t.n = kGqa ? n1 * p.kv_ratio : n1;
  return t;
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
// the multi-head instantiation keeps the plain induction variable, which the compiler holds in
// uniform registers (the loop-carried (h, i) pair cost 6 % of the backward).
// kDrop = dropout on P: with M the keep mask and s = 1 / (1 - p), the forward computed O = s (M o P) V, so
//   dV = s (M o P)^T dO   (the thinned P^T is the A operand; s is applied once, at the item's end)
//   dS = P o (s M o dP - delta)   with delta = rowsum(O o dO) as before and P the UNdropped softmax.
// kBias: 0 none, 1 one bias row per query (general), 2 one bias row for every query (row stride 0: a key mask /
// key bias -- a single load per thread and tile instead of 64)
template <int D, bool kBf16, bool kGqa, bool kDrop = false, int kBias = 0>
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
  uint64_t *s_full = bars + 7;    // MMA -> compute : S^T ready
  uint64_t *p_full = bars + 8;    // compute -> MMA : P^T in TMEM           (8 warp arrivals)
  uint64_t *dp_full = bars + 9;   // MMA -> compute : dP^T ready
  uint64_t *ds_full = bars + 10;  // compute -> MMA : dS^T in TMEM + smem   (8 warp arrivals)
  uint64_t *dq_full = bars + 11;  // MMA -> drain   : dQ ready
  uint64_t *dq_empty = bars + 12; // drain -> MMA   : dQ left TMEM          (4 warp arrivals)
  uint64_t *dkv_full = bars + 13; // MMA -> compute, producer : the item's MMAs are complete (dK, dV final; K, V free)
  uint64_t *sched_full = bars + 14;   // [2] producer -> consumers : slot holds the next item
  uint64_t *sched_empty = bars + 16;  // [2] consumers -> producer : every consumer warp has read it
  // second K / V stage (BwdKvStages): its own "full" barrier, and one "free" barrier per stage (MMA -> producer)
  auto kv_full_of = [&](uint32_t st) -> uint64_t * { return st ? bars + 18 : kv_full; };
  auto kv_free_of = [&](uint32_t st) -> uint64_t * { return bars + 19 + st; };
  uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(smem + L::kTmemPtrOff);
  volatile int *sched_item = reinterpret_cast<volatile int *>(smem + L::kSchedOff);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int num_q_tiles = (p.Lq + 127) / 128;
  long long clk0 = 0;
  unsigned long long ns0 = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) { clk0 = clock64(); ns0 = global_timer_ns(); }

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
    mbar_init(bars + 18, 1); mbar_init(bars + 19, 1); mbar_init(bars + 20, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1);
      mbar_init(&sched_full[s], 1); mbar_init(&sched_empty[s], kBwdConsumerWarps);
    }
    mbar_init(do_full, 1); mbar_init(do_empty, 1);
    mbar_init(s_full, 1); mbar_init(p_full, 8); mbar_init(dp_full, 1); mbar_init(ds_full, 8);  // per-warp arrivals
    mbar_init(dq_full, 1); mbar_init(dq_empty, 4); mbar_init(dkv_full, 1);
    fence_barrier_init();
  }
  if (warp == 14) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  griddep_launch_dependents();
  griddep_wait();  // nothing above touched global memory (sm100_ptx.cuh: programmatic dependent launch)

  // Every consumer warp fetches item number `k` of this CTA the same way: wait for the slot, read
  // it, and tell the producer the slot may be reused (item k + 2).
  auto next_item = [&](int k) -> int {
    const int s = k & 1;
    mbar_wait(&sched_full[s], (k >> 1) & 1);
    const int w = sched_item[s];
    mbar_arrive_warp(&sched_empty[s]);
    return w;
  };

  if (warp == 12) {
    // ============================== scheduler + TMA producer ==============================
    reg_dealloc<kBwdRegOther>();
    if (lane == 0) {
      int w = blockIdx.x;
      uint32_t git = 0, kc = 0;  // Q tiles loaded so far; items with at least one Q tile so far
      for (int k = 0;; ++k) {
        const int s = k & 1;
        mbar_wait(&sched_empty[s], ((k >> 1) & 1) ^ 1);
        sched_item[s] = w < p.num_items ? w : -1;
        mbar_arrive(&sched_full[s]);
        if (w >= p.num_items) break;
        const BwdItem t = bwd_decode<kGqa>(p, w, num_q_tiles);
        // claim the following item now, and pull its K / V tiles into L2 while this one runs.
        // Deterministic dQ claims only when this item is done: CTAs wait for the KV tile claimed
        // just before theirs, so a claimed item must be one that is being worked on.
        int w_next = p.dq_sem ? p.num_items : gridDim.x + atomicAdd(p.sched_counter, 1);
        if (w_next < p.num_items) {
          const BwdItem tn = bwd_decode<kGqa>(p, w_next, num_q_tiles);
          if (tn.n > 0)
            for (int hh = 0; hh < kHalves; ++hh) {
              tma_prefetch_l2_4d(&tmK, hh * 64, tn.j * 128, tn.hk, tn.b);
              tma_prefetch_l2_4d(&tmV, hh * 64, tn.j * 128, tn.hk, tn.b);
            }
        }
        for (int it = 0, h = t.h0, i = t.i0; it < t.n; ++it, ++git, bwd_next_tile<kGqa>(h, i, t.i0, num_q_tiles)) {
          const uint32_t st = git & 1;
          const float *lse_g = p.lse2 + static_cast<size_t>(t.b * p.H + h) * p.Lq_pad;
          const float *delta_g = p.delta + static_cast<size_t>(t.b * p.H + h) * p.Lq_pad;
          mbar_wait(&q_empty[st], ((git >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&q_full[st], L::kTileBytes + 1024);
          for (int hh = 0; hh < kHalves; ++hh)
            tma_load_4d_h<FA_BWD_HINT_Q>(&tmQ, &q_full[st], smem + L::kQOff + st * L::kTileBytes + hh * L::kHalfBytes,
                                         hh * 64, i * 128, h, t.b);
          bulk_load_1d(smem + L::kStatOff + st * 1024, lse_g + i * 128, 512, &q_full[st]);
          bulk_load_1d(smem + L::kStatOff + st * 1024 + 512, delta_g + i * 128, 512, &q_full[st]);
          mbar_wait(do_empty, (git & 1) ^ 1);
          mbar_arrive_expect_tx(do_full, L::kTileBytes);
          for (int hh = 0; hh < kHalves; ++hh)
            tma_load_4d_h<FA_BWD_HINT_Q>(&tmDO, do_full, smem + L::kDOOff + hh * L::kHalfBytes, hh * 64, i * 128, h, t.b);
          if (it == 0) {
            // K, V after the first Q / dO tiles: those only need a free ring stage, K / V need the
            // previous item's last MMAs to have completed
            const uint32_t ks = L::kKvStages == 2 ? (kc & 1) : 0;
            if (L::kKvStages == 2) mbar_wait(kv_free_of(ks), ((kc >> 1) & 1) ^ 1);  // the item two back has left the stage
            else if (kc > 0) mbar_wait(dkv_full, (kc - 1) & 1);
            uint64_t *full = kv_full_of(ks);
            uint8_t *kv_smem = smem + ks * L::kKvStageBytes;
            mbar_arrive_expect_tx(full, 2 * L::kTileBytes);
            for (int hh = 0; hh < kHalves; ++hh) {
              tma_load_4d_h<FA_BWD_HINT_KV>(&tmK, full, kv_smem + L::kKOff + hh * L::kHalfBytes, hh * 64, t.j * 128, t.hk, t.b);
              tma_load_4d_h<FA_BWD_HINT_KV>(&tmV, full, kv_smem + L::kVOff + hh * L::kHalfBytes, hh * 64, t.j * 128, t.hk, t.b);
            }
          }
        }
        if (t.n > 0) ++kc;
        if (p.dq_sem) w_next = gridDim.x + atomicAdd(p.sched_counter, 1);
        w = w_next;
      }
    }
  } else if (warp == 13) {
    // ============================== MMA issuer ==============================
    // Warp-uniform loop, one elected lane issues; (lo, hi) descriptors with constant offsets.
    reg_dealloc<kBwdRegOther>();
    constexpr int kFmt = kBf16 ? 1 : 0;
    constexpr uint32_t idesc_kk = make_idesc(128, 128, kFmt, 0, 0);  // S^T, dP^T
    constexpr uint32_t idesc_dv = make_idesc(128, D, kFmt, 0, 1);    // dV, dK (A in TMEM, B MN-major)
    constexpr uint32_t idesc_dq = make_idesc(128, D, kFmt, 1, 1);    // dQ (A and B MN-major)
    constexpr uint32_t hi = smem_desc_hi(1024);
    constexpr uint32_t kTile16 = L::kTileBytes >> 4;
    // K-major views (LBO unused) and MN-major views (LBO = distance between 64-column slabs)
    const uint32_t k_kmaj0 = smem_desc_lo(smem_u32(smem + L::kKOff), 16);
    const uint32_t v_kmaj0 = smem_desc_lo(smem_u32(smem + L::kVOff), 16);
    const uint32_t q_kmaj = smem_desc_lo(smem_u32(smem + L::kQOff), 16);
    const uint32_t do_kmaj = smem_desc_lo(smem_u32(smem + L::kDOOff), 16);
    const uint32_t k_mn0 = smem_desc_lo(smem_u32(smem + L::kKOff), L::kHalfBytes);
    const uint32_t q_mn = smem_desc_lo(smem_u32(smem + L::kQOff), L::kHalfBytes);
    const uint32_t do_mn = smem_desc_lo(smem_u32(smem + L::kDOOff), L::kHalfBytes);
    const uint32_t ds_mn = smem_desc_lo(smem_u32(smem + L::kDSOff), L::kHalfBytes);
    const uint32_t tS = tmem_base + kColS, tDP = tmem_base + kColDP;
    const uint32_t tDK = tmem_base + kColDK, tDV = tmem_base + kColDV;
    const bool leader = elect_one();

    auto issue_kk = [&](uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo) {  // d = a b^T, both K-major
      #pragma unroll
      for (int k = 0; k < D / 16; ++k) {
        const uint32_t off = ((k / 4) * L::kHalfBytes + (k % 4) * 32) >> 4;
        umma_ss2(d_tmem, a_lo + off, hi, b_lo + off, hi, idesc_kk, k > 0);
      }
    };
    // 16-bit A operand written by the compute warps over the first half of each warpgroup's 64 columns
    auto a_cols = [](int k) -> uint32_t { return k < 4 ? 8 * k : 64 + 8 * (k - 4); };

#ifdef FA_BWD_TRACE
    long long tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long t_in = 0, t_gap = 0, t_last = clock64();
    int n_it = 0, n_items = 0;
#endif
    uint32_t git = 0, kc = 0;
    for (int k = 0;; ++k) {
      const int w = next_item(k);
      if (w < 0) break;
      const BwdItem t = bwd_decode<kGqa>(p, w, num_q_tiles);
      if (t.n == 0) continue;
      const uint32_t ks = L::kKvStages == 2 ? (kc & 1) : 0;
      const uint32_t kv_off = ks * (L::kKvStageBytes >> 4);  // descriptor units of 16 bytes
      const uint32_t k_kmaj = k_kmaj0 + kv_off, v_kmaj = v_kmaj0 + kv_off, k_mn = k_mn0 + kv_off;
      mbar_wait(kv_full_of(ks), L::kKvStages == 2 ? (kc >> 1) & 1 : kc & 1);
      mbar_wait(&q_full[git & 1], (git >> 1) & 1);
      tc_fence_after();
      if (leader) {
        issue_kk(tS, k_kmaj, q_kmaj + (git & 1) * kTile16);  // S^T(0)
        umma_commit(s_full);
      }
#ifdef FA_BWD_TRACE
      const long long t_loop = clock64();
      t_gap += t_loop - t_last;
#endif
      for (int it = 0; it < t.n; ++it, ++git) {
        const uint32_t q_stage = (git & 1) * kTile16;
        // dP^T = V dO^T   (overwrites the previous tile's dQ: the drain must have emptied it)
        FA_BT_WAIT(0, mbar_wait(do_full, git & 1));
        if (git > 0) FA_BT_WAIT(1, mbar_wait(dq_empty, (git - 1) & 1));
        tc_fence_after();
        if (leader) {
          issue_kk(tDP, v_kmaj, do_kmaj);
          umma_commit(dp_full);
        }
        // dV += P^T dO
        FA_BT_WAIT(2, mbar_wait(p_full, git & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_ts2(tDV, tS + a_cols(kk), do_mn + kk * (2048 >> 4), hi, idesc_dv, (it > 0 || kk > 0));
          umma_commit(do_empty);
        }
        // S^T of the next Q tile (overwrites P^T, which dV has consumed by then: in-order pipe)
        if (it + 1 < t.n) {
          FA_BT_WAIT(3, mbar_wait(&q_full[(git + 1) & 1], ((git + 1) >> 1) & 1));
          tc_fence_after();
          if (leader) {
            issue_kk(tS, k_kmaj, q_kmaj + ((git + 1) & 1) * kTile16);
            umma_commit(s_full);
          }
        }
        // dK += dS^T Q (A from TMEM, over the dP^T columns) ; dQ = dS K (overwrites dP^T / dS^T)
        FA_BT_WAIT(4, mbar_wait(ds_full, git & 1));
        tc_fence_after();
        if (leader) {
          #pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_ts2(tDK, tDP + a_cols(kk), q_mn + q_stage + kk * (2048 >> 4), hi, idesc_dv, (it > 0 || kk > 0));
          umma_commit(&q_empty[git & 1]);
          #pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_ss2(tDP, ds_mn + kk * (2048 >> 4), hi, k_mn + kk * (2048 >> 4), hi, idesc_dq, kk > 0);
          umma_commit(dq_full);
        }
        __syncwarp();
      }
      if (leader) {
        umma_commit(dkv_full);
        if (L::kKvStages == 2) umma_commit(kv_free_of(ks));
      }
      ++kc;
#ifdef FA_BWD_TRACE
      t_last = clock64();
      t_in += t_last - t_loop;
      n_it += t.n;
      ++n_items;
#endif
    }
#ifdef FA_BWD_TRACE
    if (blockIdx.x == 0 && lane == 0) {
      for (int s = 0; s < 5; ++s) p.trace[s] = tr[s];
      p.trace[5] = t_in; p.trace[6] = n_it; p.trace[7] = n_items; p.trace[8] = t_gap;
    }
#endif
  } else if (warp >= 8 && warp < 12) {
    // ============================== dQ drain ==============================
    // Each warp drains its own 32 query rows: all D columns go to registers at once and TMEM is
    // released (dq_empty) BEFORE the staging slabs are touched, so dP^T of the next tile never
    // waits for the reduce engine.  Then a private ring of two 4 KB slabs (32 rows x 32 fp32,
    // SWIZZLE_128B) feeds TMA reduce-adds; a slab is re-filled once the reduce that last used it
    // has READ it.
    reg_alloc<kBwdRegDrain>();
    const int wq = warp & 3;
    const uint32_t tDQ = tmem_base + kColDP + (static_cast<uint32_t>(wq * 32) << 16);
    uint8_t *wstage = smem + L::kDQOff + wq * (L::kDrainSlabs * L::kSlabBytes);
    uint32_t git = 0;
    for (int k = 0;; ++k) {
      const int w = next_item(k);
      if (w < 0) break;
      const BwdItem t = bwd_decode<kGqa>(p, w, num_q_tiles);
      for (int it = 0, h = t.h0, i = t.i0; it < t.n; ++it, ++git, bwd_next_tile<kGqa>(h, i, t.i0, num_q_tiles)) {
        mbar_wait(dq_full, git & 1);
        tc_fence_after();
        constexpr int kSlabs = D / 32;
        uint32_t r[kSlabs][32];
        #pragma unroll
        for (int c = 0; c < kSlabs; ++c) tmem_ld32(tDQ + c * 32, r[c]);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive_warp(dq_empty);  // dQ has left TMEM: nothing below is on the tensor pipe's path
        int *sem = nullptr;
        if (p.dq_sem) {
          // ordered accumulation: wait until every KV tile above this one that sees Q tile i has added
          // its partial of these 32 rows.  j_max = last KV tile with bwd_first_q_tile(j) <= i.
          int j_max = (t.lk + 127) / 128 - 1;  // KV tiles made of padding only add nothing
          if (p.causal) j_max = min(j_max, (i * 128 + 127 + p.shift) / 128);
          sem = p.dq_sem + ((static_cast<size_t>(t.b * p.H + h) * num_q_tiles + i) * 4 + wq);
          if (lane == 0) {
            const int want = j_max - t.j;
            while (ld_acquire_gpu(sem) < want) { }
          }
          __syncwarp();
        }
        #pragma unroll
        for (int c = 0; c < kSlabs; ++c) {
          uint8_t *slab = wstage + (c % L::kDrainSlabs) * L::kSlabBytes;
          if (lane == 0) tma_store_wait_read<L::kDrainSlabs - 1>();  // the reduce that last used this slab has read it
          __syncwarp();
          #pragma unroll
          for (int q = 0; q < 8; ++q) {
            uint4 v4 = make_uint4(r[c][q * 4], r[c][q * 4 + 1], r[c][q * 4 + 2], r[c][q * 4 + 3]);
            *reinterpret_cast<uint4 *>(slab + lane * 128 + ((q ^ (lane & 7)) * 16)) = v4;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
#ifndef FA_EXPERIMENT_NO_REDUCE  // timing / power experiment only when defined (wrong results; needs FA_ALLOW_EXPERIMENTS)
            tma_reduce_add_4d_h<FA_BWD_HINT_DQ>(&tmDQ, slab, c * 32, i * 128 + wq * 32, h, t.b);
#endif
            tma_store_commit();
          }
        }
        if (sem != nullptr && lane == 0) {
          tma_store_wait<0>();       // the reduce-adds have been PERFORMED, not merely read from the slabs
          red_release_gpu_add(sem, 1);
        }
      }
    }
    // the slabs must stay valid until the last reduce has READ them; completion of the writes
    // themselves is guaranteed by the kernel boundary (the convert pass is a later launch)
    if (lane == 0) tma_store_wait_read<0>();
  } else if (warp < 8) {
    // ============================== compute: P^T, dS^T ; per item: dK, dV out ==============================
    reg_alloc<kBwdRegCompute>();
    const int wg = warp >> 2;                      // which 64 query columns of the tile
    const int row = (warp & 3) * 32 + lane;        // key row within the KV tile (= TMEM lane)
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + kColS + lane_addr + 64 * wg;
    const uint32_t tDP = tmem_base + kColDP + lane_addr + 64 * wg;
    uint8_t *ds_row = smem + L::kDSOff + wg * L::kHalfBytes + row * 128;
    uint32_t git = 0, kc = 0;

    for (int k = 0;; ++k) {
      const int w = next_item(k);
      if (w < 0) break;
      const BwdItem t = bwd_decode<kGqa>(p, w, num_q_tiles);
      const int key = t.j * 128 + row;

      for (int it = 0, hx = 0, i = t.i0; it < t.n; ++it, ++git, bwd_next_tile<kGqa>(hx, i, t.i0, num_q_tiles)) {
        const float4 *stat = reinterpret_cast<const float4 *>(smem + L::kStatOff + (git & 1) * 1024) + wg * 16;
        // -lse2 of this warpgroup's 64 query columns: loaded (broadcast LDS.128) BEFORE the S^T wait so
        // that the shared-memory latency hides behind it
        float nl[64];
        mbar_wait(&q_full[git & 1], (git >> 1) & 1);   // makes the lse2 / delta stage visible to this thread
        #pragma unroll
        for (int c4 = 0; c4 < 16; ++c4) {
          const float4 l4 = stat[c4];
          nl[4 * c4] = l4.x; nl[4 * c4 + 1] = l4.y; nl[4 * c4 + 2] = l4.z; nl[4 * c4 + 3] = l4.w;
        }
        if (kBias) {
          // bias[q][key] * log2(e) folded into -lse2 of the query column: thread <-> key, so the 32 lanes of a
          // warp read 32 neighbouring keys of one query row (one 64-byte segment per load)
          const uint16_t *bcol = p.bias + t.b * p.bias_sb + (t.h0 + hx) * p.bias_sh + key;
          const int q0 = i * 128 + 64 * wg;
          if (kBias == 2) {
            // one bias row for every query (a key mask / key bias): a single load per thread and tile
            const float bk = key < p.Lk ? unpack1<kBf16>(__ldg(bcol)) * kLog2e : 0.f;
            #pragma unroll
            for (int c = 0; c < 64; ++c) nl[c] += bk;
          } else {
          // two batches of 32 independent loads (all issued before the first is used), each under the wait for S^T
          #pragma unroll
          for (int c0 = 0; c0 < 64; c0 += 32) {
            uint16_t raw[32];
            #pragma unroll
            for (int c = 0; c < 32; ++c) {
              const bool ok = q0 + c0 + c < p.Lq && key < p.Lk;
              raw[c] = ok ? __ldg(bcol + static_cast<long long>(q0 + c0 + c) * p.bias_sl) : static_cast<uint16_t>(0);
            }
            #pragma unroll
            for (int c = 0; c < 32; ++c) nl[c0 + c] = fmaf(unpack1<kBf16>(raw[c]), kLog2e, nl[c0 + c]);
          }
          // the next Q tile's 64 rows of this warp's keys -> L2 (two rows per lane)
#ifndef FA_BIAS_PF_DIST
#define FA_BIAS_PF_DIST 1  // Q tiles ahead (without it the loads run at DRAM latency: 8.0 instead of 2.5 ms, B4 H32 T2048)
#endif
          if (i + FA_BIAS_PF_DIST < num_q_tiles) {
            #pragma unroll
            for (int r2 = 0; r2 < 2; ++r2) {
              const int qn = (i + FA_BIAS_PF_DIST) * 128 + 64 * wg + 2 * lane + r2;
              if (qn < p.Lq && key < p.Lk)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(bcol - lane + static_cast<long long>(qn) * p.bias_sl));
            }
          }
          }  // general (per-query) bias
        }
        // dropout: bit c of keepm = query column c of this warpgroup keeps this thread's key (generated
        // ahead of the S^T wait; sm100_ptx.cuh: drop_hash).  Thread <-> key, and one hash word covers the 4
        // keys of 4 neighbouring lanes for one query: each of the 4 lanes computes the words of a quarter of
        // the 64 query columns and the quad exchanges them by shuffle (a quarter of the mixing work).
        uint32_t keepm[2] = {0xffffffffu, 0xffffffffu};
        if (kDrop) {
          const uint32_t kh = drop_seed_of_group(p.drop_seed, static_cast<uint32_t>(t.b * p.H + t.h0 + hx)) ^
                              (static_cast<uint32_t>(p.drop_k40 + (key >> 2)) * 0x85EBCA77u);  // same in the 4 lanes of a quad
          const uint32_t bit = (key & 3) * 8;  // this key's byte of the word: bit 0 of the byte after the compare
          const uint32_t q0 = static_cast<uint32_t>(p.drop_q0 + i * 128 + 64 * wg);
          const int quad = lane & ~3, me = lane & 3;
          #pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t m = 0;
            #pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              // lane `me` of the quad: per-byte keep masks (0xff / 0x00) of query column 4 * c4 + me
              const uint32_t mine = drop_keep_bytes(drop_mix(kh ^ ((q0 + hf * 32 + 4 * c4 + me) * 0x9E3779B1u)), p.drop_thresh4);
              #pragma unroll
              for (int rr = 0; rr < 4; ++rr)
                m |= ((__shfl_sync(0xffffffffu, mine, quad | rr) >> bit) & 1u) << (4 * c4 + rr);
            }
            keepm[hf] = m;
          }
        }
        mbar_wait(s_full, git & 1);
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
              f2_unpack(f2_fma(f2_pack(s[c >> 5][c & 31], s[c >> 5][(c & 31) + 1]), c2, f2_pack(nl[c], nl[c + 1])),
                        x0, x1);
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
        if (p.causal && t.j * 128 + 127 > i * 128 + p.shift) {
          const int cmin = key - p.shift - i * 128 - 64 * wg;  // columns below cmin are masked
          #pragma unroll
          for (int c = 0; c < 64; ++c)
            if (c < cmin) pr[c] = 0.f;
        }
        // key rows past the end of K (last KV tile of a ragged Lk): K is zero-filled there, so
        // S = 0 and exp2(-lse2) is unbounded for rows with a strongly negative LSE.  Select, not
        // multiply: the value may be inf / polynomial garbage.  (The forward masks the same keys
        // through `limit = Lk - 1 - col0`.)
        if (t.j * 128 + 128 > t.lk && key >= t.lk) {
          #pragma unroll
          for (int c = 0; c < 64; ++c) pr[c] = 0.f;
        }
        {
          uint32_t pk[32];
          #pragma unroll
          for (int c = 0; c < 32; ++c) {
            pk[c] = pack2<kBf16>(pr[2 * c], pr[2 * c + 1]);
            if (kDrop) {  // thin the dV operand only; pr stays the undropped softmax for dS
              const uint32_t two = (keepm[c >> 4] >> ((2 * c) & 31)) & 3u;
              pk[c] &= ((two & 1u) ? 0x0000ffffu : 0u) | ((two & 2u) ? 0xffff0000u : 0u);
            }
          }
          tmem_st32(tS, pk);  // P^T over the first half of this warpgroup's own S^T columns
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive_warp(p_full);
        }
        // dS^T = P^T o (dP^T - delta); -delta of each 32-column half is fetched ahead of the
        // barrier / TMEM load it is used after
        float nd[32];
        #pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          const float4 d4 = stat[32 + c4];
          nd[4 * c4] = d4.x; nd[4 * c4 + 1] = d4.y; nd[4 * c4 + 2] = d4.z; nd[4 * c4 + 3] = d4.w;
        }
        mbar_wait(dp_full, git & 1);
        tc_fence_after();
        #pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t dp[32];
          tmem_ld32(tDP + hf * 32, dp);
          tmem_ld_wait();
          uint32_t wd[16];
          #pragma unroll
          for (int c = 0; c < 32; c += 2) {
            float a0, a1;
            if (kDrop) {
              const uint32_t d0 = (keepm[hf] >> c) & 1u ? dp[c] : 0u, d1 = (keepm[hf] >> (c + 1)) & 1u ? dp[c + 1] : 0u;
              f2_unpack(f2_mul(f2_pack(pr[hf * 32 + c], pr[hf * 32 + c + 1]),
                               f2_fma(f2_pack(d0, d1), f2_pack(p.drop_scale, p.drop_scale), f2_pack(nd[c], nd[c + 1]))), a0, a1);
            } else
            f2_unpack(f2_mul(f2_pack(pr[hf * 32 + c], pr[hf * 32 + c + 1]),
                             f2_add(f2_pack(dp[c], dp[c + 1]), f2_pack(nd[c], nd[c + 1]))), a0, a1);
            wd[c >> 1] = pack2<kBf16>(a0, a1);
          }
          if (hf == 0) {
            #pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              const float4 d4 = stat[40 + c4];
              nd[4 * c4] = d4.x; nd[4 * c4 + 1] = d4.y; nd[4 * c4 + 2] = d4.z; nd[4 * c4 + 3] = d4.w;
            }
          }
          // the TMEM copy (A operand of dK) goes over dP^T columns this thread has already read;
          // the shared-memory copy (A operand of dQ, read MN-major) is 128B-swizzled
          tmem_st16(tDP + hf * 16, wd);
          #pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {  // 8 columns -> one 16-byte chunk
            const int chunk = (hf * 4 + c8) ^ (row & 7);
            *reinterpret_cast<uint4 *>(ds_row + chunk * 16) =
                make_uint4(wd[4 * c8], wd[4 * c8 + 1], wd[4 * c8 + 2], wd[4 * c8 + 3]);
          }
        }
        fence_proxy_async_smem();
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive_warp(ds_full);
      }

      // ---------------- item end: dK * scale, dV -> smem (SW128) -> TMA store ----------------
      // Staged through the dS^T buffer (32 KB), which is free once the item's last MMAs are done:
      // one 128 x D tile per round at D = 128 (dK, then dV), both tiles in one round at D = 64.
      // The Q / dO rings and K / V already belong to the next item.
      if (t.n > 0) {
        mbar_wait(dkv_full, kc & 1);
        tc_fence_after();
        ++kc;
      }
      constexpr int kRounds = (D == 128) ? 2 : 1;
      #pragma unroll
      for (int round = 0; round < kRounds; ++round) {
        const int which = (D == 128) ? round : wg;  // 0 = dK, 1 = dV
        const int half = (D == 128) ? wg : 0;       // which 64-column slab of the tile
        const uint32_t tsrc = tmem_base + (which == 0 ? kColDK : kColDV) + lane_addr + 64 * half;
        const float mul = which == 0 ? p.scale : (kDrop ? p.drop_scale : 1.f);
        uint8_t *dst_row = smem + L::kDSOff + wg * L::kHalfBytes + row * 128;
        #pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t r[32];
          if (t.n > 0) {
            tmem_ld32(tsrc + hf * 32, r);
            tmem_ld_wait();
          } else {
            #pragma unroll
            for (int e = 0; e < 32; ++e) r[e] = 0u;
          }
          #pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {
            uint32_t wv[4];
            #pragma unroll
            for (int e = 0; e < 4; ++e)
              wv[e] = pack2<kBf16>(__uint_as_float(r[c8 * 8 + 2 * e]) * mul, __uint_as_float(r[c8 * 8 + 2 * e + 1]) * mul);
            const int chunk = (hf * 4 + c8) ^ (row & 7);
            *reinterpret_cast<uint4 *>(dst_row + chunk * 16) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
          }
        }
        fence_proxy_async_smem();
        named_bar_sync(2, 256);
        if (threadIdx.x == 0) {
          if (D == 128) {
            const CUtensorMap *tm = round == 0 ? &tmDK : &tmDV;
            for (int hh = 0; hh < 2; ++hh)
              tma_store_4d_h<FA_BWD_HINT_OUT>(tm, smem + L::kDSOff + hh * L::kHalfBytes, hh * 64, t.j * 128, t.hk, t.b);
          } else {
            tma_store_4d_h<FA_BWD_HINT_OUT>(&tmDK, smem + L::kDSOff, 0, t.j * 128, t.hk, t.b);
            tma_store_4d_h<FA_BWD_HINT_OUT>(&tmDV, smem + L::kDSOff + L::kHalfBytes, 0, t.j * 128, t.hk, t.b);
          }
          tma_store_commit();
          tma_store_wait_read<0>();  // the buffer is written again (next round / next item's dS^T)
        }
        named_bar_sync(2, 256);
      }
    }
  } else {
    reg_dealloc<kBwdRegOther>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 14) tmem_dealloc<512>(tmem_base);
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // a persistent CTA stays on one SM: both counters are that SM's
    p.clock_out[0] = static_cast<unsigned long long>(clock64() - clk0);
    p.clock_out[1] = global_timer_ns() - ns0;
  }
}

// Pre-kernel: delta = -rowsum(O o dO) (cpp/flash_attention_qkv_packed.cpp:374-377) and
// lse2 = -lse * log2(e) (both stored NEGATED so the main kernel folds them in with packed
// fma/add and no negate modifier; -inf for padded rows and rows without keys => P = 0), both padded
// to a multiple of 128 rows per (batch, head) so that the main kernel can bulk-copy 128-row slices
// without bounds checks; also zeroes the fp32 dQ accumulator row (saves a separate memset pass)
// and the work counter of the persistent main kernel.  D/8 threads per row, 16-byte loads.
template <typename T, int D>
__global__ void __launch_bounds__(256)
fa_bwd_prep_kernel(const T *o, int64_t o_sb, int64_t o_sh, int64_t o_sl, const T *dout, int64_t d_sb,
                   int64_t d_sh, int64_t d_sl, const float *lse, float *delta, float *lse2,
                   float *dq_acc, int B, int H, int Lq, int Lq_pad, int *sched_counter, int *dq_sem, int sem_count) {
  constexpr int kTpr = D / 8;  // threads per row
  griddep_launch_dependents();
  griddep_wait();
  if (blockIdx.x == 0 && threadIdx.x == 0) *sched_counter = 0;
  if (dq_sem)
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < sem_count;
         e += static_cast<int64_t>(gridDim.x) * blockDim.x)
      dq_sem[e] = 0;
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
    if (dq_acc) {  // nullptr: the caller owns the accumulator and keeps adding to it (ring hops)
      float4 *acc = reinterpret_cast<float4 *>(dq_acc + (static_cast<int64_t>(g) * Lq + r) * D + sub * 8);
      acc[0] = make_float4(0.f, 0.f, 0.f, 0.f);
      acc[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float l = lse[static_cast<int64_t>(g) * Lq + r];
    l2 = (l == -INFINITY) ? -INFINITY : -l * kLog2e;
  }
  #pragma unroll
  for (int off = kTpr / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if (in_range && sub == 0) { delta[rid] = -s; lse2[rid] = l2; }
}

}  // namespace fa
