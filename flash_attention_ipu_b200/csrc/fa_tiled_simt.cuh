// fa_tiled_simt.cuh -- shared-memory tiled fp32 attention on the CUDA cores.
//
// The reference op takes FLOAT as well as HALF (cpp/flash_attention_qkv_packed.cpp:609), its own
// tests run fp32 with 1e-5 / 1e-4 bounds (tests/cpp/test_flash_attention.cpp:245-259) and its
// nanoGPT demo trains in fp32 (BASELINE.json configs[0]).  tcgen05 has no full-precision fp32
// form, so fp32 inputs -- and 16-bit inputs whose head dim is not 64/128 -- run here: the same
// algorithm as the tensor-path kernels (forward :254-300, backward :457-472), as register-blocked
// FFMA GEMMs over 64 x 64 tiles staged in shared memory.
//
//   * 128 threads = 8 (ty) x 16 (tx); each thread owns an 8 x 4 block of the 64 x 64 score tile:
//     rows ty + 8 i, columns tx + 16 j.  The interleaved ownership makes every float4 shared
//     load conflict-free (row stride DP + 4 floats = an odd number of 16-byte units).
//   * "NT" product  C[r][c] = sum_d A[r][d] B[c][d]   (S = Q K^T, dP = dO V^T)
//     "NN" product  O[r][:] += sum_c P[r][c] Y[c][:]  (O += P V, dQ += dS K, dV += P^T dO, ...)
//     P goes through shared memory between the two; its producers and consumers are the same 16
//     lanes (one ty), so a __syncwarp() is all the hand-off needs.
//   * head dims below the padded width DP (64 or 128) are zero-filled in shared memory.
//   * backward = two kernels (dQ with Q stationary, dK/dV with K/V stationary and the score tile
//     transposed), no atomics: results are deterministic.
//   * the inner products are issued as packed fma.rn.f32x2 (SASS FFMA2): a 3-register FFMA
//     issues every other cycle per SM sub-partition, FFMA2 does two FMAs in the same slot, so it
//     is the only way to the fp32 peak.  NT packs along the contraction (even/odd partial sums,
//     the float4 operand loads are already register pairs); NN packs along the output columns
//     with the P value duplicated into both halves.
// Exponentials are exp2; softmax_scale * log2(e) is folded into the FFMA that forms the exponent, so
// every operand tile is staged unscaled (and, for fp32 inputs, by cp.async).
#pragma once
#include "fa_generic.cuh"
#include "sm100_ptx.cuh"

namespace fa {

constexpr int kTB = 64;          // tile edge: stationary rows and streamed rows per step
constexpr int kTThreads = 128;   // 8 x 16 thread grid
constexpr int kTLdp = kTB + 4;   // row stride of the P tile in shared memory (floats)
constexpr float kTLog2e = 1.4426950408889634f;
constexpr float kTLn2 = 0.6931471805599453f;

__device__ __forceinline__ float t_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float t_max16(float v) {  // over the 16 lanes that share a ty
  #pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float t_sum16(float v) {
  #pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// rows [row0, row0 + 64) of one (batch, head) slab -> dst[64][DP + 4], times `mul`, zero-filled
// past `nrows` and past D.  Consecutive threads read consecutive elements of a row.
template <typename T, int DP>
__device__ __forceinline__ void t_load_tile(float *dst, const T *slab, int64_t sl, int row0, int nrows,
                                            int D, float mul) {
  constexpr int LD = DP + 4;
  #pragma unroll 8
  for (int idx = threadIdx.x; idx < kTB * DP; idx += kTThreads) {
    const int r = idx / DP, d = idx % DP;
    const int gr = row0 + r;
    float x = 0.f;
    if (gr < nrows && d < D) x = to_f32(slab[gr * sl + d]) * mul;
    dst[r * LD + d] = x;
  }
}

__device__ __forceinline__ void t_cp_async_16(float *smem_dst, const float *gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void t_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void t_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

// A streamed tile (unscaled).  kAsync (fp32 rows, 16-byte aligned): cp.async straight into shared
// memory, one commit group per tile -- the copy lands while the caller runs a product on the
// other operand; otherwise the synchronous converting load.  The kernels' barrier structure is
// written for the asynchronous form and is also correct for the synchronous one.
template <typename T, int DP, bool kAsync>
__device__ __forceinline__ void t_stream_tile(float *dst, const T *slab, int64_t sl, int row0, int nrows, int D) {
  if constexpr (kAsync) {
    static_assert(sizeof(T) == 4, "cp.async staging is for fp32 inputs");
    constexpr int LD = DP + 4, V4 = DP / 4;
    #pragma unroll 4
    for (int idx = threadIdx.x; idx < kTB * V4; idx += kTThreads) {
      const int r = idx / V4, d = (idx % V4) * 4;
      float *to = dst + r * LD + d;
      const int gr = row0 + r;
      if (gr < nrows && d < D) t_cp_async_16(to, reinterpret_cast<const float *>(slab) + gr * sl + d);
      else *reinterpret_cast<float4 *>(to) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    t_cp_async_commit();
  } else {
    t_load_tile<T, DP>(dst, slab, sl, row0, nrows, D, 1.f);
  }
}
template <bool kAsync, int kPending>
__device__ __forceinline__ void t_stream_wait() {
  if constexpr (kAsync) t_cp_async_wait<kPending>();
}

// c[i][j] += sum_d A[ty + 8 i][d] * B[tx + 16 j][d].  kPacked: FFMA2 over (even d, odd d) partial
// sums (64 transient accumulator registers); otherwise scalar FFMA (32), for the register-bound
// dK/dV kernel at DP = 128.
template <int DP, bool kPacked>
__device__ __forceinline__ void t_gemm_nt(const float *A, const float *B, int ty, int tx, float (&c)[8][4]) {
  constexpr int LD = DP + 4;
  if constexpr (kPacked) {
    uint64_t c2[8][4];
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) c2[i][j] = 0ull;
    }
    #pragma unroll 2
    for (int d = 0; d < DP; d += 4) {
      ulonglong2 a[8], b[4];
      #pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const ulonglong2 *>(A + (ty + 8 * i) * LD + d);
      #pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const ulonglong2 *>(B + (tx + 16 * j) * LD + d);
      #pragma unroll
      for (int i = 0; i < 8; ++i) {
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
          c2[i][j] = f2_fma(a[i].x, b[j].x, c2[i][j]);
          c2[i][j] = f2_fma(a[i].y, b[j].y, c2[i][j]);
        }
      }
    }
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo, hi;
        f2_unpack(c2[i][j], lo, hi);
        c[i][j] += lo + hi;
      }
    }
  } else {
    #pragma unroll 2
    for (int d = 0; d < DP; d += 4) {
      float4 a[8], b[4];
      #pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4 *>(A + (ty + 8 * i) * LD + d);
      #pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4 *>(B + (tx + 16 * j) * LD + d);
      #pragma unroll
      for (int i = 0; i < 8; ++i) {
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
          c[i][j] = fmaf(a[i].x, b[j].x, c[i][j]);
          c[i][j] = fmaf(a[i].y, b[j].y, c[i][j]);
          c[i][j] = fmaf(a[i].z, b[j].z, c[i][j]);
          c[i][j] = fmaf(a[i].w, b[j].w, c[i][j]);
        }
      }
    }
  }
}

// o[i][:] += sum_c P[ty + 8 i][c] * Y[c][:]; the thread owns columns tx*4 .. tx*4+3 (+64 for
// DP=128), held as packed pairs: o[i][2 jj] = columns (0,1), o[i][2 jj + 1] = columns (2,3).
template <int DP>
__device__ __forceinline__ void t_gemm_nn(const float *P, const float *Y, int ty, int tx,
                                          uint64_t (&o)[8][DP / 32]) {
  constexpr int LD = DP + 4, NJ = DP / 64;
  #pragma unroll 1
  for (int c = 0; c < kTB; c += 4) {
    float4 pr[8];
    #pragma unroll
    for (int i = 0; i < 8; ++i) pr[i] = *reinterpret_cast<const float4 *>(P + (ty + 8 * i) * kTLdp + c);
    #pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      ulonglong2 y[NJ];
      #pragma unroll
      for (int jj = 0; jj < NJ; ++jj)
        y[jj] = *reinterpret_cast<const ulonglong2 *>(Y + (c + cc) * LD + tx * 4 + 64 * jj);
      #pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float pv = cc == 0 ? pr[i].x : cc == 1 ? pr[i].y : cc == 2 ? pr[i].z : pr[i].w;
        const uint64_t pp = f2_pack(pv, pv);
        #pragma unroll
        for (int jj = 0; jj < NJ; ++jj) {
          o[i][jj * 2 + 0] = f2_fma(pp, y[jj].x, o[i][jj * 2 + 0]);
          o[i][jj * 2 + 1] = f2_fma(pp, y[jj].y, o[i][jj * 2 + 1]);
        }
      }
    }
  }
}

template <int DP, typename T>
__device__ __forceinline__ void t_store_rows(T *slab, int64_t sl, int row0, int nrows, int D, int ty, int tx,
                                             const uint64_t (&acc)[8][DP / 32], const float (&mul)[8]) {
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = row0 + ty + 8 * i;
    if (r >= nrows) continue;
    T *row = slab + r * sl;
    #pragma unroll
    for (int e2 = 0; e2 < DP / 32; ++e2) {
      float x[2];
      f2_unpack(acc[i][e2], x[0], x[1]);
      #pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int d = tx * 4 + 64 * (e2 / 2) + 2 * (e2 % 2) + e;
        if (d < D) row[d] = from_f32<T>(x[e] * mul[i]);
      }
    }
  }
}

template <int DP> constexpr size_t t_fwd_smem() { return (3 * kTB * (DP + 4) + kTB * kTLdp) * sizeof(float); }
template <int DP> constexpr size_t t_bwd_smem() { return (4 * kTB * (DP + 4) + kTB * kTLdp + 2 * kTB) * sizeof(float); }

// ------------------------------------------------------------------ forward
// grid (query tiles, B*H).  Q tile stationary, K/V streamed.
template <typename T, int DP, bool kAsync>
__global__ void __launch_bounds__(kTThreads)
fa_tiled_fwd_kernel(View<const T> q, View<const T> k, View<const T> v, View<T> o, float *lse,
                    GenericParams p) {
  extern __shared__ __align__(16) float tsm[];
  constexpr int LD = DP + 4, NC = DP / 32;
  float *Qs = tsm, *Ks = Qs + kTB * LD, *Vs = Ks + kTB * LD, *Ps = Vs + kTB * LD;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  // 1-D grid (tiles x groups folded: gridDim.y stops at 65 535 groups), tile index fastest
  const int nt = (p.Lq + kTB - 1) / kTB;
  const int g = blockIdx.x / nt, b = g / p.H, h = g % p.H;
  const int m0 = (nt - 1 - static_cast<int>(blockIdx.x % nt)) * kTB;  // longest (latest) causal rows first
  // Q goes to shared memory unscaled (same asynchronous copy as the streamed tiles, in flight
  // together with K(0)); softmax_scale * log2(e) is folded into the exponent's FFMA instead.
  t_stream_tile<T, DP, kAsync>(Qs, q.row(b, h, 0), q.sl, m0, p.Lq, p.D);
  const float c2 = p.scale * kTLog2e;
  int hi = p.Lk;
  if (p.causal) hi = min(hi, m0 + kTB + p.shift);  // last row of the tile sees keys < m0+63+shift+1

  uint64_t acc[8][NC];
  float m[8], l[8];
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    m[i] = -INFINITY; l[i] = 0.f;
    #pragma unroll
    for (int c = 0; c < NC; ++c) acc[i][c] = 0ull;
  }
  const T *kslab = k.row(b, h, 0), *vslab = v.row(b, h, 0);
  // K(n) is fetched while P(n-1) V(n-1) runs, V(n) while Q K(n)^T runs.
  if (hi > 0) t_stream_tile<T, DP, kAsync>(Ks, kslab, k.sl, 0, p.Lk, p.D);
  for (int n0 = 0; n0 < hi; n0 += kTB) {
    t_stream_wait<kAsync, 0>();
    __syncthreads();  // K(n0) has landed; everyone is done with the previous V tile
    t_stream_tile<T, DP, kAsync>(Vs, vslab, v.sl, n0, p.Lk, p.D);
    float s[8][4];
    #pragma unroll
    for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
    t_gemm_nt<DP, true>(Qs, Ks, ty, tx, s);
    __syncthreads();  // everyone is done with K(n0)
    const bool has_next = n0 + kTB < hi;
    if (has_next) t_stream_tile<T, DP, kAsync>(Ks, kslab, k.sl, n0 + kTB, p.Lk, p.D);
    const bool edge = (n0 + kTB > p.Lk) || (p.causal && n0 + kTB - 1 > m0 + p.shift);
    if (edge) {
      #pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = m0 + ty + 8 * i;
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = n0 + tx + 16 * j;
          if (c >= p.Lk || (p.causal && c > r + p.shift)) s[i][j] = -INFINITY;
        }
      }
    }
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float mx = t_max16(fmaxf(fmaxf(s[i][0], s[i][1]), fmaxf(s[i][2], s[i][3])));
      const float m_new = fmaxf(m[i], mx);
      const float m_use = m_new == -INFINITY ? 0.f : m_new;  // a row with nothing visible yet
      const float corr = t_exp2((m[i] - m_use) * c2);
      const float nm = -m_use * c2;
      float part = 0.f;
      #pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float e = t_exp2(fmaf(s[i][j], c2, nm));
        part += e;
        Ps[(ty + 8 * i) * kTLdp + tx + 16 * j] = e;
      }
      l[i] = l[i] * corr + part;  // per-thread partial sum; reduced over the 16 lanes at the end
      m[i] = m_new;
      const uint64_t cc2 = f2_pack(corr, corr);
      #pragma unroll
      for (int c = 0; c < NC; ++c) acc[i][c] = f2_mul(acc[i][c], cc2);
    }
    if (has_next) t_stream_wait<kAsync, 1>(); else t_stream_wait<kAsync, 0>();  // V(n0); K(next) may be in flight
    __syncthreads();
    t_gemm_nn<DP>(Ps, Vs, ty, tx, acc);
  }
  t_stream_wait<kAsync, 0>();  // (a tile with no visible key still issued its Q copy)
  float inv[8];
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float lt = t_sum16(l[i]);
    inv[i] = lt > 0.f ? 1.f / lt : 0.f;
    const int r = m0 + ty + 8 * i;
    if (tx == 0 && r < p.Lq)
      lse[static_cast<int64_t>(g) * p.Lq + r] = lt > 0.f ? (m[i] * c2 + log2f(lt)) * kTLn2 : -INFINITY;
  }
  t_store_rows<DP, T>(o.row(b, h, 0), o.sl, m0, p.Lq, p.D, ty, tx, acc, inv);
}

// ------------------------------------------------------------------ backward: dQ
// grid (query tiles, B*H).  Q and dO stationary, K/V streamed.
// P = exp(S - LSE) (:457-459); dP = dO V^T (:465); dS = P (dP - delta) (:466-467); dQ += dS K (:470)
template <typename T, int DP, bool kAsync>
__global__ void __launch_bounds__(kTThreads)
fa_tiled_bwd_dq_kernel(View<const T> q, View<const T> k, View<const T> v, View<const T> o,
                       View<const T> dout, const float *lse, float *delta, View<T> dq, GenericParams p) {
  extern __shared__ __align__(16) float tsm[];
  constexpr int LD = DP + 4, NC = DP / 32;
  float *Qs = tsm, *Ds = Qs + kTB * LD, *Ks = Ds + kTB * LD, *Vs = Ks + kTB * LD, *Ps = Vs + kTB * LD;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int nt = (p.Lq + kTB - 1) / kTB;
  const int g = blockIdx.x / nt, b = g / p.H, h = g % p.H;
  const int m0 = (nt - 1 - static_cast<int>(blockIdx.x % nt)) * kTB;
  int hi = p.Lk;
  if (p.causal) hi = min(hi, m0 + kTB + p.shift);
  const T *kslab = k.row(b, h, 0), *vslab = v.row(b, h, 0);
  // Prologue copies, all in flight at once: Q and dO (stationary, unscaled -- softmax_scale *
  // log2(e) is folded into the exponent's FFMA), the O tile (parked in the K buffer, only needed
  // for delta) and V(0).
  t_stream_tile<T, DP, kAsync>(Qs, q.row(b, h, 0), q.sl, m0, p.Lq, p.D);
  t_stream_tile<T, DP, kAsync>(Ds, dout.row(b, h, 0), dout.sl, m0, p.Lq, p.D);
  t_stream_tile<T, DP, kAsync>(Ks, o.row(b, h, 0), o.sl, m0, p.Lq, p.D);
  if (hi > 0) t_stream_tile<T, DP, kAsync>(Vs, vslab, v.sl, 0, p.Lk, p.D);
  const float c2 = p.scale * kTLog2e;

  uint64_t acc[8][NC];
  float lse2[8], dl[8];
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = m0 + ty + 8 * i;
    float ls = INFINITY;  // exp2(s - inf) = 0: rows that saw no key, rows past the end
    if (r < p.Lq) {
      const float x = lse[static_cast<int64_t>(g) * p.Lq + r];
      if (x > -INFINITY) ls = x * kTLog2e;
    }
    lse2[i] = ls;
    #pragma unroll
    for (int c = 0; c < NC; ++c) acc[i][c] = 0ull;
  }
  t_stream_wait<kAsync, 0>();
  __syncthreads();
  // delta = rowsum(O o dO) (:374-377), 16 lanes per row from the two staged tiles; written out for
  // the dK/dV kernel that follows on the stream -- no separate pass over O and dO
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = m0 + ty + 8 * i;
    float part = 0.f;
    #pragma unroll
    for (int jj = 0; jj < DP / 64; ++jj) {
      const float4 ov = *reinterpret_cast<const float4 *>(Ks + (ty + 8 * i) * LD + tx * 4 + 64 * jj);
      const float4 dv4 = *reinterpret_cast<const float4 *>(Ds + (ty + 8 * i) * LD + tx * 4 + 64 * jj);
      part = fmaf(ov.x, dv4.x, fmaf(ov.y, dv4.y, fmaf(ov.z, dv4.z, fmaf(ov.w, dv4.w, part))));
    }
    dl[i] = t_sum16(part);
    if (tx == 0 && r < p.Lq) delta[static_cast<int64_t>(g) * p.Lq + r] = dl[i];
  }
  // V(n) is fetched while Q K(n-1)^T and dS K(n-1) run, K(n) while dO V(n)^T runs.
  for (int n0 = 0; n0 < hi; n0 += kTB) {
    t_stream_wait<kAsync, 0>();
    __syncthreads();  // V(n0) has landed; everyone is done with the previous K tile (or the O tile)
    t_stream_tile<T, DP, kAsync>(Ks, kslab, k.sl, n0, p.Lk, p.D);
    float s[8][4], dp[8][4];
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    }
    t_gemm_nt<DP, true>(Ds, Vs, ty, tx, dp);
    t_stream_wait<kAsync, 0>();
    __syncthreads();  // K(n0) has landed; everyone is done with V(n0)
    if (n0 + kTB < hi) t_stream_tile<T, DP, kAsync>(Vs, vslab, v.sl, n0 + kTB, p.Lk, p.D);
    t_gemm_nt<DP, true>(Qs, Ks, ty, tx, s);
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = m0 + ty + 8 * i;
      #pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = n0 + tx + 16 * j;
        const bool hidden = c >= p.Lk || (p.causal && c > r + p.shift);
        const float pe = hidden ? 0.f : t_exp2(fmaf(s[i][j], c2, -lse2[i]));
        Ps[(ty + 8 * i) * kTLdp + tx + 16 * j] = pe * (dp[i][j] - dl[i]);
      }
    }
    __syncwarp();
    t_gemm_nn<DP>(Ps, Ks, ty, tx, acc);
  }
  float mul[8];
  #pragma unroll
  for (int i = 0; i < 8; ++i) mul[i] = p.scale;
  t_store_rows<DP, T>(dq.row(b, h, 0), dq.sl, m0, p.Lq, p.D, ty, tx, acc, mul);
}

// ------------------------------------------------------------------ backward: dK, dV
// grid (key tiles, B*H).  K and V stationary, Q/dO streamed; the score tile is held
// transposed (rows = keys) so that dV += P^T dO (:462) and dK += dS^T Q (:472) are NN products.
template <typename T, int DP, bool kAsync>
__global__ void __launch_bounds__(kTThreads)
fa_tiled_bwd_dkv_kernel(View<const T> q, View<const T> k, View<const T> v, View<const T> dout,
                        const float *lse, const float *delta, View<T> dk, View<T> dv,
                        GenericParams p) {
  extern __shared__ __align__(16) float tsm[];
  constexpr int LD = DP + 4, NC = DP / 32;
  float *Ks = tsm, *Vs = Ks + kTB * LD, *Qs = Vs + kTB * LD, *Ds = Qs + kTB * LD, *Ps = Ds + kTB * LD;
  float *lse2s = Ps + kTB * kTLdp, *dls = lse2s + kTB;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int nt = (p.Lk + kTB - 1) / kTB;
  const int g = blockIdx.x / nt, b = g / p.H, h = g % p.H;
  const int n0 = static_cast<int>(blockIdx.x % nt) * kTB;  // earliest keys have the most causal work
  // stationary K, V: unscaled, copied asynchronously together with the first dO tile
  t_stream_tile<T, DP, kAsync>(Ks, k.row(b, h, 0), k.sl, n0, p.Lk, p.D);
  t_stream_tile<T, DP, kAsync>(Vs, v.row(b, h, 0), v.sl, n0, p.Lk, p.D);
  const float c2 = p.scale * kTLog2e;
  int m_start = 0;  // first query tile with a row that sees key n0+63... or any key of the tile
  if (p.causal) m_start = (max(0, n0 - p.shift) / kTB) * kTB;

  uint64_t acck[8][NC], accv[8][NC];
  #pragma unroll
  for (int i = 0; i < 8; ++i) {
    #pragma unroll
    for (int c = 0; c < NC; ++c) { acck[i][c] = 0ull; accv[i][c] = 0ull; }
  }
  const T *qslab = q.row(b, h, 0), *dslab = dout.row(b, h, 0);
  // dO(m) is fetched while dS^T Q(m-1) runs, Q(m) while V dO(m)^T runs.
  if (m_start < p.Lq) t_stream_tile<T, DP, kAsync>(Ds, dslab, dout.sl, m_start, p.Lq, p.D);
  for (int m0 = m_start; m0 < p.Lq; m0 += kTB) {
    t_stream_wait<kAsync, 0>();
    __syncthreads();  // dO(m0) has landed; everyone is done with the previous Q tile and its stats
    t_stream_tile<T, DP, kAsync>(Qs, qslab, q.sl, m0, p.Lq, p.D);
    if (tid < kTB) {
      const int r = m0 + tid;
      float ls = INFINITY, dl = 0.f;
      if (r < p.Lq) {
        const float x = lse[static_cast<int64_t>(g) * p.Lq + r];
        if (x > -INFINITY) ls = x * kTLog2e;
        dl = delta[static_cast<int64_t>(g) * p.Lq + r];
      }
      lse2s[tid] = ls; dls[tid] = dl;
    }
    float s[8][4], dp[8][4];
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    }
    t_gemm_nt<DP, (DP < 128)>(Vs, Ds, ty, tx, dp);
    t_stream_wait<kAsync, 0>();
    __syncthreads();  // Q(m0) and the row statistics are visible
    t_gemm_nt<DP, (DP < 128)>(Ks, Qs, ty, tx, s);    // s[i][j] = K[n0+ty+8i] . Q[m0+tx+16j]
    #pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = m0 + tx + 16 * j;
      const float ls = lse2s[tx + 16 * j], dl = dls[tx + 16 * j];
      #pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = n0 + ty + 8 * i;
        const bool hidden = c >= p.Lk || r >= p.Lq || (p.causal && c > r + p.shift);
        const float pe = hidden ? 0.f : t_exp2(fmaf(s[i][j], c2, -ls));
        s[i][j] = pe;
        dp[i][j] = pe * (dp[i][j] - dl);
      }
    }
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) Ps[(ty + 8 * i) * kTLdp + tx + 16 * j] = s[i][j];
    }
    __syncwarp();
    t_gemm_nn<DP>(Ps, Ds, ty, tx, accv);
    __syncthreads();  // everyone is done with dO(m0) (and with P^T in Ps)
    if (m0 + kTB < p.Lq) t_stream_tile<T, DP, kAsync>(Ds, dslab, dout.sl, m0 + kTB, p.Lq, p.D);
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
      #pragma unroll
      for (int j = 0; j < 4; ++j) Ps[(ty + 8 * i) * kTLdp + tx + 16 * j] = dp[i][j];
    }
    __syncwarp();
    t_gemm_nn<DP>(Ps, Qs, ty, tx, acck);
  }
  t_stream_wait<kAsync, 0>();  // (a key tile no query row sees still issued its K, V copies)
  float mulk[8], mulv[8];
  #pragma unroll
  for (int i = 0; i < 8; ++i) { mulk[i] = p.scale; mulv[i] = 1.f; }
  t_store_rows<DP, T>(dk.row(b, h, 0), dk.sl, n0, p.Lk, p.D, ty, tx, acck, mulk);
  t_store_rows<DP, T>(dv.row(b, h, 0), dv.sl, n0, p.Lk, p.D, ty, tx, accv, mulv);
}

}  // namespace fa
