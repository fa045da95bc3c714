// fa_generic.cuh -- SIMT attention kernels for the cases the tcgen05 path does not cover:
// fp32 inputs (the reference op accepts FLOAT, cpp/flash_attention_qkv_packed.cpp:609, and its
// tests use fp32 with 1e-5 / 1e-4 bounds, tests/cpp/test_flash_attention.cpp:245-259) and head
// dims other than 64/128 (e.g. the (G,L,D)=(4,6,2) cases).  Correctness path, not the headline:
// one warp per query row (fwd, dQ) or per key row (dK/dV), fp32 math, no score matrix in memory.
// Also holds the small helper kernels shared with the tensor path (delta, merge, convert).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {

template <typename T>
struct View {  // (B,H,L,D) operand with element strides
  T *ptr;
  int64_t sb, sh, sl;
  __device__ __forceinline__ T *row(int b, int h, int l) const {
    return ptr + b * sb + h * sh + static_cast<int64_t>(l) * sl;
  }
};

struct GenericParams {
  int B, H, Lq, Lk, D;
  int causal, shift;  // key c visible to query r iff c <= r + shift
  float scale;
};

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ __half from_f32<__half>(float x) { return __float2half_rn(x); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

__device__ __forceinline__ float warp_max(float v) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kGenWarps = 4;
constexpr int kGenMaxD = 256;  // each lane owns dims lane, lane+32, ... (<= 8 of them)

// ------------------------------------------------------------------ forward
template <typename T>
__global__ void __launch_bounds__(kGenWarps * 32)
fa_generic_fwd_kernel(View<const T> q, View<const T> k, View<const T> v, View<T> o, float *lse,
                      GenericParams p) {
  extern __shared__ float gsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * kGenWarps + warp;
  if (rid >= static_cast<int64_t>(p.B) * p.H * p.Lq) return;
  const int r = static_cast<int>(rid % p.Lq);
  const int g = static_cast<int>(rid / p.Lq);
  const int b = g / p.H, h = g % p.H;
  float *qs = gsm + warp * p.D;
  const T *qr = q.row(b, h, r);
  for (int d = lane; d < p.D; d += 32) qs[d] = to_f32(qr[d]) * p.scale;
  __syncwarp();

  int hi = p.Lk;
  if (p.causal) hi = min(hi, r + p.shift + 1);
  float acc[kGenMaxD / 32];
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) acc[t] = 0.f;
  float m = -INFINITY, l = 0.f;
  for (int c0 = 0; c0 < hi; c0 += 32) {
    const int c = c0 + lane;
    float s = -INFINITY;
    if (c < hi) {
      const T *kr = k.row(b, h, c);
      s = 0.f;
      for (int d = 0; d < p.D; ++d) s = fmaf(qs[d], to_f32(kr[d]), s);
    }
    const float m_new = fmaxf(m, warp_max(s));
    const float corr = __expf(m - m_new);
    const float pe = (c < hi) ? __expf(s - m_new) : 0.f;
    l = l * corr + warp_sum(pe);
    m = m_new;
    #pragma unroll
    for (int t = 0; t < kGenMaxD / 32; ++t) acc[t] *= corr;
    const int nc = min(32, hi - c0);
    for (int cc = 0; cc < nc; ++cc) {
      const float pc = __shfl_sync(0xffffffffu, pe, cc);
      const T *vr = v.row(b, h, c0 + cc);
      #pragma unroll
      for (int t = 0; t < kGenMaxD / 32; ++t) {
        const int d = lane + 32 * t;
        if (d < p.D) acc[t] = fmaf(pc, to_f32(vr[d]), acc[t]);
      }
    }
  }
  const float inv = l > 0.f ? 1.f / l : 0.f;
  T *orow = o.row(b, h, r);
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) {
    const int d = lane + 32 * t;
    if (d < p.D) orow[d] = from_f32<T>(acc[t] * inv);
  }
  if (lane == 0) lse[rid] = l > 0.f ? m + __logf(l) : -INFINITY;
}

// ------------------------------------------------------------------ delta = rowsum(O * dO)
// (cpp/flash_attention_qkv_packed.cpp:374-377).  One warp per row; used by both kernel families.
template <typename T>
__global__ void __launch_bounds__(128)
fa_delta_kernel(View<const T> o, View<const T> dout, float *delta, int B, int H, int Lq, int D) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * 4 + warp;
  if (rid >= static_cast<int64_t>(B) * H * Lq) return;
  const int r = static_cast<int>(rid % Lq);
  const int g = static_cast<int>(rid / Lq);
  const T *orow = o.row(g / H, g % H, r);
  const T *drow = dout.row(g / H, g % H, r);
  float s = 0.f;
  for (int d = lane; d < D; d += 32) s = fmaf(to_f32(orow[d]), to_f32(drow[d]), s);
  s = warp_sum(s);
  if (lane == 0) delta[rid] = s;
}

// ------------------------------------------------------------------ backward: dQ (warp per query row)
// P = exp(S - LSE) (:457-459); dP = dO V^T (:465); dS = P*(dP - D) (:466-467); dQ += dS K (:470)
template <typename T>
__global__ void __launch_bounds__(kGenWarps * 32)
fa_generic_bwd_dq_kernel(View<const T> q, View<const T> k, View<const T> v, View<const T> dout,
                         const float *lse, const float *delta, View<T> dq, GenericParams p) {
  extern __shared__ float gsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * kGenWarps + warp;
  if (rid >= static_cast<int64_t>(p.B) * p.H * p.Lq) return;
  const int r = static_cast<int>(rid % p.Lq);
  const int g = static_cast<int>(rid / p.Lq);
  const int b = g / p.H, h = g % p.H;
  float *qs = gsm + warp * 2 * p.D;
  float *ds_ = qs + p.D;
  const T *qr = q.row(b, h, r);
  const T *dr = dout.row(b, h, r);
  for (int d = lane; d < p.D; d += 32) { qs[d] = to_f32(qr[d]) * p.scale; ds_[d] = to_f32(dr[d]); }
  __syncwarp();
  const float lse_r = lse[rid], delta_r = delta[rid];
  int hi = p.Lk;
  if (p.causal) hi = min(hi, r + p.shift + 1);
  float acc[kGenMaxD / 32];
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) acc[t] = 0.f;
  for (int c0 = 0; c0 < hi; c0 += 32) {
    const int c = c0 + lane;
    float dsv = 0.f;
    if (c < hi) {
      const T *kr = k.row(b, h, c);
      const T *vr = v.row(b, h, c);
      float s = 0.f, dp = 0.f;
      for (int d = 0; d < p.D; ++d) {
        s = fmaf(qs[d], to_f32(kr[d]), s);
        dp = fmaf(ds_[d], to_f32(vr[d]), dp);
      }
      dsv = __expf(s - lse_r) * (dp - delta_r);
    }
    const int nc = min(32, hi - c0);
    for (int cc = 0; cc < nc; ++cc) {
      const float e = __shfl_sync(0xffffffffu, dsv, cc);
      const T *kr = k.row(b, h, c0 + cc);
      #pragma unroll
      for (int t = 0; t < kGenMaxD / 32; ++t) {
        const int d = lane + 32 * t;
        if (d < p.D) acc[t] = fmaf(e, to_f32(kr[d]), acc[t]);
      }
    }
  }
  T *orow = dq.row(b, h, r);
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) {
    const int d = lane + 32 * t;
    if (d < p.D) orow[d] = from_f32<T>(acc[t] * p.scale);
  }
}

// ------------------------------------------------------------------ backward: dK, dV (warp per key row)
// dV += P^T dO (:462); dK += dS^T Q (:472)
template <typename T>
__global__ void __launch_bounds__(kGenWarps * 32)
fa_generic_bwd_dkv_kernel(View<const T> q, View<const T> k, View<const T> v, View<const T> dout,
                          const float *lse, const float *delta, View<T> dk, View<T> dv,
                          GenericParams p) {
  extern __shared__ float gsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t cid = static_cast<int64_t>(blockIdx.x) * kGenWarps + warp;
  if (cid >= static_cast<int64_t>(p.B) * p.H * p.Lk) return;
  const int c = static_cast<int>(cid % p.Lk);
  const int g = static_cast<int>(cid / p.Lk);
  const int b = g / p.H, h = g % p.H;
  float *ks = gsm + warp * 2 * p.D;
  float *vs = ks + p.D;
  const T *kr = k.row(b, h, c);
  const T *vr = v.row(b, h, c);
  for (int d = lane; d < p.D; d += 32) { ks[d] = to_f32(kr[d]) * p.scale; vs[d] = to_f32(vr[d]); }
  __syncwarp();
  int lo = 0;  // first query row that sees key c
  if (p.causal) lo = max(0, c - p.shift);
  float acck[kGenMaxD / 32], accv[kGenMaxD / 32];
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) { acck[t] = 0.f; accv[t] = 0.f; }
  for (int r0 = lo; r0 < p.Lq; r0 += 32) {
    const int r = r0 + lane;
    float pe = 0.f, dsv = 0.f;
    if (r < p.Lq) {
      const T *qr = q.row(b, h, r);
      const T *dr = dout.row(b, h, r);
      float s = 0.f, dp = 0.f;
      for (int d = 0; d < p.D; ++d) {
        s = fmaf(ks[d], to_f32(qr[d]), s);
        dp = fmaf(vs[d], to_f32(dr[d]), dp);
      }
      const int64_t rid = static_cast<int64_t>(g) * p.Lq + r;
      pe = __expf(s - lse[rid]);
      dsv = pe * (dp - delta[rid]);
    }
    const int nr = min(32, p.Lq - r0);
    for (int rr = 0; rr < nr; ++rr) {
      const float pr = __shfl_sync(0xffffffffu, pe, rr);
      const float er = __shfl_sync(0xffffffffu, dsv, rr);
      const T *qr = q.row(b, h, r0 + rr);
      const T *dr = dout.row(b, h, r0 + rr);
      #pragma unroll
      for (int t = 0; t < kGenMaxD / 32; ++t) {
        const int d = lane + 32 * t;
        if (d < p.D) {
          accv[t] = fmaf(pr, to_f32(dr[d]), accv[t]);
          acck[t] = fmaf(er, to_f32(qr[d]), acck[t]);
        }
      }
    }
  }
  T *dkr = dk.row(b, h, c);
  T *dvr = dv.row(b, h, c);
  #pragma unroll
  for (int t = 0; t < kGenMaxD / 32; ++t) {
    const int d = lane + 32 * t;
    if (d < p.D) { dkr[d] = from_f32<T>(acck[t] * p.scale); dvr[d] = from_f32<T>(accv[t]); }
  }
}

// ------------------------------------------------------------------ LSE merge (ring hop)
template <typename T>
__global__ void __launch_bounds__(128)
fa_merge_kernel(T *o_a, float *lse_a, const T *o_b, const float *lse_b, int64_t rows, int D) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * 4 + warp;
  if (rid >= rows) return;
  const float la = lse_a[rid], lb = lse_b[rid];
  const float mx = fmaxf(la, lb);
  float wa = 0.f, wb = 0.f, lse = -INFINITY;
  if (mx > -INFINITY) {
    const float ea = __expf(la - mx), eb = __expf(lb - mx);
    const float inv = 1.f / (ea + eb);
    wa = ea * inv; wb = eb * inv;
    lse = mx + __logf(ea + eb);
  }
  T *ra = o_a + rid * D;
  const T *rb = o_b + rid * D;
  for (int d = lane; d < D; d += 32) ra[d] = from_f32<T>(to_f32(ra[d]) * wa + to_f32(rb[d]) * wb);
  __syncwarp();
  if (lane == 0) lse_a[rid] = lse;
}

// Ring-attention accumulator form: the running (o_acc, lse_acc) stay in fp32 for the whole ring
// (no per-hop rounding) and a hop's partial (o_b in the kernel dtype, lse_b) is folded in.  The
// partial may cover a sub-range of each group's rows: partial row (g, r) lands on accumulator row
// g * acc_outer_stride + acc_row0 + r.  One warp per row, D/32 elements per lane.
template <typename T>
__global__ void __launch_bounds__(128)
fa_merge_acc_kernel(float *o_acc, float *lse_acc, const T *o_b, const float *lse_b, int64_t n_outer,
                    int64_t n_inner, int64_t acc_outer_stride, int64_t acc_row0, int D) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t rid = static_cast<int64_t>(blockIdx.x) * 4 + warp;
  if (rid >= n_outer * n_inner) return;
  const int64_t arow = (rid / n_inner) * acc_outer_stride + acc_row0 + rid % n_inner;
  const float la = lse_acc[arow], lb = lse_b[rid];
  const float mx = fmaxf(la, lb);
  float wa = 0.f, wb = 0.f, lse = -INFINITY;
  if (mx > -INFINITY) {
    const float ea = __expf(la - mx), eb = __expf(lb - mx);
    const float inv = 1.f / (ea + eb);
    wa = ea * inv; wb = eb * inv;
    lse = mx + __logf(ea + eb);
  }
  float *ra = o_acc + arow * D;
  const T *rb = o_b + rid * D;
  for (int d = lane; d < D; d += 32) ra[d] = (wa == 0.f ? 0.f : ra[d] * wa) + to_f32(rb[d]) * wb;
  __syncwarp();
  if (lane == 0) lse_acc[arow] = lse;
}

// ------------------------------------------------------------------ fp32 accumulator -> output dtype
// dQ is accumulated in fp32 by TMA reduce-add in the tensor-path backward; this casts (and scales).
template <typename T>
__global__ void __launch_bounds__(256)
fa_convert_kernel(const float *src, View<T> dst, int B, int H, int L, int D, float scale) {
  static_assert(sizeof(T) == 2, "16-bit outputs only (8 elements = one 16-byte store)");
  // programmatic dependent launch (see sm100_ptx.cuh): no-ops when launched the plain way
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int64_t n8 = static_cast<int64_t>(B) * H * L * D / 8;  // D is 64 or 128 on this path
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float4 x = __ldcs(reinterpret_cast<const float4 *>(src) + 2 * i);
    const float4 y = __ldcs(reinterpret_cast<const float4 *>(src) + 2 * i + 1);
    const int64_t e = i * 8;
    const int d = static_cast<int>(e % D);
    const int64_t rowid = e / D;
    const int l = static_cast<int>(rowid % L);
    const int g = static_cast<int>(rowid / L);
    alignas(16) T out[8];
    out[0] = from_f32<T>(x.x * scale); out[1] = from_f32<T>(x.y * scale);
    out[2] = from_f32<T>(x.z * scale); out[3] = from_f32<T>(x.w * scale);
    out[4] = from_f32<T>(y.x * scale); out[5] = from_f32<T>(y.y * scale);
    out[6] = from_f32<T>(y.z * scale); out[7] = from_f32<T>(y.w * scale);
    *reinterpret_cast<uint4 *>(dst.row(g / H, g % H, l) + d) = *reinterpret_cast<const uint4 *>(out);
  }
}

}  // namespace fa
