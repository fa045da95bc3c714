// umma_probe.cu -- single-CTA probes of the tcgen05 / TMA building blocks the attention
// kernels rely on.  Integer-valued bf16 inputs make every result exact, so any mismatch is a
// layout / descriptor error, not rounding.  Usage: umma_probe <mode>
//   0: D[128x128] = A[128xK=128] * B[128xK=128]^T      SS, A K-major, B K-major (S = Q K^T)
//   1: D[128x128] = A[128xK=128] * V[K=128 x 128]      SS, A K-major, B MN-major (V as stored)
//   2: D[128x128] = P[128xK=128] * V[K=128 x 128]      TS, A = P written to TMEM by tcgen05.st
//   3: D[128x128] = A^T ... A MN-major: D = At[K=128 x 128]^T * V[K x 128]  (dS^T-style operand)
//   4: fp32 TMA reduce-add of a smem tile into global (dQ accumulation path)
//   5/6: D[128x128] = A[128 x K=16] * B[128 x K=16]^T with both operands in the NO-SWIZZLE K-major
//        canonical layout (8-row x 16-byte core matrices), written by st.shared: which descriptor field
//        is the K-direction core stride?  5: LBO = K stride (128), SBO = 8-row-group stride (256); 6: swapped
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "../../flash_attention_ipu_b200/csrc/sm100_ptx.cuh"

using namespace fa;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(void *ptr, int rows, int cols, bool f32) {
  void *sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q));
  CUtensorMap m;
  size_t es = f32 ? 4 : 2;
  cuuint64_t dims[4] = {(cuuint64_t)cols, (cuuint64_t)rows, 1, 1};
  cuuint64_t strides[3] = {cols * es, rows * cols * es, rows * cols * es};
  cuuint32_t box[4] = {(cuuint32_t)(128 / es), 128, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = ((EncodeTiledFn)sym)(&m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                                    4, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(2); }
  return m;
}

// smem: A at 0 (2 halves x 16 KB), B at 32 KB (2 halves x 16 KB)
__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __nv_bfloat16 *a_gmem, float *d_out, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bar_ld = (uint64_t *)(smem + 65536);
  uint64_t *bar_mma = bar_ld + 1;
  uint32_t *tptr = (uint32_t *)(bar_ld + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(bar_ld, 1); mbar_init(bar_mma, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc<256>(tptr);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = *tptr;
  const uint32_t tD = tb, tP = tb + 128;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_ld, 65536);
    for (int hh = 0; hh < 2; ++hh) {
      tma_load_4d(&tmA, bar_ld, smem + hh * 16384, hh * 64, 0, 0, 0);
      tma_load_4d(&tmB, bar_ld, smem + 32768 + hh * 16384, hh * 64, 0, 0, 0);
    }
  }
  mbar_wait(bar_ld, 0);
  if (mode == 2) {  // P = A (row-major [128][128] bf16 in global) -> TMEM, packed 2 per column
    const int row = warp * 32 + lane;
    const uint32_t *src = (const uint32_t *)(a_gmem + row * 128);
    for (int c = 0; c < 4; ++c) {
      uint32_t w[16];
      for (int e = 0; e < 16; ++e) w[e] = src[c * 16 + e];
      tmem_st16(tP + ((uint32_t)(warp * 32) << 16) + c * 16, w);
    }
    tmem_st_wait();
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
    for (int k = 0; k < 8; ++k) {
      if (mode == 0) {
        const uint32_t off = (k / 4) * 16384 + (k % 4) * 32;
        umma_ss(tD, make_smem_desc(a_addr + off, 16, 1024), make_smem_desc(b_addr + off, 16, 1024),
                make_idesc(128, 128, 1, 0, 0), k > 0);
      } else if (mode == 1) {
        const uint32_t off = (k / 4) * 16384 + (k % 4) * 32;
        umma_ss(tD, make_smem_desc(a_addr + off, 16, 1024),
                make_smem_desc(b_addr + k * 2048, 16384, 1024), make_idesc(128, 128, 1, 0, 1), k > 0);
      } else if (mode == 2) {
        umma_ts(tD, tP + k * 8, make_smem_desc(b_addr + k * 2048, 16384, 1024),
                make_idesc(128, 128, 1, 0, 1), k > 0);
      } else if (mode == 3) {  // A MN-major: A stored [K=128 rows][M=128 cols]
        umma_ss(tD, make_smem_desc(a_addr + k * 2048, 16384, 1024),
                make_smem_desc(b_addr + k * 2048, 16384, 1024), make_idesc(128, 128, 1, 1, 1), k > 0);
      }
    }
    umma_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  {
    const int row = warp * 32 + lane;
    for (int c = 0; c < 4; ++c) {
      uint32_t r[32];
      tmem_ld32(tD + ((uint32_t)(warp * 32) << 16) + c * 32, r);
      tmem_ld_wait();
      for (int e = 0; e < 32; ++e) d_out[row * 128 + c * 32 + e] = __uint_as_float(r[e]);
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tb);
}

// mode 4: each thread fills one row of a [128][32] fp32 SW128 tile with (row*32+col), reduce-add twice
__global__ void __launch_bounds__(128, 1)
reduce_probe_kernel(const __grid_constant__ CUtensorMap tmD) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int row = threadIdx.x;
  for (int half = 0; half < 4; ++half)  // 4 slabs of 32 fp32 columns = D 128
    for (int ch = 0; ch < 8; ++ch) {
      float4 v;
      const int col = half * 32 + ch * 4;
      v.x = row * 128 + col; v.y = v.x + 1; v.z = v.x + 2; v.w = v.x + 3;
      *(float4 *)(smem + half * 16384 + row * 128 + ((ch ^ (row & 7)) * 16)) = v;
    }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int rep = 0; rep < 2; ++rep)
      for (int half = 0; half < 4; ++half) tma_reduce_add_4d(&tmD, smem + half * 16384, half * 32, 0, 0, 0);
    tma_store_commit();
    tma_store_wait<0>();
  }
}


// modes 5/6: no-swizzle K-major operands.  element (r, k) of a [128 x 16] tile lives at
//   (r / 8) * 256 + (k / 8) * 128 + (r % 8) * 16 + (k % 8) * 2
__device__ __forceinline__ uint64_t make_desc_noswz(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;  // layout type 0 = no swizzle
}
__global__ void __launch_bounds__(128, 1)
noswz_probe_kernel(const __nv_bfloat16 *a, const __nv_bfloat16 *b, float *d_out, int swapped) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bar_mma = (uint64_t *)(smem + 8192);
  uint32_t *tptr = (uint32_t *)(bar_mma + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, r = threadIdx.x;
  if (threadIdx.x == 0) { mbar_init(bar_mma, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc<128>(tptr);
  for (int kc = 0; kc < 2; ++kc) {
    *(uint4 *)(smem + (r / 8) * 256 + kc * 128 + (r % 8) * 16) = *(const uint4 *)(a + r * 16 + kc * 8);
    *(uint4 *)(smem + 4096 + (r / 8) * 256 + kc * 128 + (r % 8) * 16) = *(const uint4 *)(b + r * 16 + kc * 8);
  }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = *tptr;
  if (threadIdx.x == 0) {
    const uint32_t lbo = swapped ? 256 : 128, sbo = swapped ? 128 : 256;
    umma_ss(tb, make_desc_noswz(smem_u32(smem), lbo, sbo), make_desc_noswz(smem_u32(smem + 4096), lbo, sbo),
            make_idesc(128, 128, 1, 0, 0), 0);
    umma_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  for (int c = 0; c < 4; ++c) {
    uint32_t v[32];
    tmem_ld32(tb + ((uint32_t)(warp * 32) << 16) + c * 32, v);
    tmem_ld_wait();
    for (int e = 0; e < 32; ++e) d_out[(warp * 32 + lane) * 128 + c * 32 + e] = __uint_as_float(v[e]);
  }
  tc_fence_before(); __syncthreads();
  if (warp == 1) tmem_dealloc<128>(tb);
}

int main(int argc, char **argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  const int N = 128;
  if (mode == 4) {
    float *d; CK(cudaMalloc(&d, N * N * 4)); CK(cudaMemset(d, 0, N * N * 4));
    CUtensorMap tm = make_map(d, N, N, true);
    CK(cudaFuncSetAttribute(reduce_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
    reduce_probe_kernel<<<1, 128, 70000>>>(tm);
    CK(cudaDeviceSynchronize());
    std::vector<float> h(N * N); CK(cudaMemcpy(h.data(), d, N * N * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int i = 0; i < N * N; ++i) if (h[i] != 2.f * i) { if (bad < 5) printf("  [%d] got %f want %f\n", i, h[i], 2.f * i); ++bad; }
    printf("probe mode 4: %s (%d mismatches)\n", bad ? "FAIL" : "PASS", bad);
    return bad ? 1 : 0;
  }
  if (mode == 5 || mode == 6) {
    std::vector<__nv_bfloat16> A(N * 16), B(N * 16);
    std::vector<float> Af(N * 16), Bf(N * 16);
    srand(9);
    for (int i = 0; i < N * 16; ++i) { Af[i] = (float)(rand() % 7 - 3); Bf[i] = (float)(rand() % 5 - 2); A[i] = __float2bfloat16(Af[i]); B[i] = __float2bfloat16(Bf[i]); }
    __nv_bfloat16 *dA, *dB; float *dD;
    CK(cudaMalloc(&dA, N * 16 * 2)); CK(cudaMalloc(&dB, N * 16 * 2)); CK(cudaMalloc(&dD, N * N * 4));
    CK(cudaMemcpy(dA, A.data(), N * 16 * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), N * 16 * 2, cudaMemcpyHostToDevice));
    noswz_probe_kernel<<<1, 128, 9216>>>(dA, dB, dD, mode == 6);
    CK(cudaDeviceSynchronize());
    std::vector<float> h(N * N); CK(cudaMemcpy(h.data(), dD, N * N * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int m = 0; m < N; ++m) for (int n = 0; n < N; ++n) {
      float sref = 0; for (int k = 0; k < 16; ++k) sref += Af[m * 16 + k] * Bf[n * 16 + k];
      if (h[m * N + n] != sref) { if (bad < 4) printf("  [%d,%d] got %f want %f\n", m, n, h[m * N + n], sref); ++bad; }
    }
    printf("probe mode %d (no-swizzle K-major, %s): %s (%d mismatches)\n", mode,
           mode == 5 ? "LBO=K-core stride 128, SBO=row-group stride 256" : "LBO=256, SBO=128", bad ? "FAIL" : "PASS", bad);
    return bad ? 1 : 0;
  }
  std::vector<__nv_bfloat16> A(N * N), B(N * N);
  std::vector<float> Af(N * N), Bf(N * N), ref(N * N, 0.f);
  srand(7);
  for (int i = 0; i < N * N; ++i) { Af[i] = (float)(rand() % 7 - 3); Bf[i] = (float)(rand() % 5 - 2); A[i] = __float2bfloat16(Af[i]); B[i] = __float2bfloat16(Bf[i]); }
  for (int m = 0; m < N; ++m) for (int n = 0; n < N; ++n) {
    float s = 0;
    for (int k = 0; k < N; ++k) {
      float a = mode == 3 ? Af[k * N + m] : Af[m * N + k];
      float b = mode == 0 ? Bf[n * N + k] : Bf[k * N + n];
      s += a * b;
    }
    ref[m * N + n] = s;
  }
  __nv_bfloat16 *dA, *dB; float *dD;
  CK(cudaMalloc(&dA, N * N * 2)); CK(cudaMalloc(&dB, N * N * 2)); CK(cudaMalloc(&dD, N * N * 4));
  CK(cudaMemcpy(dA, A.data(), N * N * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), N * N * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0xff, N * N * 4));
  CUtensorMap tA = make_map(dA, N, N, false), tB = make_map(dB, N, N, false);
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
  probe_kernel<<<1, 128, 70000>>>(tA, tB, dA, dD, mode);
  CK(cudaDeviceSynchronize());
  std::vector<float> h(N * N); CK(cudaMemcpy(h.data(), dD, N * N * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int i = 0; i < N * N; ++i) if (h[i] != ref[i]) { if (bad < 6) printf("  [%d,%d] got %f want %f\n", i / N, i % N, h[i], ref[i]); ++bad; }
  printf("probe mode %d: %s (%d mismatches)\n", mode, bad ? "FAIL" : "PASS", bad);
  return bad ? 1 : 0;
}
