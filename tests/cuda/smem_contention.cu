// smem_contention.cu -- what slows a tcgen05.mma when other agents use the SM's shared memory?
// One CTA per SM (all SMs at once).  Warp 0 issues a long run of 128x128x16 bf16 MMAs (SS: both
// operands from shared memory, or TS: A from TMEM) and reports clk per MMA; bystander warps 1-4
// generate ONE kind of shared-memory traffic at a rate set by `gap` (clk between operations):
//   smem_contention <kind> <gap> [ts] [grid] [region KB per SM] [mma on]
//   kind 0 none | 1 STS.128 | 2 LDS.128 (conflict-free) | 3 LDS.128 (all lanes one address: broadcast)
//        4 bulk store smem->global (TMA reads smem)   | 5 bulk reduce-add.f32 smem->global
//        6 bulk load global->smem (TMA writes smem)   | 7 tcgen05.ld 32x32b.x32 (TMEM reads)
//        8 warps 1-2 bulk reduce-add, warps 3-4 bulk load (do loads queue behind stores in the TMA unit?)
//        9 warps 1-2 idle,            warps 3-4 bulk load (the reference for 8)
// Global targets are 64 KB per SM (L2 resident).  Behind DESIGN.md section 4.6 (round 2).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "../../flash_attention_ipu_b200/csrc/sm100_ptx.cuh"

using namespace fa;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ void bulk_store(void *g, const void *s, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_reduce_f32(void *g, const void *s, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}

__device__ __forceinline__ uint32_t lds128(const void *s) {
  uint32_t a, b, c, d;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(smem_u32(s)) : "memory");
  return a ^ b ^ c ^ d;
}

// smem: A 32 KB | B 32 KB | traffic area 64 KB (4 warps x 16 KB) | barriers
template <bool kTs>
__global__ void __launch_bounds__(192, 1)
contention_kernel(int n_mma, int kind, int gap, float *gbuf, long long *clk_out, long long *bytes_out, int region_kb, int mma_on) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bar = (uint64_t *)(smem + 131072);
  uint64_t *lbar = bar + 1;  // [4] bulk-load barriers
  uint32_t *tptr = (uint32_t *)(bar + 8);
  volatile int *stop = (volatile int *)(tptr + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&lbar[i], 1);
    fence_barrier_init();
    *stop = 0;
  }
  for (int i = threadIdx.x; i < 131072 / 16; i += blockDim.x) ((uint4 *)smem)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0) tmem_alloc<512>(tptr);
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = *tptr;
  float *g = gbuf + (size_t)blockIdx.x * region_kb * 256;  // region_kb per SM
  const int n_chunks = region_kb / 16;  // 4 KB chunks per warp
  if (warp == 0) {
    if (threadIdx.x == 0) {
      const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
      const uint32_t idesc = make_idesc(128, 128, 1, 0, 0);
      const long long t0 = clock64();
      if (!mma_on) { while (clock64() - t0 < 64ll * n_mma) { } }
      for (int i = 0; i < (mma_on ? n_mma : 0); ++i) {
        const uint32_t koff = ((i >> 2) & 1) * 16384 + (i & 3) * 32;
        const uint64_t da = make_smem_desc(a_addr + koff, 16, 1024), db = make_smem_desc(b_addr + koff, 16, 1024);
        if (kTs) umma_ts(tb, tb + 128 + (i & 7) * 8, db, idesc, 1); else umma_ss(tb, da, db, idesc, 1);
      }
      umma_commit(bar);
      mbar_wait(bar, 0);
      clk_out[blockIdx.x] = clock64() - t0;
      *stop = 1;
    }
  } else if (warp <= 4 && kind > 0) {
    const int mix = kind;
    if (kind == 8) kind = warp <= 2 ? 5 : 6;
    if (kind == 9) kind = warp <= 2 ? 0 : 6;
    uint8_t *area = smem + 65536 + (warp - 1) * 16384;
    const long long t0 = clock64();
    long long ops = 0, next = t0;
    uint32_t acc = 0;
    uint32_t phase = 0;
    while (!*stop) {
      while (clock64() < next) { }
      next += gap;
      uint8_t *slab = area + (ops & 3) * 4096;
      if (kind == 1) {
        uint4 v = make_uint4(acc, 1, 2, 3);
        #pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4 *>(slab + lane * 128 + ((c ^ (lane & 7)) * 16)) = v;
      } else if (kind == 2) {
        #pragma unroll
        for (int c = 0; c < 8; ++c) acc += lds128(slab + lane * 128 + ((c ^ (lane & 7)) * 16));
      } else if (kind == 3) {
        #pragma unroll
        for (int c = 0; c < 8; ++c) acc += lds128(slab + c * 16);
      } else if (kind == 4 || kind == 5) {
        if (lane == 0) {
          tma_store_wait_read<2>();
          if (kind == 4) bulk_store(g + ((warp - 1) * n_chunks + (ops % n_chunks)) * 1024, slab, 4096);
          else bulk_reduce_f32(g + ((warp - 1) * n_chunks + (ops % n_chunks)) * 1024, slab, 4096);
          tma_store_commit();
        }
        __syncwarp();
      } else if (kind == 6) {
        if (lane == 0) {
          mbar_arrive_expect_tx(&lbar[warp - 1], 4096);
          bulk_load_1d(slab, g + ((warp - 1) * n_chunks + (ops % n_chunks)) * 1024, 4096, &lbar[warp - 1]);
        }
        mbar_wait(&lbar[warp - 1], phase);
        phase ^= 1;
      } else if (kind == 7) {
        uint32_t r[32];
        tmem_ld32(tb + 256 + (static_cast<uint32_t>((warp & 3) * 32) << 16), r);
        tmem_ld_wait();
        acc += r[0] + r[31];
      }
      ++ops;
      if (kind == 0) break;
    }
    if (lane == 0 && (kind == 4 || kind == 5)) tma_store_wait<0>();
    if (lane == 0 && (mix < 8 || kind == 6)) {
      const long long per_op = (kind == 1 || kind == 2) ? 4096 : kind == 3 ? 4096 : kind == 7 ? 4096 : 4096;
      atomicAdd((unsigned long long *)&bytes_out[blockIdx.x], (unsigned long long)(ops * per_op * 1000 / (clock64() - t0 + 1)));
    }
    if (acc == 0x12345678u) clk_out[0] = acc;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

int main(int argc, char **argv) {
  const int kind = argc > 1 ? atoi(argv[1]) : 0;
  const int gap = argc > 2 ? atoi(argv[2]) : 256;
  const bool ts = argc > 3 && atoi(argv[3]) != 0;
  const int grid_arg = argc > 4 ? atoi(argv[4]) : 0;
  const int region_kb = argc > 5 ? atoi(argv[5]) : 64;
  const int mma_on = argc > 6 ? atoi(argv[6]) : 1;
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  if (grid_arg > 0) sms = grid_arg;
  const int n_mma = 16384;
  long long *clk, *byt; float *gbuf;
  CK(cudaMalloc(&clk, sms * 8)); CK(cudaMalloc(&byt, sms * 8)); CK(cudaMalloc(&gbuf, (size_t)sms * region_kb * 1024));
  CK(cudaMemset(gbuf, 0, (size_t)sms * region_kb * 1024));
  auto kern = ts ? contention_kernel<true> : contention_kernel<false>;
  const int smem = 131072 + 256;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaMemset(clk, 0, sms * 8)); CK(cudaMemset(byt, 0, sms * 8));
    kern<<<sms, 192, smem>>>(n_mma, kind, gap, gbuf, clk, byt, region_kb, mma_on);
    CK(cudaDeviceSynchronize());
  }
  std::vector<long long> hc(sms), hb(sms);
  CK(cudaMemcpy(hc.data(), clk, sms * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hb.data(), byt, sms * 8, cudaMemcpyDeviceToHost));
  double sum = 0, bsum = 0;
  for (int i = 0; i < sms; ++i) { sum += hc[i]; bsum += hb[i] / 1000.0; }
  static const char *names[] = {"none", "STS.128", "LDS.128", "LDS.128 broadcast", "bulk store", "bulk reduce f32", "bulk load", "tcgen05.ld", "reduce + load: load rate", "load only (2 warps)"};
  printf("contention %s kind %d (%-17s) gap %5d grid %3d region %5d KB: %7.2f clk/MMA; bystander traffic %6.1f B/clk/SM\n",
         mma_on ? (ts ? "TS" : "SS") : "--", kind, names[kind], gap, sms, region_kb, sum / sms / n_mma, bsum / sms);
  return 0;
}
