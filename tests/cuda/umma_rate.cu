// umma_rate.cu -- how fast does ONE CTA per SM execute tcgen05.mma, by operand source and layout?
// Issues long runs of back-to-back MMAs on zero-filled shared memory (no TMA, no epilogue) on all
// 148 SMs at once and reports SM clocks per instruction.  The nominal cost of a 128xNx16 bf16 MMA is
// N/2 clocks; anything above it is operand delivery (shared-memory bandwidth), which is what bounds
// the attention kernels.   Usage: umma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "../../flash_attention_ipu_b200/csrc/sm100_ptx.cuh"
using namespace fa;

// spin on try_wait with nothing else in the loop (no spin counter, no trap path)
__device__ __forceinline__ void mbar_wait_lean(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "LAB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra LAB_WAIT;\n\t"
      "DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

template <int mode>
__global__ void __launch_bounds__(384, 1) rate_kernel(int iters, long long *out, int random_data) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, bar2[4];
  __shared__ uint32_t tmem_ptr;
  for (int i = threadIdx.x; i < 128 * 1024 / 4; i += blockDim.x) {
    // zeros, or pseudo-random bf16 pairs in (-2, 2): real operand bit patterns toggle the datapath
    uint32_t x = (i * 2654435761u) ^ (i >> 7) * 40503u, y = x * 2246822519u + 374761393u;
    const uint32_t v = ((x & 0x807f807fu) | 0x3f003f00u) ^ (y & 0x00800080u);
    reinterpret_cast<uint32_t *>(smem)[i] = random_data ? v : 0u;
  }
  if (threadIdx.x == 0) { mbar_init(&bar, 1); for (int i = 0; i < 4; ++i) mbar_init(&bar2[i], 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(&tmem_ptr);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (mode >= 19 && threadIdx.x >= 32) {  // bystander warps polling an mbarrier, like idle pipeline roles
    mbar_wait_lean(&bar2[2], 0);
  }
  const bool warp_mode = mode >= 11;  // smem: A at 0, B ring of 3 x 32 KB at 32 KB
  const bool leader = warp_mode ? (threadIdx.x < 32 && elect_one()) : threadIdx.x == 0;
  if (warp_mode ? threadIdx.x < 32 : threadIdx.x == 0) {
    constexpr uint32_t hi = smem_desc_hi(1024);
    const uint32_t a_k = smem_desc_lo(smem_u32(smem), 16), b_k = smem_desc_lo(smem_u32(smem + 32768), 16);
    const uint32_t a_mn = smem_desc_lo(smem_u32(smem), 16384), b_mn = smem_desc_lo(smem_u32(smem + 32768), 16384);
    const uint32_t i_kk = make_idesc(128, 128, 1, 0, 0), i_kmn = make_idesc(128, 128, 1, 0, 1),
                   i_mnmn = make_idesc(128, 128, 1, 1, 1), i_kk256 = make_idesc(128, 256, 1, 0, 0),
                   i_kk64 = make_idesc(128, 64, 1, 0, 0);
    const long long t0 = clock64();
    if (mode >= 23) {
      // straight-line rotation (no per-MMA branching): per iteration S0, PV1, S1, PV0, 8 MMAs each.
      // 23: as the forward kernel.  24: PV reads A from the other tile's columns (no WAR on the
      // S accumulator that the next GEMM overwrites).  25: all four GEMMs SS.  26: all four TS.
      for (int it = 0; it < iters / 4; ++it) {
        mbar_wait_lean(&bar2[3], 1); tc_fence_after();
        if (leader) {
          const uint32_t a0 = (mode == 24) ? tm + 128 : tm, a1 = (mode == 24) ? tm : tm + 128;
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t offk = ((k / 4) * 16384 + (k % 4) * 32) >> 4;
            if (mode == 26) umma_ts2(tm, tm + 256 + k * 8, b_k + offk, hi, i_kk, k > 0);
            else umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk, k > 0);
          }
          umma_commit(&bar2[0]);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t offk = ((k / 4) * 16384 + (k % 4) * 32) >> 4, offmn = k * (2048 >> 4);
            if (mode == 25) umma_ss2(tm + 384, a_k + offk, hi, b_mn + offmn, hi, i_kmn, 1);
            else umma_ts2(tm + 384, a1 + k * 8, b_mn + offmn, hi, i_kmn, 1);
          }
          umma_commit(&bar2[0]);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t offk = ((k / 4) * 16384 + (k % 4) * 32) >> 4;
            if (mode == 26) umma_ts2(tm + 128, tm + 320 + k * 8, b_k + offk, hi, i_kk, k > 0);
            else umma_ss2(tm + 128, a_k + offk, hi, b_k + offk, hi, i_kk, k > 0);
          }
          umma_commit(&bar2[0]);
          #pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t offk = ((k / 4) * 16384 + (k % 4) * 32) >> 4, offmn = k * (2048 >> 4);
            if (mode == 25) umma_ss2(tm + 256, a_k + offk, hi, b_mn + offmn, hi, i_kmn, 1);
            else umma_ts2(tm + 256, a0 + k * 8, b_mn + offmn, hi, i_kmn, 1);
          }
          umma_commit(&bar2[0]);
        }
        __syncwarp();
      }
    } else
    for (int it = 0; it < iters; ++it) {
      if (mode == 11 || mode == 12) tc_fence_after();
      if (mode == 12) { mbar_wait(&bar2[3], 1); }       // a wait that is already satisfied (bounded-spin form)
      if (mode == 13 || mode == 14) { mbar_wait_lean(&bar2[3], 1); }  // the same wait, lean form
      if (mode == 15) { mbar_wait_lean(&bar2[3], 1); tc_fence_after(); }
      if (mode == 16) { mbar_wait_lean(&bar2[3], 1); mbar_wait_lean(&bar2[1], 1); mbar_wait_lean(&bar2[3], 1); mbar_wait_lean(&bar2[1], 1); }
      const uint32_t stage = (mode >= 17) ? static_cast<uint32_t>((it >> 1) % 3) * (32768 >> 4) : 0u;  // ring stage
      if (mode >= 18) { mbar_wait_lean(&bar2[3], 1); tc_fence_after(); }
      if (leader && (mode < 14 || mode >= 17)) {  // mode 19: as 18 with 11 bystander warps
      #pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t offk = (((k / 4) * 16384 + (k % 4) * 32) >> 4) + stage, offmn = k * (2048 >> 4) + stage;
        switch (mode) {
          case 0: umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk, 1); break;             // S = Q K^T
          case 1: umma_ts2(tm + 256, tm + k * 8, b_mn + offmn, hi, i_kmn, 1); break;         // O += P V
          case 2: umma_ss2(tm, a_mn + offmn, hi, b_mn + offmn, hi, i_mnmn, 1); break;        // dQ = dS K
          case 3: umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk256, 1); break;           // N = 256
          case 4: umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk64, 1); break;            // N = 64
          case 5: umma_ss2(tm, a_k + offk, hi, b_mn + offmn, hi, i_kmn, 1); break;           // dK = dS^T Q
          case 6: umma_ts2(tm + 256, tm + k * 8, b_k + offk, hi, i_kk, 1); break;            // TS, B K-major
          case 8: case 10: umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk, 1); break;     // + commits
          case 9: case 7: case 11: case 12: case 13: case 20: case 21: case 22: {  // the forward's exact accumulator rotation: S0 PV1 S1 PV0
            const int ph = it & 3;
            const uint32_t acc = (mode == 21) ? 1u : (k > 0);   // 21: never start a fresh accumulator
            if (mode == 22 && (ph & 1) == 0) { umma_ss2(tm + 64 * ph, a_k + offk, hi, b_k + offk, hi, i_kk, 1); break; }  // 22: S only, acc=1
            if (ph == 0) umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk, acc);
            else if (ph == 1) umma_ts2(tm + 384, tm + 128 + k * 8, b_mn + offmn, hi, i_kmn, 1);
            else if (ph == 2) umma_ss2(tm + 128, a_k + offk + (32768 >> 4), hi, b_k + offk, hi, i_kk, acc);
            else umma_ts2(tm + 256, tm + k * 8, b_mn + offmn, hi, i_kmn, 1);
            break; }
          case 17: case 18: case 19: if (it & 1) umma_ts2(tm + 256, tm + k * 8, b_mn + offmn, hi, i_kmn, 1);    // fwd mix
                  else umma_ss2(tm, a_k + offk, hi, b_k + offk, hi, i_kk, 1); break;
        }
        if (mode == 10 && k == 3) umma_commit(&bar2[1]);
      }
      if (mode >= 8) umma_commit(&bar2[mode >= 19 ? 0 : (it & 1 ? 2 : 0)]);   // one commit per 8 MMAs, nobody waits on it
      }
      if (warp_mode) __syncwarp();
    }
    if (leader) umma_commit(&bar);
    mbar_wait(&bar, 0);
    if (leader) out[blockIdx.x] = clock64() - t0;
    if (mode >= 19 && leader) mbar_arrive(&bar2[2]);  // release the bystanders
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

int main() {
  const char *names[] = {"SS  A K-major  B K-major  N=128 (S=QK^T, S^T, dP^T)", "TS  A TMEM     B MN-major N=128 (O+=PV, dV)",
                         "SS  A MN-major B MN-major N=128 (dQ=dS K)", "SS  A K-major  B K-major  N=256",
                         "SS  A K-major  B K-major  N=64", "SS  A K-major  B MN-major N=128 (dK=dS^T Q)",
                         "TS  A TMEM     B K-major  N=128", "alternating SS(K,K) / TS(MN) GEMMs (forward mix)",
                         "mode 0 + tcgen05.commit every 8 MMAs", "mode 7 + tcgen05.commit every 8 MMAs",
                         "mode 0 + tcgen05.commit every 4 MMAs",
                         "warp loop, mode 9 + tcgen05.fence::after_thread_sync per GEMM", "warp loop, mode 11 + satisfied mbarrier wait per GEMM",
                         "warp loop, mode 9 + satisfied mbarrier wait per GEMM",
                         "NO MMAs: one satisfied mbarrier wait per iteration (clk per 8 = per wait / 8)",
                         "NO MMAs: satisfied wait + tcgen05.fence::after per iteration", "NO MMAs: 4 satisfied waits per iteration",
                         "warp loop, forward mix, B descriptor changes every GEMM (ring stage) + commit", "mode 17 + satisfied wait + fence per GEMM",
                         "mode 18 + 11 bystander warps spinning on mbarrier.try_wait",
                         "forward accumulator rotation S0 PV1 S1 PV0 (4 TMEM accumulators) + wait/fence/commit",
                         "mode 20 with accumulate=1 on every MMA (no fresh accumulator)", "mode 21 (same, control)",
                         "straight-line S0 PV1 S1 PV0 exactly as the forward kernel", "as 23 but PV reads P from the OTHER tile's S columns",
                         "as 23 with all four GEMMs SS", "as 23 with all four GEMMs TS"};
  const int nominal[] = {64, 64, 64, 128, 32, 64, 64, 64, 64, 64, 64, 64, 64, 64, 0, 0, 0, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64};
  long long *d, h[148];
  cudaMalloc(&d, sizeof(h));
  const int iters = 512;
  void (*kerns[27])(int, long long *, int) = {rate_kernel<0>, rate_kernel<1>, rate_kernel<2>, rate_kernel<3>,
                                        rate_kernel<4>, rate_kernel<5>, rate_kernel<6>, rate_kernel<7>,
                                        rate_kernel<8>, rate_kernel<9>, rate_kernel<10>,
                                        rate_kernel<11>, rate_kernel<12>, rate_kernel<13>,
                                        rate_kernel<14>, rate_kernel<15>, rate_kernel<16>, rate_kernel<17>, rate_kernel<18>, rate_kernel<19>, rate_kernel<20>, rate_kernel<21>, rate_kernel<22>, rate_kernel<23>, rate_kernel<24>, rate_kernel<25>, rate_kernel<26>};
  for (int rnd = 0; rnd < 2; ++rnd)
  for (int mode = 0; mode < 27; ++mode) {
    if (rnd && mode != 0 && mode != 1 && mode != 7) continue;
    if (rnd) printf("random operand data: ");
    cudaFuncSetAttribute(kerns[mode], cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    kerns[mode]<<<148, mode >= 19 ? 384 : 128, 200 * 1024>>>(iters * (rnd ? 8 : 1), d, rnd);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("mode %d: CUDA error %s\n", mode, cudaGetErrorString(cudaGetLastError())); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double s = 0; long long mx = 0;
    for (long long v : h) { s += v; mx = v > mx ? v : mx; }
    const double n = iters * (rnd ? 8 : 1) * 8.0;
    printf("mode %d  %-52s  %.1f clk/MMA avg, %.1f max (nominal %d)\n", mode, names[mode], s / 148 / n, mx / n, nominal[mode]);
  }
  return 0;
}
