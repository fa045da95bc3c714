// exp_rate.cu -- how fast can ONE warp per SM sub-partition turn 64 scores into 64 exponentials
// (the softmax inner loop of the forward kernel: FFMA2 argument, exp2, FADD2 row sum, bf16 pack),
// as a function of how many pairs go to the FMA-pipe polynomial instead of MUFU and in what order?
// Usage: exp_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "../../flash_attention_ipu_b200/csrc/sm100_ptx.cuh"
using namespace fa;

// POLY(pair) decides which of the 32 pairs use the polynomial
template <int kMode>
__device__ __forceinline__ bool use_poly(int pair) {
  switch (kMode) {
    case 0: return false;                 // all MUFU
    case 1: return pair % 4 == 3;         // 25 %, every 4th
    case 2: return pair % 3 == 2;         // 33 %
    case 3: return pair % 2 == 1;         // 50 %, alternating
    case 4: return pair >= 16;            // 50 %, blocked
    case 5: return pair % 8 >= 5;         // 37.5 %
    case 6: return true;                  // all polynomial
    default: return pair % 8 == 7;        // 12.5 %
  }
}

template <int kMode>
__global__ void __launch_bounds__(128, 1) exp_kernel(const float *in, float *out, long long *clk, int reps) {
  float s[64];
  #pragma unroll
  for (int i = 0; i < 64; ++i) s[i] = in[threadIdx.x * 64 + i];
  const uint64_t sc2 = f2_pack(0.125f, 0.125f), nm2 = f2_pack(-1.f, -1.f);
  uint64_t sum2[4] = {0, 0, 0, 0};
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    #pragma unroll
    for (int e = 0; e < 64; e += 2) {
      float x0, x1, p0, p1;
      f2_unpack(f2_fma(f2_pack(s[e], s[e + 1]), sc2, nm2), x0, x1);
      if (use_poly<kMode>(e >> 1)) poly_exp2_pair(x0, x1, p0, p1);
      else { p0 = fast_exp2(x0); p1 = fast_exp2(x1); }
      sum2[(e >> 1) & 3] = f2_add(sum2[(e >> 1) & 3], f2_pack(p0, p1));
      acc ^= pack2<true>(p0, p1);
      s[e] = p0 - 0.5f; s[e + 1] = p1 - 0.5f;  // feed back so that nothing is hoisted
    }
  }
  const long long t1 = clock64();
  float a, b;
  f2_unpack(f2_add(f2_add(sum2[0], sum2[1]), f2_add(sum2[2], sum2[3])), a, b);
  out[blockIdx.x * 128 + threadIdx.x] = a + b + __uint_as_float(acc & 0x3f800000u);
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
  float *in, *out; long long *clk, h[148];
  cudaMalloc(&in, 128 * 64 * 4); cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&clk, sizeof(h));
  cudaMemset(in, 0, 128 * 64 * 4);
  const char *names[] = {"all MUFU", "25% poly (every 4th pair)", "33% poly", "50% poly alternating", "50% poly blocked",
                         "37.5% poly", "all polynomial", "12.5% poly"};
  void (*k[8])(const float *, float *, long long *, int) = {exp_kernel<0>, exp_kernel<1>, exp_kernel<2>, exp_kernel<3>,
                                                            exp_kernel<4>, exp_kernel<5>, exp_kernel<6>, exp_kernel<7>};
  const int reps = 256;
  for (int m = 0; m < 8; ++m) {
    k[m]<<<148, 128>>>(in, out, clk, reps);
    cudaDeviceSynchronize();
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double s = 0; for (long long v : h) s += v;
    printf("mode %d  %-28s %.0f clk per 64 exponentials per warp (1 warp / sub-partition)\n", m, names[m], s / 148 / reps);
  }
  return 0;
}
