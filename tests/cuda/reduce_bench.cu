// reduce_bench.cu -- what is the chip-wide ceiling for accumulating fp32 dQ tiles into global
// memory?  One CTA per SM adds [128 x 128] fp32 tiles (64 KB) into a big tensor, by
//   0: TMA reduce-add (cp.reduce.async.bulk.tensor ... .add), 4 x 16 KB ops per tile
//   1: TMA plain store (upper bound of the store path)
//   2: red.global.add.v4.f32, one thread per row (row-strided, what a TMEM lane owns)
//   3: red.global.add.v4.f32, warp-coalesced (lane l -> 4 floats at column 4l)
// Usage: reduce_bench <mode> <tensor MB> [iters]
#include <cstdio>
#include <cstdlib>
#include "../../flash_attention_ipu_b200/csrc/sm100_ptx.cuh"
using namespace fa;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(2); } } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ void red_add_v4(float *p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(128, 1)
bench_kernel(const __grid_constant__ CUtensorMap tm, float *base, int num_tiles, int iters, int mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int row = threadIdx.x;
  for (int i = 0; i < 4 * 16384 / 4 / 128; ++i) reinterpret_cast<float *>(smem)[i * 128 + row] = 1.0f;
  fence_proxy_async_smem();
  __syncthreads();
  for (int it = 0; it < iters; ++it) {
    const int tile = (blockIdx.x + it * gridDim.x) % num_tiles;
    if (mode <= 1) {
      if (threadIdx.x == 0) {
        for (int s = 0; s < 4; ++s) {
          if (mode == 0) tma_reduce_add_4d(&tm, smem + s * 16384, s * 32, tile * 128, 0, 0);
          else tma_store_4d(&tm, smem + s * 16384, s * 32, tile * 128, 0, 0);
        }
        tma_store_commit();
        tma_store_wait_read<2>();
      }
    } else if (mode == 2) {
      float *p = base + (size_t)(tile * 128 + row) * 128;
      #pragma unroll
      for (int c = 0; c < 32; ++c) red_add_v4(p + c * 4, 1.f, 1.f, 1.f, 1.f);
    } else {
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      #pragma unroll 8
      for (int r = warp; r < 128; r += 4)
        red_add_v4(base + (size_t)(tile * 128 + r) * 128 + lane * 4, 1.f, 1.f, 1.f, 1.f);
    }
  }
  if (mode <= 1 && threadIdx.x == 0) tma_store_wait<0>();
}

int main(int argc, char **argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  const size_t mb = argc > 2 ? atoi(argv[2]) : 512;
  const int iters = argc > 3 ? atoi(argv[3]) : 400;
  const int num_tiles = (int)(mb * 1024 * 1024 / 65536);
  float *d; CK(cudaMalloc(&d, (size_t)num_tiles * 65536)); CK(cudaMemset(d, 0, (size_t)num_tiles * 65536));
  void *sym = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q));
  CUtensorMap tm;
  cuuint64_t dims[4] = {128, (cuuint64_t)num_tiles * 128, 1, 1};
  cuuint64_t strides[3] = {512, (cuuint64_t)num_tiles * 65536, (cuuint64_t)num_tiles * 65536};
  cuuint32_t box[4] = {32, 128, 1, 1}, estr[4] = {1, 1, 1, 1};
  CUresult r = ((EncodeTiledFn)sym)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, d, dims, strides, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed\n"); return 2; }
  CK(cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 1024));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaEventRecord(e0));
    bench_kernel<<<148, 128, 65536 + 1024>>>(tm, d, num_tiles, iters, mode);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double bytes = 148.0 * iters * 65536;
    printf("mode %d tensor %zu MB: %.3f ms, %.1f GB/s (%.1f B/clk/SM at 1.9 GHz)\n", mode, mb, ms, bytes / ms / 1e6,
           bytes / ms / 1e6 / 148 / 1.9);
  }
  return 0;
}
