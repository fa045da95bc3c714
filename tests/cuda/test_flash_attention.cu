// GPU tests of libflash_attention.so through its C ABI -- the B200 counterpart of the
// reference's Catch2 binary (tests/cpp/test_flash_attention.cpp:243-313): same case matrix and
// the same max-abs thresholds, plus bf16.  Expected values come from the plain-C oracle
// (oracle/vanilla_attention.c, a restatement of cpp/vanilla_attention.cpp) evaluated on the
// same (dtype-rounded) inputs.  Usage: build/tests [substring-filter] ; exit code = #failures.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "fa_b200.h"

extern "C" int oracle_vanilla_attention_fwd(const float *, float *, float *, int, int, int, int);
extern "C" int oracle_vanilla_attention_bwd(const float *, const float *, float *, int, int, int, int);

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(99); } } while (0)

struct Rng {  // fixed-seed N(0,1); poprand's stream (seed {40,90}, test_flash_attention.cpp:67-74)
  uint64_t s;  // is not reproducible off-IPU and the tests depend only on the distribution
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 40090) {}
  double uni() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return ((s >> 11) + 0.5) / 9007199254740992.0; }
  float normal() { return (float)(std::sqrt(-2.0 * std::log(uni())) * std::cos(6.283185307179586 * uni())); }
};

static float round_to(float x, int dtype) {
  if (dtype == FA_B200_BF16) return __bfloat162float(__float2bfloat16(x));
  if (dtype == FA_B200_F16) return __half2float(__float2half(x));
  return x;
}
static void to_device(const std::vector<float> &h, int dtype, void *d) {
  size_t n = h.size();
  if (dtype == FA_B200_F32) { CK(cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice)); return; }
  std::vector<uint16_t> t(n);
  for (size_t i = 0; i < n; ++i) {
    if (dtype == FA_B200_BF16) { __nv_bfloat16 b = __float2bfloat16(h[i]); memcpy(&t[i], &b, 2); }
    else { __half b = __float2half(h[i]); memcpy(&t[i], &b, 2); }
  }
  CK(cudaMemcpy(d, t.data(), n * 2, cudaMemcpyHostToDevice));
}
static std::vector<float> from_device(const void *d, size_t n, int dtype) {
  std::vector<float> h(n);
  if (dtype == FA_B200_F32) { CK(cudaMemcpy(h.data(), d, n * 4, cudaMemcpyDeviceToHost)); return h; }
  std::vector<uint16_t> t(n);
  CK(cudaMemcpy(t.data(), d, n * 2, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < n; ++i) {
    if (dtype == FA_B200_BF16) { __nv_bfloat16 b; memcpy(&b, &t[i], 2); h[i] = __bfloat162float(b); }
    else { __half b; memcpy(&b, &t[i], 2); h[i] = __half2float(b); }
  }
  return h;
}
static double max_abs_err(const std::vector<float> &a, const float *b, size_t n) {
  double m = 0;
  for (size_t i = 0; i < n; ++i) {
    double d = std::fabs((double)a[i] - (double)b[i]);
    if (!(d <= m)) m = d;  // NaN propagates to the result
  }
  return m;
}

static const char *dname(int dt) { return dt == FA_B200_F32 ? "fp32" : dt == FA_B200_F16 ? "fp16" : "bf16"; }
static std::string g_filter;
static int g_failed = 0, g_run = 0;

// One case: forward (and optionally backward) vs the oracle.  num_chunks_* only feed the hints.
static void run_case(const char *what, int G, int L, int D, int ncq, int nckv, int dtype, int causal,
                     bool backward, double tol_fwd, double tol_bwd) {
  char name[160];
  snprintf(name, sizeof name, "%s G=%d L=%d D=%d chunks=%dx%d %s %s", what, G, L, D, ncq, nckv,
           dname(dtype), causal ? "causal" : "full");
  if (!g_filter.empty() && std::string(name).find(g_filter) == std::string::npos) return;
  ++g_run;
  const size_t GLD = (size_t)G * L * D, es = dtype == FA_B200_F32 ? 4 : 2;
  Rng rng(1234 + L * 31 + D);
  std::vector<float> qkv(3 * GLD), grad(GLD);
  const float qs = 1.0f / std::sqrt((float)D);  // the only in-repo caller pre-scales q
  for (size_t i = 0; i < 3 * GLD; ++i) qkv[i] = round_to(rng.normal() * (i < GLD ? qs : 1.f), dtype);
  for (size_t i = 0; i < GLD; ++i) grad[i] = round_to(rng.normal(), dtype);
  std::vector<float> o_ref(GLD), lse_ref((size_t)G * L), dqkv_ref(3 * GLD);
  oracle_vanilla_attention_fwd(qkv.data(), o_ref.data(), lse_ref.data(), G, L, D, causal);
  if (backward) oracle_vanilla_attention_bwd(grad.data(), qkv.data(), dqkv_ref.data(), G, L, D, causal);

  void *d_qkv, *d_o, *d_do, *d_dqkv, *d_ws; float *d_lse;
  CK(cudaMalloc(&d_qkv, 3 * GLD * es)); CK(cudaMalloc(&d_o, GLD * es)); CK(cudaMalloc(&d_do, GLD * es));
  CK(cudaMalloc(&d_dqkv, 3 * GLD * es)); CK(cudaMalloc(&d_lse, (size_t)G * L * 4));
  CK(cudaMemset(d_o, 0xff, GLD * es)); CK(cudaMemset(d_dqkv, 0xff, 3 * GLD * es));
  to_device(qkv, dtype, d_qkv); to_device(grad, dtype, d_do);
  fa_b200_problem p{};
  p.B = 1; p.H = G; p.Lq = L; p.Lk = L; p.D = D; p.dtype = dtype; p.causal = causal;
  p.softmax_scale = 1.f; p.block_q_hint = L / ncq; p.block_kv_hint = L / nckv;
  auto T = [&](void *base, int which) {
    return fa_b200_tensor{(char *)base + which * GLD * es, (int64_t)GLD, (int64_t)L * D, (int64_t)D};
  };
  fa_b200_tensor q = T(d_qkv, 0), k = T(d_qkv, 1), v = T(d_qkv, 2), o = T(d_o, 0), dout = T(d_do, 0);
  fa_b200_tensor dq = T(d_dqkv, 0), dk = T(d_dqkv, 1), dv = T(d_dqkv, 2);
  bool ok = true;
  int rc = fa_b200_fwd(&p, &q, &k, &v, &o, d_lse, nullptr);
  if (rc) { printf("  fa_b200_fwd rc=%d: %s\n", rc, fa_b200_last_error()); ok = false; }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("FAILED %s : CUDA error after fwd: %s\n", name, cudaGetErrorString(e)); exit(98); }
  double ef = max_abs_err(from_device(d_o, GLD, dtype), o_ref.data(), GLD);
  double el = max_abs_err(from_device(d_lse, (size_t)G * L, FA_B200_F32), lse_ref.data(), (size_t)G * L);
  if (!(ef <= tol_fwd) || !(el <= std::max(tol_fwd, 1e-3))) ok = false;
  double eb[3] = {0, 0, 0};
  if (backward && ok) {
    size_t wsb = fa_b200_bwd_workspace_bytes(&p);
    CK(cudaMalloc(&d_ws, wsb));
    rc = fa_b200_bwd(&p, &q, &k, &v, &o, &dout, d_lse, &dq, &dk, &dv, d_ws, wsb, nullptr);
    if (rc) { printf("  fa_b200_bwd rc=%d: %s\n", rc, fa_b200_last_error()); ok = false; }
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("FAILED %s : CUDA error after bwd: %s\n", name, cudaGetErrorString(e)); exit(98); }
    std::vector<float> got = from_device(d_dqkv, 3 * GLD, dtype);
    for (int t = 0; t < 3; ++t) {
      std::vector<float> part(got.begin() + t * GLD, got.begin() + (t + 1) * GLD);
      eb[t] = max_abs_err(part, dqkv_ref.data() + t * GLD, GLD);
      if (!(eb[t] <= tol_bwd)) ok = false;
    }
    CK(cudaFree(d_ws));
  }
  printf("%s %s : family=%d  max|dO|=%.3g (tol %.0e)  max|dLSE|=%.3g", ok ? "ok    " : "FAILED", name,
         fa_b200_kernel_family(&p), ef, tol_fwd, el);
  if (backward) printf("  dQ=%.3g dK=%.3g dV=%.3g (tol %.0e)", eb[0], eb[1], eb[2], tol_bwd);
  printf("\n");
  fflush(stdout);
  if (!ok) ++g_failed;
  CK(cudaFree(d_qkv)); CK(cudaFree(d_o)); CK(cudaFree(d_do)); CK(cudaFree(d_dqkv)); CK(cudaFree(d_lse));
}

int main(int argc, char **argv) {
  if (argc > 1) g_filter = argv[1];
  // --- reference matrix, fp32 (test_flash_attention.cpp:245-259 fwd 1e-5; :273-287 bwd 1e-4)
  const int small[][2] = {{1, 1}, {2, 1}, {1, 2}, {2, 2}, {3, 2}, {2, 3}, {3, 3}};
  for (auto &c : small) run_case("ref-matrix", 4, 6, 2, c[0], c[1], FA_B200_F32, 1, true, 1e-5, 1e-4);
  const int mid[][2] = {{2, 2}, {4, 2}, {2, 4}, {4, 4}};
  for (auto &c : mid) run_case("ref-matrix", 8, 256, 128, c[0], c[1], FA_B200_F32, 1, true, 1e-5, 1e-4);
  // --- reference matrix, half precision (:261-267 fwd 1e-2; :289-295 bwd 1.0); fp16 and bf16.
  // G reduced 8 -> 2 to keep the scalar CPU oracle in seconds.
  for (int dt : {FA_B200_F16, FA_B200_BF16})
    for (auto &c : mid) run_case("ref-matrix", 2, 2048, 128, c[0], c[1], dt, 1, true, 1e-2, 1.0);
  // --- tensor-path shapes: single tile, tile pair, ragged tails, D=64, non-causal extension
  for (int dt : {FA_B200_BF16, FA_B200_F16}) {
    run_case("tile", 1, 128, 128, 1, 1, dt, 0, true, 1e-2, 1e-1);
    run_case("tile", 1, 128, 128, 1, 1, dt, 1, true, 1e-2, 1e-1);
    run_case("tile", 2, 256, 128, 1, 1, dt, 1, true, 1e-2, 1e-1);
    run_case("tile", 2, 512, 64, 1, 1, dt, 1, true, 1e-2, 1e-1);
    run_case("tile", 3, 1024, 64, 4, 4, dt, 1, true, 1e-2, 1e-1);
    run_case("tile", 2, 512, 64, 4, 2, dt, 0, true, 1e-2, 1e-1);
    run_case("ragged", 2, 200, 128, 1, 1, dt, 1, true, 1e-2, 1e-1);
    run_case("ragged", 2, 331, 64, 1, 1, dt, 0, true, 1e-2, 1e-1);
    run_case("ragged", 1, 8, 128, 1, 1, dt, 1, true, 1e-2, 1e-1);
  }
  printf("%d cases run, %d failed\n", g_run, g_failed);
  return g_failed;
}
