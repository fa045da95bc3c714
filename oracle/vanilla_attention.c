/* CPU oracle, plain C.  TEST INFRASTRUCTURE ONLY -- never linked into the product
 * library (libflash_attention.so); used by tests/, smoke() and bench.py's CPU leg.
 *
 * Restates the reference's un-chunked attention, which is what the reference's own
 * C++ tests use as the expected value:
 *   forward : flash_attention_ipu/cpp/vanilla_attention.cpp:33-66
 *   backward: flash_attention_ipu/cpp/vanilla_attention.cpp:68-124
 *   mask    : additive -10000 above the diagonal (:52-54, triu :15-31)
 * Parity status: pinned (checked in tests/test_oracle.py against the tests/golden npz fixtures,
 * which are outputs of the reference's own Python code; see oracle/gen_golden.py).
 *
 * Layout: qkv [3][G][L][D] row-major fp32, out [G][L][D], lse [G][L], grad [G][L][D],
 * dqkv [3][G][L][D].  Accumulation in double.  causal==0 is the config-4 extension.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define MASK_VALUE (-10000.0)

/* one row of softmax(q_i . K^T + mask): writes p[0..L) and returns logsumexp */
static double softmax_row(const float *q_i, const float *K, int i, int L, int D,
                          int causal, double *p) {
    double m = -INFINITY;
    for (int j = 0; j < L; ++j) {
        double s = 0.0;
        const float *k_j = K + (size_t)j * D;
        for (int d = 0; d < D; ++d) s += (double)q_i[d] * (double)k_j[d];
        if (causal && j > i) s += MASK_VALUE;
        p[j] = s;
        if (s > m) m = s;
    }
    double l = 0.0;
    for (int j = 0; j < L; ++j) { p[j] = exp(p[j] - m); l += p[j]; }
    for (int j = 0; j < L; ++j) p[j] /= l;
    return m + log(l);
}

int oracle_vanilla_attention_fwd(const float *qkv, float *out, float *lse,
                                 int G, int L, int D, int causal) {
    size_t GLD = (size_t)G * L * D;
    const float *Q = qkv, *K = qkv + GLD, *V = qkv + 2 * GLD;
    double *p = (double *)malloc(sizeof(double) * L);
    double *acc = (double *)malloc(sizeof(double) * D);
    if (!p || !acc) return -1;
    for (int g = 0; g < G; ++g) {
        const float *Kg = K + (size_t)g * L * D, *Vg = V + (size_t)g * L * D;
        for (int i = 0; i < L; ++i) {
            const float *q_i = Q + ((size_t)g * L + i) * D;
            double r = softmax_row(q_i, Kg, i, L, D, causal, p);
            if (lse) lse[(size_t)g * L + i] = (float)r;
            memset(acc, 0, sizeof(double) * D);
            for (int j = 0; j < L; ++j) {
                const float *v_j = Vg + (size_t)j * D;
                for (int d = 0; d < D; ++d) acc[d] += p[j] * (double)v_j[d];
            }
            float *o_i = out + ((size_t)g * L + i) * D;
            for (int d = 0; d < D; ++d) o_i[d] = (float)acc[d];
        }
    }
    free(p); free(acc);
    return 0;
}

int oracle_vanilla_attention_bwd(const float *grad, const float *qkv, float *dqkv,
                                 int G, int L, int D, int causal) {
    size_t GLD = (size_t)G * L * D;
    const float *Q = qkv, *K = qkv + GLD, *V = qkv + 2 * GLD;
    double *p = (double *)malloc(sizeof(double) * L);
    double *ds = (double *)malloc(sizeof(double) * L);
    double *dq = (double *)malloc(sizeof(double) * D);
    double *dk = (double *)calloc((size_t)L * D, sizeof(double));
    double *dv = (double *)calloc((size_t)L * D, sizeof(double));
    if (!p || !ds || !dq || !dk || !dv) return -1;
    for (int g = 0; g < G; ++g) {
        const float *Kg = K + (size_t)g * L * D, *Vg = V + (size_t)g * L * D;
        memset(dk, 0, sizeof(double) * L * D);
        memset(dv, 0, sizeof(double) * L * D);
        for (int i = 0; i < L; ++i) {
            const float *q_i = Q + ((size_t)g * L + i) * D;
            const float *do_i = grad + ((size_t)g * L + i) * D;
            softmax_row(q_i, Kg, i, L, D, causal, p);
            /* dV += P^T dO (:110); dP = dO V^T (:111); dP *= P (:114); s = rowsum (:115) */
            double s = 0.0;
            for (int j = 0; j < L; ++j) {
                const float *v_j = Vg + (size_t)j * D;
                double dp = 0.0;
                for (int d = 0; d < D; ++d) {
                    dv[(size_t)j * D + d] += p[j] * (double)do_i[d];
                    dp += (double)do_i[d] * (double)v_j[d];
                }
                ds[j] = dp * p[j];
                s += ds[j];
            }
            /* dS = P*dP - P*s (:116-117); dQ = dS K (:120); dK = dS^T Q (:122) */
            memset(dq, 0, sizeof(double) * D);
            for (int j = 0; j < L; ++j) {
                double e = ds[j] - p[j] * s;
                const float *k_j = Kg + (size_t)j * D;
                for (int d = 0; d < D; ++d) {
                    dq[d] += e * (double)k_j[d];
                    dk[(size_t)j * D + d] += e * (double)q_i[d];
                }
            }
            float *dq_i = dqkv + ((size_t)g * L + i) * D;
            for (int d = 0; d < D; ++d) dq_i[d] = (float)dq[d];
        }
        float *dKg = dqkv + GLD + (size_t)g * L * D, *dVg = dqkv + 2 * GLD + (size_t)g * L * D;
        for (size_t t = 0; t < (size_t)L * D; ++t) { dKg[t] = (float)dk[t]; dVg[t] = (float)dv[t]; }
    }
    free(p); free(ds); free(dq); free(dk); free(dv);
    return 0;
}
