/*
 * irbfn_oracle.c -- plain-C, float32, multi-threaded CPU restatement of the IRBFN hot path.
 * TEST INFRASTRUCTURE ONLY (the checker and the timed CPU baseline): only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product path (irbfn_b200/) never does.
 *
 * PARITY UNPINNED by reference tests (the reference has none for this path);
 * pinned against the NumPy oracle and the committed golden fixtures in tests/test_oracle_c.py.
 *
 * Reference lines followed (paths relative to the reference checkout):
 *   gate      src/irbfn_mpc/model.py:70-95    (tanh form, product over split dims)
 *   layer     deprecated/f1tenth_gym/examples/flax_rbf/flax_rbf/flax_rbf.py:275-283
 *             (sqrt of the squared distance, divided by exp(log_sig))
 *   basis     flax_rbf.py:34-37 (gaussian), :49-52 (inverse_quadratic)
 *   blend     src/irbfn_mpc/model.py:187-193   readout  src/irbfn_mpc/model.py:196
 *   VJP       what jax.value_and_grad of those lines returns (scripts/train_nmpc.py:297),
 *             written on d^2 (finite at x == centre)
 *
 * Build: gcc -O3 -march=native -fopenmp -shared -fPIC oracle/irbfn_oracle.c -o oracle/_build/libirbfn_oracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXD 16
#define MAXB 256 /* max gate bounds */

typedef struct {
    int S, num_ranges, num_bounds;
    const float* lower;
    const float* upper;
    const int32_t* offset; /* S+1 */
    const float* delta;    /* S   */
    const int32_t* ranges; /* num_ranges x S */
} OracleGate;

static float basis_of(float alpha, int basis) {
    if (basis == 0) return expf(-1.0f * alpha * alpha); /* flax_rbf.py:36 */
    return 1.0f / (1.0f + alpha * alpha);               /* flax_rbf.py:51 */
}

/* gamma[r] for one query (model.py:70-95) */
static void gate_one(const float* x, const OracleGate* g, int R, float* fac, float* gamma) {
    for (int d = 0; d < g->S; ++d)
        for (int i = g->offset[d]; i < g->offset[d + 1]; ++i) {
            float lo = x[d] - g->lower[i], up = g->upper[i] - x[d];
            fac[i] = ((tanhf(g->delta[d] * lo) + 1.0f) / 2.0f) * ((tanhf(g->delta[d] * up) + 1.0f) / 2.0f);
        }
    for (int r = 0; r < R; ++r) {
        if (r >= g->num_ranges) { gamma[r] = 0.0f; continue; }
        float cur = fac[g->offset[0] + g->ranges[r * g->S]];
        for (int d = 1; d < g->S; ++d) cur *= fac[g->offset[d] + g->ranges[r * g->S + d]];
        gamma[r] = cur;
    }
}

int irbfn_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* y (B,O) and/or hidden (B,K); gate_ext (B,R) or the tanh gate g. Returns 0 / -1. */
int irbfn_oracle_forward_f32(const float* x, const float* centers, const float* log_sigs,
                             const float* W, const float* bias, const float* gate_ext,
                             const OracleGate* g, float* y, float* hidden, int64_t B, int R, int K,
                             int D, int O, int basis, int nthreads) {
    if (D > MAXD || (g && g->num_bounds > MAXB) || (basis != 0 && basis != 1)) return -1;
    float* sig = (float*)malloc(sizeof(float) * (size_t)R * K);
    for (int i = 0; i < R * K; ++i) sig[i] = expf(log_sigs[i]);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        float* h = (float*)malloc(sizeof(float) * K);
        float* gamma = (float*)malloc(sizeof(float) * R);
        float fac[MAXB];
#pragma omp for schedule(static)
        for (int64_t b = 0; b < B; ++b) {
            const float* xb = x + b * D;
            if (gate_ext) memcpy(gamma, gate_ext + b * R, sizeof(float) * R);
            else gate_one(xb, g, R, fac, gamma);
            for (int k = 0; k < K; ++k) h[k] = 0.0f;
            for (int r = 0; r < R; ++r) {
                const float* c = centers + (size_t)r * K * D;
                for (int k = 0; k < K; ++k) {
                    float acc = 0.0f;
                    for (int j = 0; j < D; ++j) {
                        float t = xb[j] - c[k * D + j];
                        acc += t * t;
                    }
                    float alpha = sqrtf(acc) / sig[r * K + k]; /* flax_rbf.py:280 */
                    h[k] += gamma[r] * basis_of(alpha, basis); /* model.py:193    */
                }
            }
            if (hidden) memcpy(hidden + b * K, h, sizeof(float) * K);
            if (y) {
                for (int o = 0; o < O; ++o) {
                    float acc = 0.0f;
                    for (int k = 0; k < K; ++k) acc += h[k] * W[k * O + o];
                    y[b * O + o] = acc + bias[o]; /* model.py:196 */
                }
            }
        }
        free(h);
        free(gamma);
    }
    free(sig);
    return 0;
}

/* VJP for cotangent dy (B,O): dcenters (R,K,D), dlog_sigs (R,K), dW (K,O), dbias (O), and
 * optionally dx (B,D) (RBF-layer term + tanh-gate term).  Per-thread partials are summed in a
 * fixed order.  Any output pointer may be NULL. */
int irbfn_oracle_backward_f32(const float* x, const float* centers, const float* log_sigs,
                              const float* W, const float* gate_ext, const OracleGate* g,
                              const float* dy, float* dcenters, float* dlog_sigs, float* dW,
                              float* dbias, float* dx, int64_t B, int R, int K, int D, int O,
                              int basis, int nthreads) {
    if (D > MAXD || (g && g->num_bounds > MAXB) || (basis != 0 && basis != 1)) return -1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    const int T = omp_get_max_threads();
#else
    const int T = 1;
#endif
    const size_t nC = (size_t)R * K * D, nL = (size_t)R * K, nW = (size_t)K * O;
    const size_t per = nC + nL + nW + (size_t)O;
    double* part = (double*)calloc(per * T, sizeof(double));
    float* einv = (float*)malloc(sizeof(float) * nL);
    for (size_t i = 0; i < nL; ++i) einv[i] = expf(-2.0f * log_sigs[i]);
#pragma omp parallel num_threads(T)
    {
#ifdef _OPENMP
        const int tid = omp_get_thread_num();
#else
        const int tid = 0;
#endif
        double* pC = part + per * tid;
        double* pL = pC + nC;
        double* pW = pL + nL;
        double* pB = pW + nW;
        float* h = (float*)malloc(sizeof(float) * K);
        float* dh = (float*)malloc(sizeof(float) * K);
        float* gamma = (float*)malloc(sizeof(float) * R);
        float* dgamma = (float*)malloc(sizeof(float) * R);
        float fac[MAXB], dfac[MAXB];
#pragma omp for schedule(static)
        for (int64_t b = 0; b < B; ++b) {
            const float* xb = x + b * D;
            if (gate_ext) memcpy(gamma, gate_ext + b * R, sizeof(float) * R);
            else gate_one(xb, g, R, fac, gamma);
            for (int k = 0; k < K; ++k) {
                float a = 0.0f;
                for (int o = 0; o < O; ++o) a += dy[b * O + o] * W[k * O + o];
                dh[k] = a;
                h[k] = 0.0f;
            }
            float dxl[MAXD];
            for (int j = 0; j < D; ++j) dxl[j] = 0.0f;
            for (int r = 0; r < R; ++r) {
                const float* c = centers + (size_t)r * K * D;
                float dg = 0.0f;
                for (int k = 0; k < K; ++k) {
                    float diff[MAXD], d2 = 0.0f;
                    for (int j = 0; j < D; ++j) { diff[j] = xb[j] - c[k * D + j]; d2 += diff[j] * diff[j]; }
                    const float e = einv[r * K + k], a2 = d2 * e;
                    float phi, dphi;
                    if (basis == 0) { phi = expf(-a2); dphi = -phi; }
                    else { phi = 1.0f / (1.0f + a2); dphi = -phi * phi; }
                    h[k] += gamma[r] * phi;
                    dg += dh[k] * phi;
                    const float da2 = dh[k] * gamma[r] * dphi, dd2 = da2 * e;
                    for (int j = 0; j < D; ++j) {
                        pC[((size_t)r * K + k) * D + j] += (double)(-2.0f * dd2 * diff[j]);
                        dxl[j] += 2.0f * dd2 * diff[j];
                    }
                    pL[(size_t)r * K + k] += (double)(-2.0f * da2 * a2);
                }
                dgamma[r] = dg;
            }
            for (int k = 0; k < K; ++k)
                for (int o = 0; o < O; ++o) pW[(size_t)k * O + o] += (double)(h[k] * dy[b * O + o]);
            for (int o = 0; o < O; ++o) pB[o] += (double)dy[b * O + o];
            if (dx) {
                if (!gate_ext) { /* gate's own dependence on x */
                    for (int d = 0; d < g->S; ++d)
                        for (int i = g->offset[d]; i < g->offset[d + 1]; ++i) {
                            float tl = tanhf(g->delta[d] * (xb[d] - g->lower[i]));
                            float tu = tanhf(g->delta[d] * (g->upper[i] - xb[d]));
                            float sl = (tl + 1.0f) / 2.0f, su = (tu + 1.0f) / 2.0f;
                            fac[i] = sl * su;
                            dfac[i] = g->delta[d] * (1.0f - tl * tl) / 2.0f * su - sl * g->delta[d] * (1.0f - tu * tu) / 2.0f;
                        }
                    for (int r = 0; r < R && r < g->num_ranges; ++r)
                        for (int d = 0; d < g->S; ++d) {
                            float prod = dfac[g->offset[d] + g->ranges[r * g->S + d]];
                            for (int dd = 0; dd < g->S; ++dd)
                                if (dd != d) prod *= fac[g->offset[dd] + g->ranges[r * g->S + dd]];
                            dxl[d] += dgamma[r] * prod;
                        }
                }
                for (int j = 0; j < D; ++j) dx[b * D + j] = dxl[j];
            }
        }
        free(h); free(dh); free(gamma); free(dgamma);
    }
    for (size_t i = 0; i < per; ++i) {
        double s = 0.0;
        for (int t = 0; t < T; ++t) s += part[per * t + i];
        float v = (float)s;
        if (i < nC) { if (dcenters) dcenters[i] = v; }
        else if (i < nC + nL) { if (dlog_sigs) dlog_sigs[i - nC] = v; }
        else if (i < nC + nL + nW) { if (dW) dW[i - nC - nL] = v; }
        else if (dbias) dbias[i - nC - nL - nW] = v;
    }
    free(part);
    free(einv);
    return 0;
}
