// IRBFN per-query Jacobian dy/dx (B,O,D) in ONE pass (sm_100a) -- SURVEY.md 8f rank 2.
//
// Reference: no call site; the definition is jax.jacrev / jax.jacfwd of
// WCRBFNet.__call__ (src/irbfn_mpc/model.py:169-198 with flax_rbf.py:275-283), i.e. what the
// paper's differentiable trajectory generation needs.  jacrev over the VJP costs O backward passes;
// this kernel produces all O rows (and, optionally, y itself) from one evaluation of the activations.
//
//   J[b,o,j] = sum_r gamma_r sum_k W[k,o] dphi_rk/dx_j  +  sum_r (dgamma_r/dx_j) sum_k W[k,o] phi_rk
//   dphi/dx_j = phi' * 2 exp(-2 ls) (x_j - c_j),  phi' = -phi (gaussian) | -phi^2 (inverse quadratic)
//   dgamma_r/dx_j = gamma_r * 2 delta_j (sig_up - sig_lo)     for the tanh gate (model.py:70-95)
//
// Design: query-stationary like the forward, one packed PAIR of queries per thread
// so that the O x D packed Jacobian accumulators fit in registers; parameters come from the
// forward's transposed 4-kernel shared-memory tiles (rows -c_0..-c_{D-1}, width scale) as scalar
// broadcast operands.  Loop order is region outer, kernel inner: the per-region readout
// yr[o] = sum_k W[k,o] phi_rk is complete when the region ends and carries the gate term.  Written on
// d^2, so x == centre gives the finite limit (the reference's sqrt form gives NaN, F8).
// FP32-pipe bound: per (query, region, kernel) 2D + 3 + OT (D + 2) lane-operations and one MUFU.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct JacArgs {
    const float* x;         // (B,Dsrc)
    const float* centers;   // (R,K,Dsrc)
    const float* log_sigs;  // (R,K)
    const float* W;         // (K,O)
    const float* bias;      // (O) (only for y)
    const float* gate_ext;  // (B,R) or nullptr -> fused tanh gate from gd
    GateDev gd;
    float* jac;             // (B,O,Dsrc)
    float* y;               // (B,O) or nullptr
    float* region_y;        // (B,R,O) or nullptr: yr[o] = sum_k W[k,o] phi_rk (external-gate chain rule)
    long long B;
    int R, K, Dsrc, O;
    int o0, ocur;           // output column window of this launch
    int RC, KC;             // regions / kernels per stage; RC == R (everything resident) or RC == 1
    int num_tiles;
    int stage_once;
    // ---- read by the SPARSE instantiations only (irbfn_jacobian_sparse) ----
    const int* perm;        // (B) queries grouped by their set of non-negligible gate intervals
    const float* edge;      // (num_bounds, 2): the open interval on which each gate factor exceeds gate_tol
};

struct JacShape {
    static constexpr int NT = 256, Q = 2, TQ = NT * Q, KT = 4;
};

// SPARSE = 1 (irbfn_jacobian_sparse; tanh gate, resident parameters, at most 64 gate intervals): a (query, region)
// pair is dropped when one of the region's interval factors is <= gate_tol for the query (then gamma <= gate_tol, and
// every term of J and y that the pair contributes carries the factor gamma).  The test is a bit mask per query, built
// once per tile from the interval edges, so a dropped region costs neither its gate nor its kernels; queries are
// visited in the binned order, so a warp's 64 queries drop the same regions.
template <int D, int OT, int BASIS, int SPARSE = 0>
__global__ void __launch_bounds__(JacShape::NT) irbfn_jac_kernel(const JacArgs a) {
    constexpr int NT = JacShape::NT, TQ = JacShape::TQ, KT = JacShape::KT;
    extern __shared__ __align__(16) float smem[];
    const int KCT = a.KC / KT;
    float4* stage4 = reinterpret_cast<float4*>(smem);                     // [RC][KCT][D+1] float4
    float* stage = smem;
    float* Ws = smem + (size_t)a.RC * KCT * (D + 1) * 4;                  // [KC][OT]
    float2* dlg = reinterpret_cast<float2*>(Ws + (size_t)a.KC * OT);      // [D][NT] thread-private columns

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, KC = a.KC;
    const int n_kch = (K + KC - 1) / KC;
    const int S = a.gate_ext != nullptr ? 0 : a.gd.S;
    bool staged = false;
    // phi' * 2 exp(-2 ls) from the staged width scale: gaussian row = -s log2e, inverse quadratic row = s
    const float mfac = BASIS == IRBFN_BASIS_GAUSSIAN ? 2.0f / kLog2e : -2.0f;

    // parameters of regions [r_lo, r_lo + nr), kernels [k0, k0 + kc) -> shared memory, transposed per 4-kernel tile
    auto do_stage = [&](int r_lo, int nr, int k0, int kc, int kcp) {
        __syncthreads();
        for (int idx = tid; idx < nr * kcp; idx += NT) {
            const int sr = idx / kcp, kk = idx - sr * kcp;
            float* t = stage + ((size_t)sr * KCT + (kk >> 2)) * ((D + 1) * 4) + (kk & 3);
            if (kk < kc) {
                const size_t rk = (size_t)(r_lo + sr) * K + (k0 + kk);
#pragma unroll
                for (int j = 0; j < D; ++j)
                    t[4 * j] = j < a.Dsrc ? -__ldg(a.centers + rk * a.Dsrc + j) : 0.f;
                t[4 * D] = width_scale<BASIS>(__ldg(a.log_sigs + rk));
            } else {
#pragma unroll
                for (int j = 0; j <= D; ++j) t[4 * j] = 0.f;
            }
        }
        for (int idx = tid; idx < kcp * OT; idx += NT) {
            const int kk = idx / OT, o = idx - kk * OT;
            Ws[idx] = (kk < kc && o < a.ocur) ? __ldg(a.W + (size_t)(k0 + kk) * a.O + a.o0 + o) : 0.f;
        }
        __syncthreads();
    };
    if (SPARSE) {
        // resident parameters are staged BEFORE any warp can skip a region: the staging barriers are CTA-wide
        do_stage(0, R, 0, K, ((K + KT - 1) / KT) * KT);
        staged = true;
    }

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        long long qa = (long long)tile * TQ + 2 * tid, qb = qa + 1;
        if (SPARSE) {   // rows in the binned order; a row index >= B means "no query"
            qa = qa < a.B ? (long long)__ldg(a.perm + qa) : a.B;
            qb = qb < a.B ? (long long)__ldg(a.perm + qb) : a.B;
        }
        float2 x2[D];
#pragma unroll
        for (int j = 0; j < D; ++j) {
            x2[j].x = (qa < a.B && j < a.Dsrc) ? __ldg(a.x + qa * a.Dsrc + j) : 0.f;
            x2[j].y = (qb < a.B && j < a.Dsrc) ? __ldg(a.x + qb * a.Dsrc + j) : 0.f;
        }
        unsigned long long sigx = 0ull, sigy = 0ull;   // SPARSE: bit i = interval factor i exceeds gate_tol
        if (SPARSE) {
#pragma unroll
            for (int j = 0; j < D; ++j) {
                if (j < S) {
                    for (int i = __ldg(a.gd.offset + j); i < __ldg(a.gd.offset + j + 1); ++i) {
                        const float lo = __ldg(a.edge + 2 * i), hi = __ldg(a.edge + 2 * i + 1);
                        if (qa < a.B && x2[j].x > lo && x2[j].x < hi) sigx |= 1ull << i;
                        if (qb < a.B && x2[j].y > lo && x2[j].y < hi) sigy |= 1ull << i;
                    }
                }
            }
        }
        float2 J[OT][D], y2[OT];
#pragma unroll
        for (int o = 0; o < OT; ++o) {
            const float bv = (a.y != nullptr && o < a.ocur) ? __ldg(a.bias + a.o0 + o) : 0.f;
            y2[o] = dup2(bv);
#pragma unroll
            for (int j = 0; j < D; ++j) J[o][j] = dup2(0.f);
        }

        for (int r = 0; r < R; ++r) {
            // ---- gate of region r for this thread's two queries, and d log gamma / d x_j ----
            float2 g2;
            if (a.gate_ext != nullptr) {
                g2.x = qa < a.B ? __ldg(a.gate_ext + qa * R + r) : 0.f;
                g2.y = qb < a.B ? __ldg(a.gate_ext + qb * R + r) : 0.f;
            } else {
                if (r >= a.gd.num_ranges) continue;      // gamma == 0 (uniform across the CTA)
                bool ax = true, ay = true;
                if (SPARSE) {
                    unsigned long long req = 0ull;
                    for (int j = 0; j < S; ++j) req |= 1ull << (__ldg(a.gd.offset + j) + __ldg(a.gd.ranges + r * S + j));
                    ax = (sigx & req) == req;
                    ay = (sigy & req) == req;
                    if (!__any_sync(0xffffffffu, ax || ay)) continue;   // warp-uniform; no barrier inside (resident parameters)
                }
                g2 = dup2(1.f);
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    if (j < S) {
                        const float dl = __ldg(a.gd.delta + j);
                        const float k2 = 2.0f * kLog2e * dl;
                        const int i = __ldg(a.gd.offset + j) + __ldg(a.gd.ranges + r * S + j);
                        const float lb = __ldg(a.gd.lower + i), ub = __ldg(a.gd.upper + i);
                        const float slx = gate_sigmoid(k2, x2[j].x - lb), sux = gate_sigmoid(k2, ub - x2[j].x);
                        const float sly = gate_sigmoid(k2, x2[j].y - lb), suy = gate_sigmoid(k2, ub - x2[j].y);
                        g2.x *= slx * sux;
                        g2.y *= sly * suy;
                        dlg[j * NT + tid] = make_float2(2.f * dl * (sux - slx), 2.f * dl * (suy - sly));
                    }
                }
                if (SPARSE) {   // the drop rule is per query: a query's row does not depend on its warp-mates
                    g2.x = ax ? g2.x : 0.f;
                    g2.y = ay ? g2.y : 0.f;
                }
            }
            float2 yr[OT];
#pragma unroll
            for (int o = 0; o < OT; ++o) yr[o] = dup2(0.f);

            const int rr = a.stage_once ? r : 0;
            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kch * KC;
                const int kc = min(KC, K - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;
                if (!(a.stage_once && staged)) {
                    do_stage(a.stage_once ? 0 : r, a.stage_once ? R : 1, k0, kc, kcp);
                    staged = true;
                }

                const float4* rec = stage4 + (size_t)rr * KCT * (D + 1);
                for (int kt = 0; kt < kcp; kt += KT, rec += (D + 1)) {
                    const float4 s4 = rec[D];
                    const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                    for (int half = 0; half < 2; ++half) {      // two kernels at a time: d stays in registers
                        float2 d[2][D], acc[2];
#pragma unroll
                        for (int j = 0; j < D; ++j) {
                            const float4 c4 = rec[j];
                            const float c0 = half ? c4.z : c4.x, c1 = half ? c4.w : c4.y;
                            d[0][j] = add2(x2[j], dup2(c0));
                            d[1][j] = add2(x2[j], dup2(c1));
                            acc[0] = j == 0 ? mul2(d[0][0], d[0][0]) : fma2(d[0][j], d[0][j], acc[0]);
                            acc[1] = j == 0 ? mul2(d[1][0], d[1][0]) : fma2(d[1][j], d[1][j], acc[1]);
                        }
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            const float s = sc[2 * half + u];
                            float2 ph, wv;
                            if (BASIS == IRBFN_BASIS_GAUSSIAN) {
                                const float2 ea = mul2(acc[u], dup2(s));
                                ph = make_float2(ex2_approx(ea.x), ex2_approx(ea.y));
                                wv = mul2(ph, mul2(g2, dup2(s * mfac)));
                            } else {
                                const float2 ea = fma2(acc[u], dup2(s), dup2(1.f));
                                ph = make_float2(rcp_approx(ea.x), rcp_approx(ea.y));
                                wv = mul2(mul2(ph, ph), mul2(g2, dup2(s * mfac)));
                            }
                            const float* wrow = Ws + (size_t)(kt + 2 * half + u) * OT;
#pragma unroll
                            for (int o = 0; o < OT; ++o) {
                                const float2 w2 = dup2(wrow[o]);
                                yr[o] = fma2(w2, ph, yr[o]);
                                const float2 t = mul2(w2, wv);
#pragma unroll
                                for (int j = 0; j < D; ++j) J[o][j] = fma2(t, d[u][j], J[o][j]);
                            }
                        }
                    }
                }
            }

            // ---- region closes: blend its readout into y, add the gate term to J ----
#pragma unroll
            for (int o = 0; o < OT; ++o) {
                y2[o] = fma2(g2, yr[o], y2[o]);
                if (a.gate_ext == nullptr) {
                    const float2 gy = mul2(g2, yr[o]);
#pragma unroll
                    for (int j = 0; j < D; ++j)
                        if (j < S) J[o][j] = fma2(gy, dlg[j * NT + tid], J[o][j]);
                }
            }
            if (a.region_y != nullptr) {
#pragma unroll
                for (int o = 0; o < OT; ++o) {
                    if (o < a.ocur) {
                        if (qa < a.B) a.region_y[(qa * R + r) * a.O + a.o0 + o] = yr[o].x;
                        if (qb < a.B) a.region_y[(qb * R + r) * a.O + a.o0 + o] = yr[o].y;
                    }
                }
            }
        }

        // ---- store ----
#pragma unroll
        for (int o = 0; o < OT; ++o) {
            if (o < a.ocur) {
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    if (j < a.Dsrc) {
                        if (qa < a.B) a.jac[(qa * a.O + a.o0 + o) * a.Dsrc + j] = J[o][j].x;
                        if (qb < a.B) a.jac[(qb * a.O + a.o0 + o) * a.Dsrc + j] = J[o][j].y;
                    }
                }
                if (a.y != nullptr) {
                    if (qa < a.B) a.y[qa * a.O + a.o0 + o] = y2[o].x;
                    if (qb < a.B) a.y[qb * a.O + a.o0 + o] = y2[o].y;
                }
            }
        }
    }
}

}  // namespace irbfn
