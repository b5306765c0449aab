// IRBFN backward, query-side gradients (sm_100a): dx (B,D) and dgate (B,R).
//
// Reference: no call site differentiates w.r.t. the query states; the definition
// is autodiff of the forward, model.py:169-198 + flax_rbf.py:275-283, written on d^2 so that
// x == centre yields the finite analytic limit (F8).
//
// Design ("backward kernel X"): the forward's layout -- query-stationary, Q = 4
// queries per thread as two packed pairs, parameters broadcast from shared-memory records
// [-c_0,-c_0,...,-c_{D-1},-c_{D-1}, s,s, m,m] with m = -2 exp(-2 log_sig).  With
//   w = dh_k * gamma_r * dphi/d(alpha^2) * 2 exp(-2 ls)      (per query, region, kernel)
// the input gradient is dx_j = x_j * sum(w) + sum(w * (-c_j)), so only sum(w) and D packed
// accumulators are carried and the differences need not stay live.  dh_k = dy . W[k,:] (+dhidden)
// is formed per kernel tile in registers; d gamma_r = sum_k dh_k phi_rk goes to a thread-private
// shared-memory column and is either the dgate output (external gate) or, for the tanh gate, is
// written to scratch and folded into dx by gate_bwd_kernel.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct BwdXArgs {
    const float* x;         // (B,Dsrc)
    const float* centers;   // (R,K,Dsrc)
    const float* log_sigs;  // (R,K)
    const float* W;         // (K,O) or nullptr
    const float* gate_ext;  // (B,R) or nullptr -> fused tanh gate from gd
    GateDev gd;
    const float* dy;        // (B,O) or nullptr
    const float* dhidden;   // (B,K) or nullptr
    float* dx;              // (B,Dsrc) or nullptr
    float* dgate;           // (B,R) or nullptr
    long long B;
    int R, K, Dsrc, O;
    int RC, KC;
    int num_tiles;
    int stage_once;
};

struct BwdXShape {
    static constexpr int NT = 256, Q = 4, TQ = NT * Q, KT = 4;
    static __host__ __device__ constexpr int rec_floats(int D) { return ((2 * D + 4 + 3) / 4) * 4; }
};

template <int D, int OT, int BASIS>
__global__ void __launch_bounds__(BwdXShape::NT, 1) irbfn_bwd_x_kernel(const BwdXArgs a) {
    constexpr int NT = BwdXShape::NT, Q = BwdXShape::Q, TQ = BwdXShape::TQ, KT = BwdXShape::KT;
    constexpr int RS = BwdXShape::rec_floats(D);
    constexpr int OTS = OT > 0 ? OT : 1;

    extern __shared__ __align__(16) float smem[];
    float* stage = smem;                                       // [RC][KC][RS]
    float* Ws = stage + (size_t)a.RC * a.KC * RS;              // [KC][2*OT]  (w,w) duplicated
    float* gtab = Ws + (size_t)a.KC * 2 * (OT > 0 ? OT : 0);   // [RC][TQ]
    float* dgtab = gtab + (size_t)a.RC * TQ;                   // [RC][TQ]

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, RC = a.RC, KC = a.KC, O = a.O;
    const int n_rch = (R + RC - 1) / RC;
    const int n_kch = (K + KC - 1) / KC;
    const bool want_dg = a.dgate != nullptr;
    bool staged = false;

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const long long q0 = (long long)tile * TQ + (long long)tid * Q;

        float2 x2[2][D];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                x2[p][j].x = (qa < a.B && j < a.Dsrc) ? __ldg(a.x + qa * a.Dsrc + j) : 0.f;
                x2[p][j].y = (qb < a.B && j < a.Dsrc) ? __ldg(a.x + qb * a.Dsrc + j) : 0.f;
            }
        }
        // cotangent rows, packed over the query pair
        float2 dy2[2][OTS];
        if (OT > 0) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
                for (int o = 0; o < OT; ++o) {
                    dy2[p][o].x = (qa < a.B && o < O) ? __ldg(a.dy + qa * O + o) : 0.f;
                    dy2[p][o].y = (qb < a.B && o < O) ? __ldg(a.dy + qb * O + o) : 0.f;
                }
            }
        }

        float2 sw[2], dxn[2][D];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            sw[p] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < D; ++j) dxn[p][j] = make_float2(0.f, 0.f);
        }

        for (int rch = 0; rch < n_rch; ++rch) {
            const int r0 = rch * RC;
            const int rc = min(RC, R - r0);
            float4* gcol = reinterpret_cast<float4*>(gtab) + tid;
            float4* dgcol = reinterpret_cast<float4*>(dgtab) + tid;

            // ---- gate columns (thread-private) ----
            if (a.gate_ext != nullptr) {
                for (int rr = 0; rr < rc; ++rr) {
                    float4 g;
                    g.x = (q0 + 0 < a.B) ? __ldg(a.gate_ext + (q0 + 0) * R + r0 + rr) : 0.f;
                    g.y = (q0 + 1 < a.B) ? __ldg(a.gate_ext + (q0 + 1) * R + r0 + rr) : 0.f;
                    g.z = (q0 + 2 < a.B) ? __ldg(a.gate_ext + (q0 + 2) * R + r0 + rr) : 0.f;
                    g.w = (q0 + 3 < a.B) ? __ldg(a.gate_ext + (q0 + 3) * R + r0 + rr) : 0.f;
                    gcol[rr * (TQ / 4)] = g;
                }
            } else {
                const int S = a.gd.S;
                for (int rr = 0; rr < rc; ++rr) {
                    const float one = (r0 + rr < a.gd.num_ranges) ? 1.f : 0.f;
                    gcol[rr * (TQ / 4)] = make_float4(one, one, one, one);
                }
                const int rlim = min(rc, a.gd.num_ranges - r0);
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    if (j < S) {
                        const float k2 = 2.0f * kLog2e * __ldg(a.gd.delta + j);
                        const int ob = __ldg(a.gd.offset + j), oe = __ldg(a.gd.offset + j + 1);
                        for (int i = ob; i < oe; ++i) {
                            const float lb = __ldg(a.gd.lower + i), ub = __ldg(a.gd.upper + i);
                            float4 f;
                            f.x = gate_sigmoid(k2, x2[0][j].x - lb) * gate_sigmoid(k2, ub - x2[0][j].x);
                            f.y = gate_sigmoid(k2, x2[0][j].y - lb) * gate_sigmoid(k2, ub - x2[0][j].y);
                            f.z = gate_sigmoid(k2, x2[1][j].x - lb) * gate_sigmoid(k2, ub - x2[1][j].x);
                            f.w = gate_sigmoid(k2, x2[1][j].y - lb) * gate_sigmoid(k2, ub - x2[1][j].y);
                            for (int rr = 0; rr < rlim; ++rr) {
                                if (__ldg(a.gd.ranges + (r0 + rr) * S + j) == i - ob) {
                                    float4 g = gcol[rr * (TQ / 4)];
                                    g.x *= f.x; g.y *= f.y; g.z *= f.z; g.w *= f.w;
                                    gcol[rr * (TQ / 4)] = g;
                                }
                            }
                        }
                    }
                }
            }
            if (want_dg)
                for (int rr = 0; rr < rc; ++rr) dgcol[rr * (TQ / 4)] = make_float4(0.f, 0.f, 0.f, 0.f);

            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kch * KC;
                const int kc = min(KC, K - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;

                if (!(a.stage_once && staged)) {
                    __syncthreads();
                    for (int idx = tid; idx < rc * kcp; idx += NT) {
                        const int rr = idx / kcp, kk = idx - rr * kcp;
                        float* rec = stage + ((size_t)rr * KC + kk) * RS;
                        if (kk < kc) {
                            const size_t rk = (size_t)(r0 + rr) * K + (k0 + kk);
#pragma unroll
                            for (int j = 0; j < D; ++j) {
                                const float c = j < a.Dsrc ? -__ldg(a.centers + rk * a.Dsrc + j) : 0.f;
                                rec[2 * j] = c;
                                rec[2 * j + 1] = c;
                            }
                            const float ls = __ldg(a.log_sigs + rk);
                            const float s = width_scale<BASIS>(ls);
                            const float m = -2.0f * expf(-2.0f * ls);
                            rec[2 * D] = s; rec[2 * D + 1] = s;
                            rec[2 * D + 2] = m; rec[2 * D + 3] = m;
                        } else {
#pragma unroll
                            for (int j = 0; j < 2 * D + 4; ++j) rec[j] = 0.f;
                        }
#pragma unroll
                        for (int j = 2 * D + 4; j < RS; ++j) rec[j] = 0.f;
                    }
                    if (OT > 0) {
                        for (int idx = tid; idx < kcp * OTS; idx += NT) {
                            const int kk = idx / OTS, o = idx - kk * OTS;
                            const float w = (kk < kc && o < O) ? __ldg(a.W + (size_t)(k0 + kk) * O + o) : 0.f;
                            Ws[2 * idx] = w;
                            Ws[2 * idx + 1] = w;
                        }
                    }
                    __syncthreads();
                    staged = true;
                }

                for (int kt = 0; kt < kcp; kt += KT) {
                    // ---- dh for the 4 kernels of the tile x 2 query pairs ----
                    float2 dh2[KT][2];
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        dh2[kk][0] = dh2[kk][1] = make_float2(0.f, 0.f);
                        if (a.dhidden != nullptr) {
                            const int k = k0 + kt + kk;
                            if (k < K) {
#pragma unroll
                                for (int p = 0; p < 2; ++p) {
                                    const long long qa = q0 + 2 * p, qb = qa + 1;
                                    dh2[kk][p].x = qa < a.B ? __ldg(a.dhidden + qa * K + k) : 0.f;
                                    dh2[kk][p].y = qb < a.B ? __ldg(a.dhidden + qb * K + k) : 0.f;
                                }
                            }
                        }
                        if (OT > 0) {
                            const float2* w2 = reinterpret_cast<const float2*>(Ws + (size_t)(kt + kk) * 2 * OTS);
#pragma unroll
                            for (int o = 0; o < OT; ++o) {
                                const float2 wv = w2[o];
                                dh2[kk][0] = fma2(dy2[0][o], wv, dh2[kk][0]);
                                dh2[kk][1] = fma2(dy2[1][o], wv, dh2[kk][1]);
                            }
                        }
                    }

                    for (int rr = 0; rr < rc; ++rr) {
                        const float4 g4 = gcol[rr * (TQ / 4)];
                        const float2 g2[2] = {make_float2(g4.x, g4.y), make_float2(g4.z, g4.w)};
                        float2 dg2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
                        const float4* rec4 =
                            reinterpret_cast<const float4*>(stage + ((size_t)rr * KC + kt) * RS);
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
                            float v[RS];
#pragma unroll
                            for (int i = 0; i < RS / 4; ++i) {
                                const float4 t = rec4[kk * (RS / 4) + i];
                                v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
                            }
                            const float2 sc = make_float2(v[2 * D], v[2 * D + 1]);
                            const float2 m2 = make_float2(v[2 * D + 2], v[2 * D + 3]);
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                float2 t = add2(x2[p][0], make_float2(v[0], v[1]));
                                float2 acc = mul2(t, t);
#pragma unroll
                                for (int j = 1; j < D; ++j) {
                                    t = add2(x2[p][j], make_float2(v[2 * j], v[2 * j + 1]));
                                    acc = fma2(t, t, acc);
                                }
                                float2 phi;
                                if (BASIS == IRBFN_BASIS_GAUSSIAN) {
                                    const float2 e = mul2(acc, sc);
                                    phi = make_float2(ex2_approx(e.x), ex2_approx(e.y));
                                } else {
                                    const float2 e = fma2(acc, sc, make_float2(1.f, 1.f));
                                    phi = make_float2(rcp_approx(e.x), rcp_approx(e.y));
                                }
                                dg2[p] = fma2(dh2[kk][p], phi, dg2[p]);
                                // w = dh * gamma * dphi/d(alpha^2) * 2 e  = dh gamma phi (-2e)  [* phi for IQ]
                                float2 w = mul2(mul2(dh2[kk][p], g2[p]), mul2(phi, m2));
                                if (BASIS == IRBFN_BASIS_INVERSE_QUADRATIC) w = mul2(w, phi);
                                sw[p] = add2(sw[p], w);
#pragma unroll
                                for (int j = 0; j < D; ++j)
                                    dxn[p][j] = fma2(w, make_float2(v[2 * j], v[2 * j + 1]), dxn[p][j]);
                            }
                        }
                        if (want_dg) {
                            float4 d = dgcol[rr * (TQ / 4)];
                            d.x += dg2[0].x; d.y += dg2[0].y; d.z += dg2[1].x; d.w += dg2[1].y;
                            dgcol[rr * (TQ / 4)] = d;
                        }
                    }
                }
            }

            if (want_dg) {
                for (int rr = 0; rr < rc; ++rr) {
                    const float4 d = dgcol[rr * (TQ / 4)];
                    if (q0 + 0 < a.B) a.dgate[(q0 + 0) * R + r0 + rr] = d.x;
                    if (q0 + 1 < a.B) a.dgate[(q0 + 1) * R + r0 + rr] = d.y;
                    if (q0 + 2 < a.B) a.dgate[(q0 + 2) * R + r0 + rr] = d.z;
                    if (q0 + 3 < a.B) a.dgate[(q0 + 3) * R + r0 + rr] = d.w;
                }
            }
        }

        // ---- dx_j = x_j * sum(w) + sum(w * -c_j) ----
        if (a.dx != nullptr) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    const float2 r = fma2(x2[p][j], sw[p], dxn[p][j]);
                    if (j < a.Dsrc) {
                        if (qa < a.B) a.dx[qa * a.Dsrc + j] = r.x;
                        if (qb < a.B) a.dx[qb * a.Dsrc + j] = r.y;
                    }
                }
            }
        }
    }
}

template <int D, int OT>
inline int launch_bwd_x_ot(const BwdXArgs& a, int basis, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e;
    if (basis == IRBFN_BASIS_GAUSSIAN) {
        auto kern = irbfn_bwd_x_kernel<D, OT, IRBFN_BASIS_GAUSSIAN>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, BwdXShape::NT, smem, st>>>(a);
    } else {
        auto kern = irbfn_bwd_x_kernel<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, BwdXShape::NT, smem, st>>>(a);
    }
    return (int)cudaGetLastError();
}

template <int D>
inline int launch_bwd_x(const BwdXArgs& a, int OT, int basis, int grid, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return launch_bwd_x_ot<D, 0>(a, basis, grid, smem, st);
        case 2: return launch_bwd_x_ot<D, 2>(a, basis, grid, smem, st);
        case 4: return launch_bwd_x_ot<D, 4>(a, basis, grid, smem, st);
        case 8: return launch_bwd_x_ot<D, 8>(a, basis, grid, smem, st);
        case 10: return launch_bwd_x_ot<D, 10>(a, basis, grid, smem, st);
        case 16: return launch_bwd_x_ot<D, 16>(a, basis, grid, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
