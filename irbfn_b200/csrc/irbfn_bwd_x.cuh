// IRBFN backward, query-side gradients (sm_100a): dx (B,D) and dgate (B,R).
//
// Reference: no call site differentiates w.r.t. the query states; the definition
// is autodiff of the forward, model.py:169-198 + flax_rbf.py:275-283, written on d^2 so that
// x == centre yields the finite analytic limit (F8).
//
// Design: the forward's layout -- query-stationary, Q = 4 queries per thread as two
// packed pairs, parameters broadcast from transposed 4-kernel shared-memory tiles with rows
// -c_0 .. -c_{D-1}, s, m  (s the width scale of the basis, m = -2 exp(-2 log_sig)).  With
//   w = dh_k * gamma_r * dphi/d(alpha^2) * 2 exp(-2 ls)      (per query, region, kernel)
// the input gradient is dx_j = x_j * sum(w) + sum(w * (-c_j)), so only sum(w) and D packed
// accumulators are carried and the differences need not stay live.  Each (kernel tile, region) step
// is software-pipelined like the forward: its MUFUs are issued during the next step's distance
// loop, and its w / dx / dgamma updates close that next step.  dh_k = dy . W[k,:] (+dhidden) is formed
// per kernel tile in registers; d gamma_r = sum_k dh_k phi_rk goes to a thread-private shared-memory
// column and is either the dgate output (external gate) or, for the tanh gate, folded into dx at
// the end of the region chunk by the same thread (chain rule through model.py:70-95), so no
// (B,R) scratch and no second kernel are needed for up to 4 split dimensions.
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
    float* dgate;           // (B,R) or nullptr: dgate output, or scratch for gate_bwd_kernel
    long long B;
    int R, K, Dsrc, O;
    int RC, KC;
    int num_tiles;
    int stage_once;
    int fuse_gate;          // 1: tanh gate, S <= 4: d gamma is folded into dx inside this kernel
    // ---- read by the SPARSE instantiations only (irbfn_backward_sparse); behind the dense fields ----
    const int* perm;        // (B): queries are visited in this order (grouped by their set of non-negligible regions)
    float gate_thr;         // a (query, region) pair with gamma <= gate_thr is dropped; a warp skips the regions
                            // that are dropped for all of its 128 queries
};

struct BwdXShape {
    static constexpr int NT = 256, Q = 4, TQ = NT * Q, KT = 4;
    static constexpr int MAX_FUSED_S = 4;
    static __host__ __device__ constexpr int tile_floats(int D) { return (D + 2) * 4; }
};

// SPARSE = 1: gate-sparse, region-binned evaluation exactly as the forward's (irbfn_fwd_modes.cuh, irbfn_sparse.cu):
// rows gathered and results scattered through perm, the drop rule applied per query, the hot loop run over the warp's
// active regions only.  Every dropped term of dx carries the factor gamma (the gate path too: d gamma / dx = gamma *
// d log gamma / dx), so |dx - dx(dense)| is of the order of gate_thr.  Tanh gate, at most 64 regions per stage.
template <int D, int OT, int BASIS, int SPARSE = 0>
__global__ void __launch_bounds__(BwdXShape::NT, 1) irbfn_bwd_x_kernel(const BwdXArgs a) {
    constexpr int NT = BwdXShape::NT, Q = BwdXShape::Q, TQ = BwdXShape::TQ, KT = BwdXShape::KT;
    constexpr int OTS = OT > 0 ? OT : 1;
    constexpr int TR = D + 2;             // rows per tile

    extern __shared__ __align__(16) float smem[];
    const int KCT = a.KC / KT;
    float4* stage4 = reinterpret_cast<float4*>(smem);              // [RC][KCT][D+2] float4
    float* stage = smem;
    float* Ws = smem + (size_t)a.RC * KCT * TR * 4;                // [KC][OT]
    float* gtab = Ws + (size_t)a.KC * (OT > 0 ? OT : 0);           // [RC][NT] float4
    float* dgtab = gtab + (size_t)a.RC * TQ;                       // [RC][NT] float4

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, RC = a.RC, KC = a.KC, O = a.O;
    const int n_rch = (R + RC - 1) / RC;
    const int n_kch = (K + KC - 1) / KC;
    const bool want_dg = a.dgate != nullptr || a.fuse_gate;
    bool staged = false;

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const long long q0 = (long long)tile * TQ + (long long)tid * Q;

        float2 x2[2][D];
        long long qi[Q];                      // the rows this thread evaluates (SPARSE: permuted order), -1 = none
#pragma unroll
        for (int i = 0; i < Q; ++i) {
            if (SPARSE) qi[i] = q0 + i < a.B ? (long long)__ldg(a.perm + q0 + i) : -1ll;
            else qi[i] = q0 + i < a.B ? q0 + i : -1ll;
        }
        if (SPARSE) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    x2[p][j].x = (qi[2 * p] >= 0 && j < a.Dsrc) ? __ldg(a.x + qi[2 * p] * a.Dsrc + j) : 0.f;
                    x2[p][j].y = (qi[2 * p + 1] >= 0 && j < a.Dsrc) ? __ldg(a.x + qi[2 * p + 1] * a.Dsrc + j) : 0.f;
                }
            }
        } else {
            load_queries4<D>(a.x, q0, a.B, a.Dsrc, x2);
        }
        // cotangent rows, packed over the query pair
        float2 dy2[2][OTS];
        if (OT > 0) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const long long qa = SPARSE ? qi[2 * p] : q0 + 2 * p, qb = SPARSE ? qi[2 * p + 1] : qa + 1;
                const bool va = SPARSE ? qa >= 0 : qa < a.B, vb = SPARSE ? qb >= 0 : qb < a.B;
#pragma unroll
                for (int o = 0; o < OT; ++o) {
                    dy2[p][o].x = (va && o < O) ? __ldg(a.dy + qa * O + o) : 0.f;
                    dy2[p][o].y = (vb && o < O) ? __ldg(a.dy + qb * O + o) : 0.f;
                }
            }
        }

        float2 sw[2], dxn[2][D];
        float4 dxg[BwdXShape::MAX_FUSED_S];   // gate-path contribution to dx_j, j < S (fused gate only)
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            sw[p] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < D; ++j) dxn[p][j] = make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int j = 0; j < BwdXShape::MAX_FUSED_S; ++j) dxg[j] = make_float4(0.f, 0.f, 0.f, 0.f);

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
                    gcol[rr * NT] = g;
                }
            } else if (SPARSE) {
                gate_columns_rows(a.gd, a.x, qi, a.Dsrc, r0, rc, gcol, NT);
            } else {
                gate_columns(a.gd, a.x, q0, a.B, a.Dsrc, r0, rc, gcol, NT);
            }
            if (want_dg)
                for (int rr = 0; rr < rc; ++rr) dgcol[rr * NT] = make_float4(0.f, 0.f, 0.f, 0.f);
            // ---- SPARSE: the drop rule per query, then the regions this warp still needs ----
            unsigned long long wmask = 0ull;
            if (SPARSE) {
                for (int rr = 0; rr < rc; ++rr) {
                    float4 g = gcol[rr * NT];
                    g.x = g.x > a.gate_thr ? g.x : 0.f;
                    g.y = g.y > a.gate_thr ? g.y : 0.f;
                    g.z = g.z > a.gate_thr ? g.z : 0.f;
                    g.w = g.w > a.gate_thr ? g.w : 0.f;
                    gcol[rr * NT] = g;
                    const bool act = fmaxf(fmaxf(g.x, g.y), fmaxf(g.z, g.w)) > 0.f;
                    if (__any_sync(0xffffffffu, act)) wmask |= 1ull << rr;
                }
            }

            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kch * KC;
                const int kc = min(KC, K - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;

                // ---- stage parameters, transposed per 4-kernel tile: rows -c_j, s, m ----
                if (!(a.stage_once && staged)) {
                    __syncthreads();
                    for (int idx = tid; idx < rc * kcp; idx += NT) {
                        const int rr = idx / kcp, kk = idx - rr * kcp;
                        float* t = stage + ((size_t)rr * KCT + (kk >> 2)) * (TR * 4) + (kk & 3);
                        if (kk < kc) {
                            const size_t rk = (size_t)(r0 + rr) * K + (k0 + kk);
#pragma unroll
                            for (int j = 0; j < D; ++j)
                                t[4 * j] = j < a.Dsrc ? -__ldg(a.centers + rk * a.Dsrc + j) : 0.f;
                            const float ls = __ldg(a.log_sigs + rk);
                            t[4 * D] = width_scale<BASIS>(ls);
                            t[4 * (D + 1)] = -2.0f * expf(-2.0f * ls);
                        } else {
#pragma unroll
                            for (int j = 0; j < TR; ++j) t[4 * j] = 0.f;
                        }
                    }
                    if (OT > 0) {
                        for (int idx = tid; idx < kcp * OTS; idx += NT) {
                            const int kk = idx / OTS, o = idx - kk * OTS;
                            Ws[idx] = (kk < kc && o < O) ? __ldg(a.W + (size_t)(k0 + kk) * O + o) : 0.f;
                        }
                    }
                    __syncthreads();
                    staged = true;
                }

                // ---- hot loop: (kernel tile, region) steps, software-pipelined by one step ----
                float2 dh2[KT][2], ea[KT][2], ph[KT][2];
                float2 gp[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
                const float4* recp = stage4;   // tile of the previous step
                int rrp = 0;                   // region (within the chunk) of the previous step
                const float kEmpty = BASIS == IRBFN_BASIS_GAUSSIAN ? -INFINITY : INFINITY;  // basis value 0
#pragma unroll
                for (int kk = 0; kk < KT; ++kk) {
                    dh2[kk][0] = dh2[kk][1] = make_float2(0.f, 0.f);
                    ea[kk][0] = ea[kk][1] = make_float2(kEmpty, kEmpty);
                }

                auto mufu = [&](int i) {
                    const int kk = i >> 2, p = (i >> 1) & 1;
                    float& src = (i & 1) ? ea[kk][p].y : ea[kk][p].x;
                    float& dst = (i & 1) ? ph[kk][p].y : ph[kk][p].x;
                    dst = BASIS == IRBFN_BASIS_GAUSSIAN ? ex2_approx(src) : rcp_approx(src);
                };
                // d^2 of one step, feature-outer; the previous step's 16 MUFUs ride between the features
                auto distances = [&](const float4* rec, float2 (&acc)[KT][2]) {
#pragma unroll
                    for (int j = 0; j < D; ++j) {
                        const float4 c4 = rec[j];
                        const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const float2 d = add2(x2[p][j], dup2(c[kk]));
                                acc[kk][p] = j == 0 ? mul2(d, d) : fma2(d, d, acc[kk][p]);
                            }
                        }
#pragma unroll
                        for (int i = (16 * j) / D; i < (16 * (j + 1)) / D; ++i) mufu(i);
                    }
                };
                // close the PREVIOUS step: w = dh gamma phi' (2e), dx and d gamma updates
                auto finish = [&]() {
                    const float4 m4 = recp[D + 1];
                    const float m[4] = {m4.x, m4.y, m4.z, m4.w};
                    float2 wv[KT][2];
                    float2 dg2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
                        for (int p = 0; p < 2; ++p) {
                            dg2[p] = fma2(dh2[kk][p], ph[kk][p], dg2[p]);
                            float2 w = mul2(mul2(dh2[kk][p], gp[p]), mul2(ph[kk][p], dup2(m[kk])));
                            if (BASIS == IRBFN_BASIS_INVERSE_QUADRATIC) w = mul2(w, ph[kk][p]);
                            sw[p] = add2(sw[p], w);
                            wv[kk][p] = w;
                        }
                    }
#pragma unroll
                    for (int j = 0; j < D; ++j) {
                        const float4 c4 = recp[j];
                        const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
                            dxn[0][j] = fma2(wv[kk][0], dup2(c[kk]), dxn[0][j]);
                            dxn[1][j] = fma2(wv[kk][1], dup2(c[kk]), dxn[1][j]);
                        }
                    }
                    if (want_dg) {
                        float4 d = dgcol[rrp * NT];
                        if (SPARSE) {
                            // a (query, region) pair that the drop rule removed contributes nothing to the gate path
                            // either, whether or not the query's warp evaluated the region for other queries: a
                            // query's gradient must not depend on which queries share its warp
                            dg2[0].x = gp[0].x > 0.f ? dg2[0].x : 0.f;
                            dg2[0].y = gp[0].y > 0.f ? dg2[0].y : 0.f;
                            dg2[1].x = gp[1].x > 0.f ? dg2[1].x : 0.f;
                            dg2[1].y = gp[1].y > 0.f ? dg2[1].y : 0.f;
                        }
                        d.x += dg2[0].x; d.y += dg2[0].y; d.z += dg2[1].x; d.w += dg2[1].y;
                        dgcol[rrp * NT] = d;
                    }
                };
                // basis arguments of this step; evaluated during the next step
                auto scale = [&](const float4* rec, const float2 (&acc)[KT][2], const float4 g4, int rr) {
                    const float4 s4 = rec[D];
                    const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
                        for (int p = 0; p < 2; ++p) {
                            if (BASIS == IRBFN_BASIS_GAUSSIAN) ea[kk][p] = mul2(acc[kk][p], dup2(sc[kk]));
                            else ea[kk][p] = fma2(acc[kk][p], dup2(sc[kk]), make_float2(1.f, 1.f));
                        }
                    }
                    gp[0] = make_float2(g4.x, g4.y);
                    gp[1] = make_float2(g4.z, g4.w);
                    recp = rec;
                    rrp = rr;
                };
                // dh of the 4 kernels of tile kt x 2 query pairs
                auto load_dh = [&](int kt) {
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        dh2[kk][0] = dh2[kk][1] = make_float2(0.f, 0.f);
                        if (a.dhidden != nullptr) {
                            const int k = k0 + kt + kk;
                            if (k < K) {
#pragma unroll
                                for (int p = 0; p < 2; ++p) {
                                    const long long qa = qi[2 * p], qb = qi[2 * p + 1];
                                    dh2[kk][p].x = qa >= 0 ? __ldg(a.dhidden + qa * K + k) : 0.f;
                                    dh2[kk][p].y = qb >= 0 ? __ldg(a.dhidden + qb * K + k) : 0.f;
                                }
                            }
                        }
                        if (OT > 0) {
                            const float* wr = Ws + (size_t)(kt + kk) * OTS;
#pragma unroll
                            for (int o = 0; o < OT; ++o) {
                                const float w = wr[o];
                                dh2[kk][0] = fma2(dy2[0][o], dup2(w), dh2[kk][0]);
                                dh2[kk][1] = fma2(dy2[1][o], dup2(w), dh2[kk][1]);
                            }
                        }
                    }
                };

                const size_t rstride = (size_t)KCT * TR;
                if (SPARSE) {
                    // the same pipeline over the warp's active regions only (warp-uniform trip counts; the CTA's
                    // barriers all sit in the staging above, which every warp still joins)
                    if (wmask != 0ull) {
                        const int rr0 = __ffsll((long long)wmask) - 1;
                        const unsigned long long rest = wmask & (wmask - 1ull);
                        for (int kt = 0; kt < kcp; kt += KT) {
                            const float4* rec0 = stage4 + (size_t)(kt >> 2) * TR;
                            {
                                const float4* rec = rec0 + (size_t)rr0 * rstride;
                                const float4 g4 = gcol[rr0 * NT];
                                float2 acc[KT][2];
                                distances(rec, acc);
                                finish();
                                load_dh(kt);
                                scale(rec, acc, g4, rr0);
                            }
                            unsigned long long m = rest;
                            while (m) {
                                const int rr = __ffsll((long long)m) - 1;
                                m &= m - 1ull;
                                const float4* rec = rec0 + (size_t)rr * rstride;
                                const float4 g4 = gcol[rr * NT];
                                float2 acc[KT][2];
                                distances(rec, acc);
                                finish();
                                scale(rec, acc, g4, rr);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 16; ++i) mufu(i);
                        finish();
                    }
                    continue;
                }
                for (int kt = 0; kt < kcp; kt += KT) {
                    const float4* rec = stage4 + (size_t)(kt >> 2) * TR;
                    {   // first region of the tile (peeled): the previous step belongs to tile kt - KT,
                        // so it is closed with the OLD dh before the new one is formed
                        const float4 g4 = gcol[0];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        finish();
                        load_dh(kt);
                        scale(rec, acc, g4, 0);
                    }
                    for (int rr = 1; rr < rc; ++rr) {   // branch-free steady state
                        rec += rstride;
                        const float4 g4 = gcol[rr * NT];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        finish();
                        scale(rec, acc, g4, rr);
                    }
                }
                // drain the pipeline
#pragma unroll
                for (int i = 0; i < 16; ++i) mufu(i);
                finish();
            }

            if (a.fuse_gate) {
                // ---- dx_j += sum_r dgamma_r * d gamma_r / d x_j   (tanh gate, model.py:70-95) ----
                // s(z) = 1/(1+exp(-2 delta z)):  f = s(x-lb) s(ub-x),  f' = 2 delta f (s(ub-x) - s(x-lb))
                const int S = a.gd.S;
                const int rlim = min(rc, a.gd.num_ranges - r0);
                for (int rr = 0; rr < rlim; ++rr) {
                    if (SPARSE && !((wmask >> rr) & 1ull)) continue;   // d gamma of a skipped region is 0
                    const float4 dgv = dgcol[rr * NT];
                    float4 f[BwdXShape::MAX_FUSED_S], fd[BwdXShape::MAX_FUSED_S];
#pragma unroll
                    for (int j = 0; j < BwdXShape::MAX_FUSED_S; ++j) {
                        f[j] = make_float4(1.f, 1.f, 1.f, 1.f);
                        fd[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (j < S) {
                            const float dl = __ldg(a.gd.delta + j);
                            const float k2 = 2.0f * kLog2e * dl;
                            const int i = __ldg(a.gd.offset + j) + __ldg(a.gd.ranges + (r0 + rr) * S + j);
                            const float lb = __ldg(a.gd.lower + i), ub = __ldg(a.gd.upper + i);
                            const float xq[4] = {x2[0][j].x, x2[0][j].y, x2[1][j].x, x2[1][j].y};
                            float fv[4], fdv[4];
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float sl = gate_sigmoid(k2, xq[q] - lb), su = gate_sigmoid(k2, ub - xq[q]);
                                fv[q] = sl * su;
                                fdv[q] = 2.0f * dl * sl * su * (su - sl);
                            }
                            f[j] = make_float4(fv[0], fv[1], fv[2], fv[3]);
                            fd[j] = make_float4(fdv[0], fdv[1], fdv[2], fdv[3]);
                        }
                    }
                    float4 pre[BwdXShape::MAX_FUSED_S];
                    float4 run = make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll
                    for (int j = 0; j < BwdXShape::MAX_FUSED_S; ++j) {
                        pre[j] = run;
                        run.x *= f[j].x; run.y *= f[j].y; run.z *= f[j].z; run.w *= f[j].w;
                    }
                    run = dgv;
#pragma unroll
                    for (int j = BwdXShape::MAX_FUSED_S - 1; j >= 0; --j) {
                        dxg[j].x = fmaf(fd[j].x * pre[j].x, run.x, dxg[j].x);
                        dxg[j].y = fmaf(fd[j].y * pre[j].y, run.y, dxg[j].y);
                        dxg[j].z = fmaf(fd[j].z * pre[j].z, run.z, dxg[j].z);
                        dxg[j].w = fmaf(fd[j].w * pre[j].w, run.w, dxg[j].w);
                        run.x *= f[j].x; run.y *= f[j].y; run.z *= f[j].z; run.w *= f[j].w;
                    }
                }
            } else if (a.dgate != nullptr) {
                for (int rr = 0; rr < rc; ++rr) {
                    const float4 d = dgcol[rr * NT];
                    if (qi[0] >= 0) a.dgate[qi[0] * R + r0 + rr] = d.x;
                    if (qi[1] >= 0) a.dgate[qi[1] * R + r0 + rr] = d.y;
                    if (qi[2] >= 0) a.dgate[qi[2] * R + r0 + rr] = d.z;
                    if (qi[3] >= 0) a.dgate[qi[3] * R + r0 + rr] = d.w;
                }
            }
        }

        // ---- dx_j = x_j * sum(w) + sum(w * -c_j)  (+ the gate path) ----
        if (a.dx != nullptr) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const long long qa = qi[2 * p], qb = qi[2 * p + 1];
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    float2 r = fma2(x2[p][j], sw[p], dxn[p][j]);
                    if (j < BwdXShape::MAX_FUSED_S) {
                        r.x += p == 0 ? dxg[j].x : dxg[j].z;
                        r.y += p == 0 ? dxg[j].y : dxg[j].w;
                    }
                    if (j < a.Dsrc) {
                        if (qa >= 0) a.dx[qa * a.Dsrc + j] = r.x;
                        if (qb >= 0) a.dx[qb * a.Dsrc + j] = r.y;
                    }
                }
            }
        }
    }
}

template <int D, int OT, int SPARSE = 0>
inline int launch_bwd_x_ot(const BwdXArgs& a, int basis, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e;
    if (basis == IRBFN_BASIS_GAUSSIAN) {
        auto kern = irbfn_bwd_x_kernel<D, OT, IRBFN_BASIS_GAUSSIAN, SPARSE>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, BwdXShape::NT, smem, st>>>(a);
    } else {
        auto kern = irbfn_bwd_x_kernel<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, SPARSE>;
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

// gate-sparse instantiations (irbfn_bwd_xs_inst.cu)
template <int D>
inline int launch_bwd_x_sparse(const BwdXArgs& a, int OT, int basis, int grid, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return launch_bwd_x_ot<D, 0, 1>(a, basis, grid, smem, st);
        case 2: return launch_bwd_x_ot<D, 2, 1>(a, basis, grid, smem, st);
        case 4: return launch_bwd_x_ot<D, 4, 1>(a, basis, grid, smem, st);
        case 8: return launch_bwd_x_ot<D, 8, 1>(a, basis, grid, smem, st);
        case 10: return launch_bwd_x_ot<D, 10, 1>(a, basis, grid, smem, st);
        case 16: return launch_bwd_x_ot<D, 16, 1>(a, basis, grid, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
