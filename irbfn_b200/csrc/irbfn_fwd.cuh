// IRBFN forward kernel (sm_100a): gate -> RBF activations -> blend -> readout, fused.
//
// Reference arithmetic (paths relative to the reference checkout):
//   gate    src/irbfn_mpc/model.py:70-95      blend   src/irbfn_mpc/model.py:187-193
//   layer   flax_rbf/flax_rbf.py:275-283      readout src/irbfn_mpc/model.py:196
//
// Design:
//   * query-stationary: each thread owns Q = 4 queries, held in registers as two packed pairs
//     (x_{q0,j}, x_{q1,j}); every fp32 operation of the distance loop is a packed FADD2/FFMA2
//     whose centre operand is a scalar register broadcast to both lanes (SASS `Rn.F32`), so a
//     (query, region, kernel) evaluation costs D+2 issue slots for 2D+2 FP32-pipe cycles and the
//     kernel is bound by the FP32 pipe, not by instruction issue.
//   * parameters are staged in shared memory once per block (or once per (region,kernel) chunk
//     when they do not fit) as tiles of KT = 4 kernels, TRANSPOSED: row j of a tile holds -c_j of
//     its four kernels, row D their width scales s = exp(-2 log_sig) (times -log2 e for the
//     gaussian).  One broadcast LDS.128 per feature feeds 8 independent packed chains
//     (4 kernels x 2 query pairs): feature-outer, chain-inner, so latency hides inside one warp.
//   * the gate lives in a thread-private column of a shared-memory table gtab[region][thread],
//     computed once per query tile from the packed bound tables (no (B,R) array in HBM).
//   * loop order: kernel tile outer, region inner, so the blended activation h_k is complete in
//     registers before the readout touches it; (B,R,K) and (B,K) never exist.  The basis function
//     (MUFU) of one (tile, region) step is evaluated DURING the next step, two MUFUs after each
//     feature of its distance loop, and blended at that step's end (software pipeline): the XU
//     never sees a burst and its latency is covered by packed arithmetic of the same warp.
//   * persistent grid: one NT-thread CTA per SM walks query tiles of NT*4 queries.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct FwdShape {
    static constexpr int Q = 4;           // queries per thread
    static constexpr int KT = 4;          // kernels per register tile / per shared-memory tile
    static __host__ __device__ constexpr int tile_floats(int D) { return (D + 1) * 4; }
};

template <int D, int OT, int BASIS, int NT>
__global__ void __launch_bounds__(NT, 1) irbfn_fwd_kernel(const FwdArgs a) {
    constexpr int Q = FwdShape::Q, TQ = NT * Q, KT = FwdShape::KT;
    constexpr int OP = OT / 2;            // packed output pairs
    constexpr int OTS = OT > 0 ? OT : 1;
    constexpr int OPS = OP > 0 ? OP : 1;

    extern __shared__ __align__(16) float smem[];
    const int KCT = a.KC / KT;                                  // kernel tiles per region in a stage
    float4* stage4 = reinterpret_cast<float4*>(smem);           // [RC][KCT][D+1] float4
    float* Ws = smem + (size_t)a.RC * KCT * (D + 1) * 4;        // [KC][OT]
    float* gtab = Ws + (size_t)a.KC * (OT > 0 ? OT : 0);        // [RC][NT] float4
    float* stage = smem;

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, RC = a.RC, KC = a.KC;
    const int n_rch = (R + RC - 1) / RC;
    const int kb = blockIdx.y * a.k_per_split;              // this CTA's kernel range (split-K for small batches)
    const int ke = min(K, kb + a.k_per_split);
    const int n_kch = (ke - kb + KC - 1) / KC;
    float* const ybase = a.y != nullptr ? a.y + (size_t)blockIdx.y * a.split_stride : nullptr;
    const bool add_bias = blockIdx.y == 0;
    bool staged = false;

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const long long q0 = (long long)tile * TQ + (long long)tid * Q;

        // ---- queries -> registers, packed in pairs over queries ----
        float2 x2[2][D];
        load_queries4<D>(a.x, q0, a.B, a.Dsrc, x2);

        // ---- output accumulators start at the bias ----
        float2 y2[Q][OPS];
        if (OT > 0) {
#pragma unroll
            for (int op = 0; op < OP; ++op) {
                float2 bv;
                bv.x = (add_bias && 2 * op < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op) : 0.f;
                bv.y = (add_bias && 2 * op + 1 < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op + 1) : 0.f;
#pragma unroll
                for (int q = 0; q < Q; ++q) y2[q][op] = bv;
            }
        }

        for (int rch = 0; rch < n_rch; ++rch) {
            const int r0 = rch * RC;
            const int rc = min(RC, R - r0);
            float4* gcol = reinterpret_cast<float4*>(gtab) + tid;  // + rr * NT

            // ---- gate table for this region chunk: thread-private columns, no barrier ----
            if (a.gate_ext != nullptr) {
                for (int rr = 0; rr < rc; ++rr) {
                    float4 g;
                    g.x = (q0 + 0 < a.B) ? __ldg(a.gate_ext + (q0 + 0) * R + r0 + rr) : 0.f;
                    g.y = (q0 + 1 < a.B) ? __ldg(a.gate_ext + (q0 + 1) * R + r0 + rr) : 0.f;
                    g.z = (q0 + 2 < a.B) ? __ldg(a.gate_ext + (q0 + 2) * R + r0 + rr) : 0.f;
                    g.w = (q0 + 3 < a.B) ? __ldg(a.gate_ext + (q0 + 3) * R + r0 + rr) : 0.f;
                    gcol[rr * NT] = g;
                }
            } else {
                gate_columns(a.gd, a.x, q0, a.B, a.Dsrc, r0, rc, gcol, NT);
            }

            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kb + kch * KC;
                const int kc = min(KC, ke - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;

                // ---- stage parameters of (region chunk, kernel chunk), transposed per 4-kernel tile ----
                if (!(a.stage_once && staged)) {
                    __syncthreads();  // everyone is done with the previous stage
                    for (int idx = tid; idx < rc * kcp; idx += NT) {
                        const int rr = idx / kcp, kk = idx - rr * kcp;
                        float* t = stage + ((size_t)rr * KCT + (kk >> 2)) * ((D + 1) * 4) + (kk & 3);
                        if (kk < kc) {
                            const size_t rk = (size_t)(r0 + rr) * K + (k0 + kk);
#pragma unroll
                            for (int j = 0; j < D; ++j)
                                t[4 * j] = j < a.Dsrc ? -__ldg(a.centers + rk * a.Dsrc + j) : 0.f;
                            t[4 * D] = width_scale<BASIS>(__ldg(a.log_sigs + rk));
                        } else {
#pragma unroll
                            for (int j = 0; j <= D; ++j) t[4 * j] = 0.f;
                        }
                    }
                    if (OT > 0) {
                        for (int idx = tid; idx < kcp * OTS; idx += NT) {
                            const int kk = idx / OTS, o = idx - kk * OTS;
                            Ws[idx] = (kk < kc && o < a.ocur)
                                          ? __ldg(a.W + (size_t)(k0 + kk) * a.ldw + a.o0 + o) : 0.f;
                        }
                    }
                    __syncthreads();
                    staged = true;
                }

                // ---- hot loop: (kernel tile, region) steps, blend pipelined by one step ----
                float2 h2[KT][2], ph[KT][2], ea[KT][2];   // blend accumulators, basis values, basis arguments
                float2 gp[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
                for (int kk = 0; kk < KT; ++kk) {
                    h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);
                    ea[kk][0] = ea[kk][1] = make_float2(1.f, 1.f);   // finite basis value, blended with gate 0
                }

                // y += h . W[kt .. kt+3, :] for the finished kernel tile at chunk offset kt
                auto readout = [&](int kt) {
                    if (OT > 0) {
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
                            const float2* w2 = reinterpret_cast<const float2*>(Ws + (size_t)(kt + kk) * OTS);
                            float2 wv[OPS];
#pragma unroll
                            for (int op = 0; op < OP; ++op) wv[op] = w2[op];
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const float2 hx = dup2(h2[kk][p].x), hy = dup2(h2[kk][p].y);
#pragma unroll
                                for (int op = 0; op < OP; ++op) {
                                    y2[2 * p][op] = fma2(hx, wv[op], y2[2 * p][op]);
                                    y2[2 * p + 1][op] = fma2(hy, wv[op], y2[2 * p + 1][op]);
                                }
                            }
                        }
                    }
                };
                auto store_hidden = [&](int kt) {
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        const int k = k0 + kt + kk;
                        if (k < K) {
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const long long qa = q0 + 2 * p, qb = qa + 1;
                                if (qa < a.B) {
                                    float* hp = a.hidden + qa * K + k;
                                    *hp = (rch > 0 ? *hp : 0.f) + h2[kk][p].x;
                                }
                                if (qb < a.B) {
                                    float* hp = a.hidden + qb * K + k;
                                    *hp = (rch > 0 ? *hp : 0.f) + h2[kk][p].y;
                                }
                            }
                        }
                    }
                };
                // basis function of the PREVIOUS step's arguments: MUFU number i of 16 (4 kernels x 4 queries)
                auto mufu = [&](int i) {
                    const int kk = i >> 2, p = (i >> 1) & 1;
                    float& src = (i & 1) ? ea[kk][p].y : ea[kk][p].x;
                    float& dst = (i & 1) ? ph[kk][p].y : ph[kk][p].x;
                    dst = BASIS == IRBFN_BASIS_GAUSSIAN ? ex2_approx(src) : rcp_approx(src);
                };
                // distances of one step: feature-outer, 8 independent packed chains; the 16 MUFUs of the
                // previous step are spread between the features so that the XU never sees a burst and the
                // issue gaps a MUFU leaves are filled with this step's packed arithmetic.
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
                // blend of the PREVIOUS step
                auto blend = [&]() {
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        h2[kk][0] = fma2(gp[0], ph[kk][0], h2[kk][0]);
                        h2[kk][1] = fma2(gp[1], ph[kk][1], h2[kk][1]);
                    }
                };
                // arguments of this step's basis function; evaluated during the next step's distances
                auto scale = [&](const float4* rec, const float2 (&acc)[KT][2], const float4 g4) {
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
                };

                const size_t rstride = (size_t)KCT * (D + 1);
                for (int kt = 0; kt < kcp; kt += KT) {
                    const float4* rec = stage4 + (size_t)(kt >> 2) * (D + 1);
                    {   // first region of the tile (peeled): also closes the previous kernel tile.
                        // At kt == 0 the pipeline is empty (h and gp are zero): blend and readout add 0.
                        const float4 g4 = gcol[0];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        blend();
                        readout(kt > 0 ? kt - KT : 0);
                        scale(rec, acc, g4);
                        if (a.hidden != nullptr && kt > 0) store_hidden(kt - KT);
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);
                    }
                    for (int rr = 1; rr < rc; ++rr) {   // branch-free steady state
                        rec += rstride;
                        const float4 g4 = gcol[rr * NT];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        blend();
                        scale(rec, acc, g4);
                    }
                }
                // drain the pipeline: basis + blend of the last step, close the last kernel tile
#pragma unroll
                for (int i = 0; i < 16; ++i) mufu(i);
                blend();
                readout(kcp - KT);
                if (a.hidden != nullptr) store_hidden(kcp - KT);
            }
        }

        // ---- store y ----
        if (OT > 0) {
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const long long qq = q0 + q;
                if (qq < a.B) {
                    float* yp = ybase + qq * a.ldw + a.o0;
#pragma unroll
                    for (int op = 0; op < OP; ++op) {
                        if (2 * op < a.ocur) yp[2 * op] = y2[q][op].x;
                        if (2 * op + 1 < a.ocur) yp[2 * op + 1] = y2[q][op].y;
                    }
                }
            }
        }
    }
}

// Host-side: pick the instantiation for (OT, basis, threads) and launch.  Returns a cudaError_t as int.
template <int D, int OT, int BASIS, int NT>
inline int launch_fwd_inst(const FwdArgs& a, int grid, size_t smem, cudaStream_t st) {
    auto kern = irbfn_fwd_kernel<D, OT, BASIS, NT>;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<dim3(grid, a.nsplit > 0 ? a.nsplit : 1), NT, smem, st>>>(a);
    return (int)cudaGetLastError();
}

template <int D, int OT>
inline int launch_fwd_ot(const FwdArgs& a, int basis, int nt, int grid, size_t smem, cudaStream_t st) {
    if (nt != 256) return (int)cudaErrorInvalidValue;   // the one compiled CTA size (fwd_threads_for)
    if (basis == IRBFN_BASIS_GAUSSIAN) return launch_fwd_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 256>(a, grid, smem, st);
    return launch_fwd_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 256>(a, grid, smem, st);
}

template <int D>
inline int launch_fwd(const FwdArgs& a, int OT, int basis, int nt, int grid, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return launch_fwd_ot<D, 0>(a, basis, nt, grid, smem, st);
        case 2: return launch_fwd_ot<D, 2>(a, basis, nt, grid, smem, st);
        case 4: return launch_fwd_ot<D, 4>(a, basis, nt, grid, smem, st);
        case 8: return launch_fwd_ot<D, 8>(a, basis, nt, grid, smem, st);
        case 10: return launch_fwd_ot<D, 10>(a, basis, nt, grid, smem, st);
        case 16: return launch_fwd_ot<D, 16>(a, basis, nt, grid, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
