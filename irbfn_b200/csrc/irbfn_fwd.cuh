// IRBFN forward kernel (sm_100a): gate -> RBF activations -> blend -> readout, fused.
//
// Reference arithmetic (paths relative to the reference checkout):
//   gate    src/irbfn_mpc/model.py:70-95      blend   src/irbfn_mpc/model.py:187-193
//   layer   flax_rbf/flax_rbf.py:275-283      readout src/irbfn_mpc/model.py:196
//
// Design ("forward kernel"):
//   * query-stationary: each thread owns Q = 4 queries, held in registers as two packed pairs
//     (x_{q0,j}, x_{q1,j}); every fp32 operation of the distance loop is a packed FADD2/FFMA2,
//     so an (query, region, kernel) evaluation costs D+2 issue slots for 2D+2 FP32-pipe cycles
//     and the kernel is bound by the FP32 pipe, not by instruction issue.
//   * parameters are staged in shared memory once per block (or once per (region,kernel) chunk
//     when they do not fit) as records  [-c_0,-c_0, ..., -c_{D-1},-c_{D-1}, s, s, pad]  so a
//     broadcast LDS.128 delivers two ready-made packed operands; s already contains
//     exp(-2 log_sig) (times -log2 e for the gaussian), so the hot loop has no divide / sqrt.
//   * the gate lives in a thread-private column of a shared-memory table gtab[region][query],
//     computed once per query tile from the packed bound tables (no (B,R) array in HBM).
//   * loop order: kernel tile (KT=4) outer, region inner, so the blended activation h_k is
//     complete in registers before the readout touches it; (B,R,K) and (B,K) never exist.
//   * persistent grid: one 256-thread CTA per SM walks query tiles of 1024 queries.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct FwdShape {
    static constexpr int NT = 256;        // threads per CTA
    static constexpr int Q = 4;           // queries per thread
    static constexpr int TQ = NT * Q;     // queries per CTA tile
    static constexpr int KT = 4;          // kernels per register tile
    static __host__ __device__ constexpr int rec_floats(int D) { return ((2 * D + 2 + 3) / 4) * 4; }
};

template <int D, int OT, int BASIS>
__global__ void __launch_bounds__(FwdShape::NT, 1) irbfn_fwd_kernel(const FwdArgs a) {
    constexpr int NT = FwdShape::NT, Q = FwdShape::Q, TQ = FwdShape::TQ, KT = FwdShape::KT;
    constexpr int RS = FwdShape::rec_floats(D);
    constexpr int OP = OT / 2;            // packed output pairs
    constexpr int OTS = OT > 0 ? OT : 1;

    extern __shared__ __align__(16) float smem[];
    float* stage = smem;                                  // [RC][KC][RS]
    float* Ws = stage + (size_t)a.RC * a.KC * RS;         // [KC][OT]
    float* gtab = Ws + (size_t)a.KC * (OT > 0 ? OT : 0);  // [RC][TQ]

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, RC = a.RC, KC = a.KC;
    const int n_rch = (R + RC - 1) / RC;
    const int n_kch = (K + KC - 1) / KC;
    bool staged = false;

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const long long q0 = (long long)tile * TQ + (long long)tid * Q;

        // ---- queries -> registers, packed in pairs over queries ----
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

        // ---- output accumulators start at the bias ----
        float2 y2[Q][OP > 0 ? OP : 1];
        if (OT > 0) {
#pragma unroll
            for (int op = 0; op < OP; ++op) {
                float2 bv;
                bv.x = (2 * op < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op) : 0.f;
                bv.y = (2 * op + 1 < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op + 1) : 0.f;
#pragma unroll
                for (int q = 0; q < Q; ++q) y2[q][op] = bv;
            }
        }

        for (int rch = 0; rch < n_rch; ++rch) {
            const int r0 = rch * RC;
            const int rc = min(RC, R - r0);
            float4* gcol = reinterpret_cast<float4*>(gtab) + tid;  // + rr * (TQ/4)

            // ---- gate table for this region chunk: thread-private columns, no barrier ----
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

            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kch * KC;
                const int kc = min(KC, K - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;

                // ---- stage parameters of (region chunk, kernel chunk) ----
                if (!(a.stage_once && staged)) {
                    __syncthreads();  // everyone is done with the previous stage
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
                            const float s = width_scale<BASIS>(__ldg(a.log_sigs + rk));
                            rec[2 * D] = s;
                            rec[2 * D + 1] = s;
                        } else {
#pragma unroll
                            for (int j = 0; j < 2 * D + 2; ++j) rec[j] = 0.f;
                        }
#pragma unroll
                        for (int j = 2 * D + 2; j < RS; ++j) rec[j] = 0.f;
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

                // ---- hot loop ----
                for (int kt = 0; kt < kcp; kt += KT) {
                    float2 h2[KT][2];
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);

                    for (int rr = 0; rr < rc; ++rr) {
                        const float4 g4 = gcol[rr * (TQ / 4)];
                        const float2 g2[2] = {make_float2(g4.x, g4.y), make_float2(g4.z, g4.w)};
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
                                h2[kk][p] = fma2(g2[p], phi, h2[kk][p]);
                            }
                        }
                    }

                    // ---- readout of the finished kernel tile: y += h_k * W[k,:] ----
                    if (OT > 0) {
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
                            const float2* w2 = reinterpret_cast<const float2*>(Ws + (size_t)(kt + kk) * OTS);
                            float2 wv[OP > 0 ? OP : 1];
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
                    if (a.hidden != nullptr) {
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
                    }
                }
            }
        }

        // ---- store y ----
        if (OT > 0) {
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const long long qq = q0 + q;
                if (qq < a.B) {
                    float* yp = a.y + qq * a.ldw + a.o0;
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

// Host-side: pick the instantiation for (OT, basis) and launch.  Returns a cudaError_t as int.
template <int D, int OT>
inline int launch_fwd_ot(const FwdArgs& a, int basis, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e;
    if (basis == IRBFN_BASIS_GAUSSIAN) {
        auto kern = irbfn_fwd_kernel<D, OT, IRBFN_BASIS_GAUSSIAN>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, FwdShape::NT, smem, st>>>(a);
    } else {
        auto kern = irbfn_fwd_kernel<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, FwdShape::NT, smem, st>>>(a);
    }
    return (int)cudaGetLastError();
}

template <int D>
inline int launch_fwd(const FwdArgs& a, int OT, int basis, int grid, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return launch_fwd_ot<D, 0>(a, basis, grid, smem, st);
        case 2: return launch_fwd_ot<D, 2>(a, basis, grid, smem, st);
        case 4: return launch_fwd_ot<D, 4>(a, basis, grid, smem, st);
        case 8: return launch_fwd_ot<D, 8>(a, basis, grid, smem, st);
        case 10: return launch_fwd_ot<D, 10>(a, basis, grid, smem, st);
        case 16: return launch_fwd_ot<D, 16>(a, basis, grid, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
