// FP32 / XU issue-rate micro-benchmark for sm_100a.
//
// The IRBFN kernels are bound by the SM's FP32 pipe, for which
// MEASURED_PEAKS.json has no entry.  This file measures the denominators the
// roofline in bench.py uses: sustained lane-operations per second of
//   mode 0: FFMA            (scalar fp32 fused multiply-add)
//   mode 1: FFMA2           (packed fma.rn.f32x2, two fp32 lanes per thread-instruction)
//   mode 2: FADD2 + FFMA2   (the distance inner loop's instruction pair)
//   mode 3: MUFU.EX2
//   mode 4: packed IRBFN-like mix: D x (FADD2, FFMA2), FMUL2, 2 x MUFU.EX2, FFMA2   (D = 8)
//   mode 5: the same mix with scalar instructions
//   mode 6: FFMA2 with MUFU.EX2 interleaved 1 per 9 (does the XU pipe co-issue for free?)
//   mode 7: DFMA            (fp64 fused multiply-add: the FP64 pipe alone)
//   mode 8: FFMA2 + DFMA interleaved 1:1 (do the FP32 and FP64 pipes run concurrently?)
//   mode 9: FFMA2 + DFMA interleaved 2:1
//   mode 10: IRBFN-like mix on both pipes: 8 packed fp32 (pair,kernel) evaluations + 8 fp64
//            (query,kernel) evaluations whose d^2 is narrowed to fp32 with integer (ALU) ops
//   mode 11: the same with cvt.rn.f32.f64 doing the narrowing
//   modes 12-17: the dual-pipe forward mix: NP packed fp32 query pairs (direct-form distance) plus
//            ND fp64 queries (expanded-form distance: D DFMA + 1 DADD) against the same kernel.
//            12: NP=1 ND=2 cvt   13: NP=1 ND=2 integer narrowing   14: NP=2 ND=2 cvt
//            15: NP=1 ND=3 cvt   16: NP=2 ND=4 cvt                 17: NP=2 ND=4 integer narrowing
//            22-27: feature-outer variant (mix2): 22: NP=2 ND=0  23: NP=0 ND=4  24: NP=1 ND=2
//            25: NP=2 ND=2   26: NP=1 ND=3   27: NP=2 ND=4
//            20: NP=2 ND=0 (the fp32-only kernel's mix, records from shared memory)   21: NP=0 ND=4 cvt
//   mode 18: DMMA m8n8k4 (fp64 tensor path) alone      mode 19: DMMA + FFMA2 interleaved 1:4
// Every mode runs ILP independent dependency chains per thread so that latency is hidden
// and only the pipe / issue rate shows.
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

__device__ __forceinline__ float ex2_approx(float v) {
    float r;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

template <int MODE>
__global__ void __launch_bounds__(256) peak_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int ILP = 8;
    float2 v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) v[i] = make_float2(a + i + threadIdx.x * 1e-3f, b - i);
    const float2 a2 = make_float2(a, a * 1.0001f);
    const float2 b2 = make_float2(b, b * 0.9999f);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) {
                    v[i].x = fmaf(v[i].x, a2.x, b2.x);
                    v[i].y = fmaf(v[i].y, a2.y, b2.y);
                }
        } else if (MODE == 1) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) v[i] = __ffma2_rn(v[i], a2, b2);
        } else if (MODE == 2) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) {
                    float2 t = __fadd2_rn(v[i], b2);
                    v[i] = __ffma2_rn(t, t, v[i]);
                }
        } else if (MODE == 3) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) {
                    v[i].x = ex2_approx(v[i].x);
                    v[i].y = ex2_approx(v[i].y);
                }
        } else if (MODE == 4) {
            // ILP/2 query pairs x 2 kernels
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                float2 acc = make_float2(0.f, 0.f);
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    float2 c = make_float2(b2.x + d, b2.x + d);
                    float2 t = __fadd2_rn(v[i], c);
                    acc = __ffma2_rn(t, t, acc);
                }
                acc = __fmul2_rn(acc, a2);
                float2 p = make_float2(ex2_approx(acc.x), ex2_approx(acc.y));
                v[i] = __ffma2_rn(p, b2, v[i]);
            }
        } else if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                float ax = 0.f, ay = 0.f;
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    float c = b2.x + d;
                    float tx = v[i].x - c, ty = v[i].y - c;
                    ax = fmaf(tx, tx, ax);
                    ay = fmaf(ty, ty, ay);
                }
                ax *= a2.x; ay *= a2.y;
                v[i].x = fmaf(ex2_approx(ax), b2.x, v[i].x);
                v[i].y = fmaf(ex2_approx(ay), b2.y, v[i].y);
            }
        } else if (MODE == 6) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                float2 w = v[i];
#pragma unroll
                for (int d = 0; d < 9; ++d) w = __ffma2_rn(w, a2, b2);
                v[i] = make_float2(ex2_approx(w.x), ex2_approx(w.y));
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// fp64 -> fp32 narrowing of a positive value with integer instructions only (truncating):
// float bits = ((hi - (896 << 20)) : lo) << 3, with the exponent clamped into fp32's normal range.
__device__ __forceinline__ float narrow_bits(double v) {
    int hi = __double2hiint(v);
    const unsigned lo = (unsigned)__double2loint(v);
    hi = max(min(hi, 0x47d00000), 0x38100000) - 0x38000000;
    return __uint_as_float(__funnelshift_l(lo, (unsigned)hi, 3));
}

template <int MODE>
__global__ void __launch_bounds__(256) peak64_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int ILP = 8;
    float2 v[ILP];
    double w[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        v[i] = make_float2(a + i + threadIdx.x * 1e-3f, b - i);
        w[i] = (double)(a + i) + threadIdx.x * 1e-3;
    }
    const float2 a2 = make_float2(a, a * 1.0001f);
    const float2 b2 = make_float2(b, b * 0.9999f);
    const double ad = (double)a, bd = (double)b;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 7) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) w[i] = fma(w[i], ad, bd);
        } else if (MODE == 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) {
                    v[i] = __ffma2_rn(v[i], a2, b2);
                    w[i] = fma(w[i], ad, bd);
                }
        } else if (MODE == 9) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < ILP; ++i) {
                    v[i] = __ffma2_rn(v[i], a2, b2);
                    w[i] = fma(w[i], ad, bd);
                    v[i] = __ffma2_rn(v[i], b2, a2);
                }
        } else {
            // 8 packed fp32 (pair,kernel) evaluations ...
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                float2 acc = make_float2(0.f, 0.f);
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    float2 c = make_float2(b2.x + d, b2.x + d);
                    float2 t = __fadd2_rn(v[i], c);
                    acc = __ffma2_rn(t, t, acc);
                }
                acc = __fmul2_rn(acc, a2);
                float2 p = make_float2(ex2_approx(acc.x), ex2_approx(acc.y));
                v[i] = __ffma2_rn(p, b2, v[i]);
            }
            // ... and 8 fp64 (query,kernel) evaluations
            float hs = 0.f;
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                double acc = 0x1p-100;
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    const double t = w[i] + (bd + d);
                    acc = fma(t, t, acc);
                }
                acc *= ad;
                const float e = (MODE == 10) ? narrow_bits(acc) : (float)acc;
                hs = fmaf(ex2_approx(-e), b2.x, hs);
            }
            w[0] += (double)hs * 1e-30;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += v[i].x + v[i].y + (float)w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Dual-pipe forward mix (see DESIGN.md "forward kernel"): per kernel, NP packed query pairs take the
// direct form on the FP32 pipe and ND queries the expanded form |x|^2 + |c|^2 - 2 x.c on the FP64 pipe.
template <int NP, int ND, int NARROW_INT>
__global__ void __launch_bounds__(256) mix_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int D = 8, KT = 4, NK = 64, RS = 12, RD = 10;
    __shared__ __align__(16) float rec32[NK * RS];    // [-c_0..-c_7, s, pad x3]
    __shared__ __align__(16) double rec64[NK * RD];   // [|c|^2, -2c_0..-2c_7, pad]
    for (int i = threadIdx.x; i < NK * RS; i += blockDim.x) rec32[i] = b + (i % 7) * 0.25f;
    for (int i = threadIdx.x; i < NK * RD; i += blockDim.x) rec64[i] = (double)b + (i % 5) * 0.25;
    __syncthreads();
    float2 x2[NP > 0 ? NP : 1][D];
    double xd[ND > 0 ? ND : 1][D], xx[ND > 0 ? ND : 1];
    float2 h2[KT][NP > 0 ? NP : 1];
    float hd[KT][ND > 0 ? ND : 1];
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int j = 0; j < D; ++j) x2[p][j] = make_float2(a + j + p + threadIdx.x * 1e-3f, b - j);
#pragma unroll
    for (int q = 0; q < ND; ++q) {
        xx[q] = 0.0;
#pragma unroll
        for (int j = 0; j < D; ++j) { xd[q][j] = (double)(a + j * 0.25f + q) + threadIdx.x * 1e-3; xx[q] += xd[q][j] * xd[q][j]; }
    }
#pragma unroll
    for (int k = 0; k < KT; ++k) {
#pragma unroll
        for (int p = 0; p < NP; ++p) h2[k][p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < ND; ++q) hd[k][q] = 0.f;
    }
    const float2 g2 = make_float2(b, b * 0.999f);
    for (int it = 0; it < iters; ++it) {
        const int k0 = (it * KT) & (NK - 1);
#pragma unroll
        for (int k = 0; k < KT; ++k) {
            const float4* r4 = reinterpret_cast<const float4*>(rec32 + (k0 + k) * RS);
            const float4 ca = r4[0], cb = r4[1], cc = r4[2];
            const float c[9] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w, cc.x};
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                float2 t = __fadd2_rn(x2[p][0], make_float2(c[0], c[0]));
                float2 acc = __fmul2_rn(t, t);
#pragma unroll
                for (int j = 1; j < D; ++j) {
                    t = __fadd2_rn(x2[p][j], make_float2(c[j], c[j]));
                    acc = __ffma2_rn(t, t, acc);
                }
                acc = __fmul2_rn(acc, make_float2(c[8], c[8]));
                h2[k][p] = __ffma2_rn(g2, make_float2(ex2_approx(acc.x), ex2_approx(acc.y)), h2[k][p]);
            }
            if (ND > 0) {
                const double2* d2p = reinterpret_cast<const double2*>(rec64 + (k0 + k) * RD);
                double m[10];
#pragma unroll
                for (int i = 0; i < 5; ++i) { const double2 v = d2p[i]; m[2 * i] = v.x; m[2 * i + 1] = v.y; }
                float e[ND > 0 ? ND : 1];
#pragma unroll
                for (int q = 0; q < ND; ++q) {
                    double acc = xx[q] + m[0];
#pragma unroll
                    for (int j = 0; j < D; ++j) acc = fma(xd[q][j], m[1 + j], acc);
                    e[q] = NARROW_INT ? narrow_bits(acc) : (float)acc;
                }
#pragma unroll
                for (int q = 0; q < ND; ++q) hd[k][q] = fmaf(g2.x, ex2_approx(e[q] * c[8]), hd[k][q]);
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < KT; ++k) {
#pragma unroll
        for (int p = 0; p < NP; ++p) s += h2[k][p].x + h2[k][p].y;
#pragma unroll
        for (int q = 0; q < ND; ++q) s += hd[k][q];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// The same mix, written feature-outer / chain-inner so that every step exposes 4 NP + 4 ND
// independent chains, with the records transposed in shared memory ([feature][kernel] within a
// 4-kernel tile) so one LDS.128 serves one step of 4 fp32 (or 2 fp64) kernels.
template <int NP, int ND, int NARROW_INT>
__global__ void __launch_bounds__(256) mix2_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int D = 8, KT = 4, NT = 16;
    __shared__ __align__(16) float rec32[NT][D + 1][KT];    // rows 0..D-1: -c_j ; row D: scale
    __shared__ __align__(16) double rec64[NT][D + 1][KT];   // row 0: |c|^2 ; rows 1..D: -2 c_j
    for (int i = threadIdx.x; i < NT * (D + 1) * KT; i += blockDim.x) {
        (&rec32[0][0][0])[i] = b + (i % 7) * 0.25f;
        (&rec64[0][0][0])[i] = (double)b + (i % 5) * 0.25;
    }
    __syncthreads();
    float2 x2[NP > 0 ? NP : 1][D];
    double xd[ND > 0 ? ND : 1][D], xx[ND > 0 ? ND : 1];
    float2 h2[KT][NP > 0 ? NP : 1];
    float hd[KT][ND > 0 ? ND : 1];
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int j = 0; j < D; ++j) x2[p][j] = make_float2(a + j + p + threadIdx.x * 1e-3f, b - j);
#pragma unroll
    for (int q = 0; q < ND; ++q) {
        xx[q] = 0.0;
#pragma unroll
        for (int j = 0; j < D; ++j) { xd[q][j] = (double)(a + j * 0.25f + q) + threadIdx.x * 1e-3; xx[q] += xd[q][j] * xd[q][j]; }
    }
#pragma unroll
    for (int k = 0; k < KT; ++k) {
#pragma unroll
        for (int p = 0; p < NP; ++p) h2[k][p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < ND; ++q) hd[k][q] = 0.f;
    }
    const float2 g2 = make_float2(b, b * 0.999f);
    for (int it = 0; it < iters; ++it) {
        const int t = it & (NT - 1);
        float2 acc[KT][NP > 0 ? NP : 1];
        double dac[KT][ND > 0 ? ND : 1];
        if (ND > 0) {
            const double2 c0 = *reinterpret_cast<const double2*>(&rec64[t][0][0]);
            const double2 c1 = *reinterpret_cast<const double2*>(&rec64[t][0][2]);
            const double cc[4] = {c0.x, c0.y, c1.x, c1.y};
#pragma unroll
            for (int k = 0; k < KT; ++k)
#pragma unroll
                for (int q = 0; q < ND; ++q) dac[k][q] = xx[q] + cc[k];
        }
#pragma unroll
        for (int j = 0; j < D; ++j) {
            if (NP > 0) {
                const float4 c4 = *reinterpret_cast<const float4*>(&rec32[t][j][0]);
                const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int p = 0; p < NP; ++p) {
                        const float2 d = __fadd2_rn(x2[p][j], make_float2(c[k], c[k]));
                        acc[k][p] = j == 0 ? __fmul2_rn(d, d) : __ffma2_rn(d, d, acc[k][p]);
                    }
            }
            if (ND > 0) {
                const double2 m0 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][0]);
                const double2 m1 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][2]);
                const double m[4] = {m0.x, m0.y, m1.x, m1.y};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int q = 0; q < ND; ++q) dac[k][q] = fma(xd[q][j], m[k], dac[k][q]);
            }
        }
        const float4 s4 = *reinterpret_cast<const float4*>(&rec32[t][D][0]);
        const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
        for (int k = 0; k < KT; ++k) {
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const float2 e = __fmul2_rn(acc[k][p], make_float2(sc[k], sc[k]));
                h2[k][p] = __ffma2_rn(g2, make_float2(ex2_approx(e.x), ex2_approx(e.y)), h2[k][p]);
            }
#pragma unroll
            for (int q = 0; q < ND; ++q)
                hd[k][q] = fmaf(g2.x, ex2_approx((NARROW_INT ? narrow_bits(dac[k][q]) : (float)dac[k][q]) * sc[k]), hd[k][q]);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < KT; ++k) {
#pragma unroll
        for (int p = 0; p < NP; ++p) s += h2[k][p].x + h2[k][p].y;
#pragma unroll
        for (int q = 0; q < ND; ++q) s += hd[k][q];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Operand-freshness probes: PAT 0: w = fma(u, a, w) (2 fresh 64-bit operands, 1 reused)
//   1: w = fma(u, z, w) (3 fresh)   2: w = fma(u, u, w) (direct-form pattern)
//   3: FFMA2 v = fma2(p, q, v) (3 fresh pairs)   4: FADD2 t = p + c.F32 ; FFMA2 v = fma2(t, t, v)
template <int PAT>
__global__ void __launch_bounds__(256) operand_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int ILP = 8;
    double w[ILP], u[ILP], z[ILP];
    float2 v[ILP], p2[ILP], q2[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        w[i] = i; u[i] = (double)a + i * 1e-3 + threadIdx.x * 1e-6; z[i] = (double)b - i * 1e-3;
        v[i] = make_float2(i, -i); p2[i] = make_float2(a + i * 1e-3f, a - i * 1e-3f); q2[i] = make_float2(b + i * 1e-3f, b);
    }
    const double ad = (double)a;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                if (PAT == 0) w[i] = fma(u[(i + r) % ILP], ad, w[i]);
                if (PAT == 1) w[i] = fma(u[(i + r) % ILP], z[(i + 2 * r) % ILP], w[i]);
                if (PAT == 2) w[i] = fma(u[(i + r) % ILP], u[(i + r) % ILP], w[i]);
                if (PAT == 3) v[i] = __ffma2_rn(p2[(i + r) % ILP], q2[(i + 2 * r) % ILP], v[i]);
                if (PAT == 4) {
                    const float2 t = __fadd2_rn(p2[(i + r) % ILP], make_float2(q2[r].x, q2[r].x));
                    v[i] = __ffma2_rn(t, t, v[i]);
                }
            }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += (float)w[i] + v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// fp64-only dissection of the expanded-form path: TAIL = narrowing + scale + MUFU + blend present,
// LDSOP = centre operands come from shared memory (else from registers).
template <int TAIL, int LDSOP>
__global__ void __launch_bounds__(256) dp_probe_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int D = 8, KT = 4, NT = 16, ND = 4;
    __shared__ __align__(16) double rec64[NT][D + 1][KT];
    for (int i = threadIdx.x; i < NT * (D + 1) * KT; i += blockDim.x) (&rec64[0][0][0])[i] = (double)b + (i % 5) * 0.25;
    __syncthreads();
    double xd[ND][D], xx[ND], mreg[KT];
    float hd[KT][ND];
#pragma unroll
    for (int k = 0; k < KT; ++k) mreg[k] = (double)b + k * 0.125 + threadIdx.x * 1e-9;
#pragma unroll
    for (int q = 0; q < ND; ++q) {
        xx[q] = 0.0;
#pragma unroll
        for (int j = 0; j < D; ++j) { xd[q][j] = (double)(a + j * 0.25f + q) + threadIdx.x * 1e-3; xx[q] += xd[q][j] * xd[q][j]; }
#pragma unroll
        for (int k = 0; k < KT; ++k) hd[k][q] = 0.f;
    }
    for (int it = 0; it < iters; ++it) {
        const int t = it & (NT - 1);
        double dac[KT][ND];
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int q = 0; q < ND; ++q) dac[k][q] = xx[q] + mreg[k];
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double m[4];
            if (LDSOP) {
                const double2 m0 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][0]);
                const double2 m1 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][2]);
                m[0] = m0.x; m[1] = m0.y; m[2] = m1.x; m[3] = m1.y;
            } else {
#pragma unroll
                for (int k = 0; k < KT; ++k) m[k] = mreg[(k + j) & 3];
            }
#pragma unroll
            for (int k = 0; k < KT; ++k)
#pragma unroll
                for (int q = 0; q < ND; ++q) dac[k][q] = fma(xd[q][j], m[k], dac[k][q]);
        }
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int q = 0; q < ND; ++q) {
                if (TAIL) hd[k][q] = fmaf(b, ex2_approx((float)dac[k][q] * a), hd[k][q]);
                else xx[q] += dac[k][q] * 1e-30;
            }
    }
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < ND; ++q) {
        s += (float)xx[q];
#pragma unroll
        for (int k = 0; k < KT; ++k) s += hd[k][q];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Warp-specialised dual-pipe mix: in a 512-thread CTA the first NW32 warps run the packed fp32
// direct-form loop (2 query pairs per thread), the others the fp64 expanded-form loop (4 queries per
// thread) for PCT % as many iterations.  Each warp's instruction stream targets one math pipe only.
template <int NW32, int PCT>
__global__ void __launch_bounds__(512, 1) species_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int D = 8, KT = 4, NT = 16;
    __shared__ __align__(16) float rec32[NT][D + 1][KT];
    __shared__ __align__(16) double rec64[NT][D + 1][KT];
    for (int i = threadIdx.x; i < NT * (D + 1) * KT; i += blockDim.x) {
        (&rec32[0][0][0])[i] = b + (i % 7) * 0.25f;
        (&rec64[0][0][0])[i] = (double)b + (i % 5) * 0.25;
    }
    __syncthreads();
    float s = 0.f;
    if ((threadIdx.x >> 5) < NW32) {
        constexpr int NP = 2;
        float2 x2[NP][D], h2[KT][NP];
#pragma unroll
        for (int p = 0; p < NP; ++p)
#pragma unroll
            for (int j = 0; j < D; ++j) x2[p][j] = make_float2(a + j + p + threadIdx.x * 1e-3f, b - j);
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int p = 0; p < NP; ++p) h2[k][p] = make_float2(0.f, 0.f);
        const float2 g2 = make_float2(b, b * 0.999f);
        for (int it = 0; it < iters; ++it) {
            const int t = it & (NT - 1);
            float2 acc[KT][NP];
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float4 c4 = *reinterpret_cast<const float4*>(&rec32[t][j][0]);
                const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int p = 0; p < NP; ++p) {
                        const float2 d = __fadd2_rn(x2[p][j], make_float2(c[k], c[k]));
                        acc[k][p] = j == 0 ? __fmul2_rn(d, d) : __ffma2_rn(d, d, acc[k][p]);
                    }
            }
            const float4 s4 = *reinterpret_cast<const float4*>(&rec32[t][D][0]);
            const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int k = 0; k < KT; ++k)
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    const float2 e = __fmul2_rn(acc[k][p], make_float2(sc[k], sc[k]));
                    h2[k][p] = __ffma2_rn(g2, make_float2(ex2_approx(e.x), ex2_approx(e.y)), h2[k][p]);
                }
        }
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int p = 0; p < NP; ++p) s += h2[k][p].x + h2[k][p].y;
    } else {
        constexpr int ND = 4;
        double xd[ND][D], xx[ND];
        float hd[KT][ND];
#pragma unroll
        for (int q = 0; q < ND; ++q) {
            xx[q] = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j) { xd[q][j] = (double)(a + j * 0.25f + q) + threadIdx.x * 1e-3; xx[q] += xd[q][j] * xd[q][j]; }
#pragma unroll
            for (int k = 0; k < KT; ++k) hd[k][q] = 0.f;
        }
        const int n64 = (int)((long long)iters * PCT / 100);
        for (int it = 0; it < n64; ++it) {
            const int t = it & (NT - 1);
            double dac[KT][ND];
            {
                const double2 c0 = *reinterpret_cast<const double2*>(&rec64[t][0][0]);
                const double2 c1 = *reinterpret_cast<const double2*>(&rec64[t][0][2]);
                const double cc[4] = {c0.x, c0.y, c1.x, c1.y};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int q = 0; q < ND; ++q) dac[k][q] = xx[q] + cc[k];
            }
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const double2 m0 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][0]);
                const double2 m1 = *reinterpret_cast<const double2*>(&rec64[t][1 + j][2]);
                const double m[4] = {m0.x, m0.y, m1.x, m1.y};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int q = 0; q < ND; ++q) dac[k][q] = fma(xd[q][j], m[k], dac[k][q]);
            }
            const float4 s4 = *reinterpret_cast<const float4*>(&rec32[t][D][0]);
            const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int k = 0; k < KT; ++k)
#pragma unroll
                for (int q = 0; q < ND; ++q) hd[k][q] = fmaf(b, ex2_approx((float)dac[k][q] * sc[k]), hd[k][q]);
        }
#pragma unroll
        for (int q = 0; q < ND; ++q)
#pragma unroll
            for (int k = 0; k < KT; ++k) s += hd[k][q];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Warp-specialised mix with the fp64 side on DMMA (expanded form as an 8x8x8 product per
// (row block of 8 queries, column block of 8 kernels)): NW32 warps of a 512-thread CTA run the packed
// fp32 loop, the others RB row blocks x DMMA for PCT % as many 8-kernel column blocks.
template <int NW32, int PCT, int NARROW_INT>
__global__ void __launch_bounds__(512, 1) species_dmma_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int D = 8, KT = 4, NT = 16, RB = 8;
    __shared__ __align__(16) float rec32[NT][D + 1][KT];
    __shared__ __align__(16) double recB[NT][2][32];     // B fragments: [step][lane] = -2 c[n = lane>>2][4 step + (lane&3)]
    __shared__ __align__(16) double recC[NT][8];         // |c|^2 per column
    __shared__ __align__(16) float recS[NT][8];          // scale per column
    for (int i = threadIdx.x; i < NT * (D + 1) * KT; i += blockDim.x) (&rec32[0][0][0])[i] = b + (i % 7) * 0.25f;
    for (int i = threadIdx.x; i < NT * 64; i += blockDim.x) (&recB[0][0][0])[i] = (double)b + (i % 5) * 0.25;
    for (int i = threadIdx.x; i < NT * 8; i += blockDim.x) { (&recC[0][0])[i] = 1.0 + i; (&recS[0][0])[i] = -0.01f; }
    __syncthreads();
    float s = 0.f;
    const int lane = threadIdx.x & 31;
    if ((threadIdx.x >> 5) < NW32) {
        constexpr int NP = 2;
        float2 x2[NP][D], h2[KT][NP];
#pragma unroll
        for (int p = 0; p < NP; ++p)
#pragma unroll
            for (int j = 0; j < D; ++j) x2[p][j] = make_float2(a + j + p + threadIdx.x * 1e-3f, b - j);
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int p = 0; p < NP; ++p) h2[k][p] = make_float2(0.f, 0.f);
        const float2 g2 = make_float2(b, b * 0.999f);
        for (int it = 0; it < iters; ++it) {
            const int t = it & (NT - 1);
            float2 acc[KT][NP];
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float4 c4 = *reinterpret_cast<const float4*>(&rec32[t][j][0]);
                const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                for (int k = 0; k < KT; ++k)
#pragma unroll
                    for (int p = 0; p < NP; ++p) {
                        const float2 d = __fadd2_rn(x2[p][j], make_float2(c[k], c[k]));
                        acc[k][p] = j == 0 ? __fmul2_rn(d, d) : __ffma2_rn(d, d, acc[k][p]);
                    }
            }
            const float4 s4 = *reinterpret_cast<const float4*>(&rec32[t][D][0]);
            const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int k = 0; k < KT; ++k)
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    const float2 e = __fmul2_rn(acc[k][p], make_float2(sc[k], sc[k]));
                    h2[k][p] = __ffma2_rn(g2, make_float2(ex2_approx(e.x), ex2_approx(e.y)), h2[k][p]);
                }
        }
#pragma unroll
        for (int k = 0; k < KT; ++k)
#pragma unroll
            for (int p = 0; p < NP; ++p) s += h2[k][p].x + h2[k][p].y;
    } else {
        // A fragments of RB row blocks (two k-steps), |x|^2 of this lane's row in each block, gates
        double xa[RB][2], xx[RB];
        float2 hh[RB];
        float gq[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) {
            xa[r][0] = (double)a + r * 0.125 + lane * 1e-3;
            xa[r][1] = (double)b - r * 0.125 + lane * 1e-3;
            xx[r] = xa[r][0] * xa[r][0] + xa[r][1] * xa[r][1];
            hh[r] = make_float2(0.f, 0.f);
            gq[r] = b + r * 1e-3f;
        }
        const int n64 = (int)((long long)iters * PCT / 100);
        for (int it = 0; it < n64; ++it) {
            const int t = it & (NT - 1);
            const double b0 = recB[t][0][lane], b1 = recB[t][1][lane];
            const double2 cc = *reinterpret_cast<const double2*>(&recC[t][2 * (lane & 3)]);
            const float2 sc = *reinterpret_cast<const float2*>(&recS[t][2 * (lane & 3)]);
#pragma unroll
            for (int r = 0; r < RB; ++r) {
                double c0 = xx[r] + cc.x, c1 = xx[r] + cc.y;
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c0), "+d"(c1) : "d"(xa[r][0]), "d"(b0));
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c0), "+d"(c1) : "d"(xa[r][1]), "d"(b1));
                float2 e;
                e.x = NARROW_INT ? narrow_bits(c0) : (float)c0;
                e.y = NARROW_INT ? narrow_bits(c1) : (float)c1;
                e = __fmul2_rn(e, sc);
                hh[r] = __ffma2_rn(make_float2(gq[r], gq[r]), make_float2(ex2_approx(e.x), ex2_approx(e.y)), hh[r]);
            }
        }
#pragma unroll
        for (int r = 0; r < RB; ++r) s += hh[r].x + hh[r].y;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// DMMA m8n8k4 (fp64 mma.sync): 8 independent accumulator pairs per warp; optionally interleaved with FFMA2.
template <int WITH_FFMA2>
__global__ void __launch_bounds__(256) dmma_kernel(float* __restrict__ out, int iters, float a, float b) {
    constexpr int ILP = 8;
    double c0[ILP], c1[ILP];
    float2 v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c0[i] = i; c1[i] = -i; v[i] = make_float2(a + i, b - i); }
    const double fa = (double)a + threadIdx.x * 1e-6, fb = (double)b * 1e-3;
    const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.9999f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(fa), "d"(fb));
            if (WITH_FFMA2) {
#pragma unroll
                for (int u = 0; u < 4; ++u) v[(i + u) % ILP] = __ffma2_rn(v[(i + u) % ILP], a2, b2);
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += (float)(c0[i] + c1[i]) + v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// fp32-pipe lane-operations (FADD/FMUL/FFMA lanes) and MUFU lane-operations per thread per iteration
void mode_ops(int mode, double* fp32_ops, double* mufu_ops, double* fp64_ops) {
    *fp64_ops = 0;
    switch (mode) {
        case 0: *fp32_ops = 8 * 8 * 2; *mufu_ops = 0; break;
        case 1: *fp32_ops = 8 * 8 * 2; *mufu_ops = 0; break;
        case 2: *fp32_ops = 4 * 8 * 4; *mufu_ops = 0; break;
        case 3: *fp32_ops = 0; *mufu_ops = 4 * 8 * 2; break;
        case 4: *fp32_ops = 8 * (8 * 4 + 2 + 2); *mufu_ops = 8 * 2; break;
        case 5: *fp32_ops = 8 * (8 * 4 + 2 + 2); *mufu_ops = 8 * 2; break;
        case 6: *fp32_ops = 8 * 9 * 2; *mufu_ops = 8 * 2; break;
        case 7: *fp32_ops = 0; *mufu_ops = 0; *fp64_ops = 8 * 8; break;
        case 8: *fp32_ops = 8 * 8 * 2; *mufu_ops = 0; *fp64_ops = 8 * 8; break;
        case 9: *fp32_ops = 4 * 8 * 4; *mufu_ops = 0; *fp64_ops = 4 * 8; break;
        case 10:
        case 11: *fp32_ops = 8 * (8 * 4 + 2 + 2) + 8; *mufu_ops = 8 * 2 + 8; *fp64_ops = 8 * 17; break;
        // dual-pipe mixes: per iteration KT=4 kernels; fp32 lanes: NP pairs x 2 x 18 + ND x 2 (scale, blend);
        // MUFU: NP*2 + ND; fp64: ND x 9.  "evaluations" = (2 NP + ND) x 4 = MUFU count.
        case 12: case 13: *fp32_ops = 4 * (1 * 36 + 2 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 2 * 9; break;
        case 14: *fp32_ops = 4 * (2 * 36 + 2 * 2); *mufu_ops = 4 * 6; *fp64_ops = 4 * 2 * 9; break;
        case 15: *fp32_ops = 4 * (1 * 36 + 3 * 2); *mufu_ops = 4 * 5; *fp64_ops = 4 * 3 * 9; break;
        case 16: case 17: *fp32_ops = 4 * (2 * 36 + 4 * 2); *mufu_ops = 4 * 8; *fp64_ops = 4 * 4 * 9; break;
        case 28: case 29: case 30: *fp32_ops = 0; *mufu_ops = 0; *fp64_ops = 64; break;
        case 31: *fp32_ops = 64 * 2; *mufu_ops = 0; break;
        case 32: *fp32_ops = 64 * 4; *mufu_ops = 0; break;
        case 33: *fp32_ops = 4 * (4 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 4 * 9; break;
        case 34: *fp32_ops = 4 * (1 * 36 + 2 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 2 * 9; break;
        case 35: *fp32_ops = 4 * (2 * 36 + 2 * 2); *mufu_ops = 4 * 6; *fp64_ops = 4 * 2 * 9; break;
        case 36: *fp32_ops = 4 * (1 * 36 + 3 * 2); *mufu_ops = 4 * 5; *fp64_ops = 4 * 3 * 9; break;
        case 37: *fp32_ops = 4 * (2 * 36 + 4 * 2); *mufu_ops = 4 * 8; *fp64_ops = 4 * 4 * 9; break;
        case 38: case 39: *fp32_ops = 0; *mufu_ops = 0; *fp64_ops = 4 * 4 * 10; break;
        case 40: case 41: *fp32_ops = 4 * 4 * 2; *mufu_ops = 4 * 4; *fp64_ops = 4 * 4 * 9; break;
        case 42: *fp32_ops = 288 + 32 * 40 / 100.0; *mufu_ops = 16 + 16 * 40 / 100.0; *fp64_ops = 144 * 40 / 100.0; break;
        case 43: *fp32_ops = 288 + 32 * 50 / 100.0; *mufu_ops = 16 + 16 * 50 / 100.0; *fp64_ops = 144 * 50 / 100.0; break;
        case 44: *fp32_ops = 288 + 32 * 60 / 100.0; *mufu_ops = 16 + 16 * 60 / 100.0; *fp64_ops = 144 * 60 / 100.0; break;
        case 45: *fp32_ops = 288 + 32 * 70 / 100.0; *mufu_ops = 16 + 16 * 70 / 100.0; *fp64_ops = 144 * 70 / 100.0; break;
        case 46: *fp32_ops = 288 + 32 * 0 / 100.0; *mufu_ops = 16 + 16 * 0 / 100.0; *fp64_ops = 144 * 0 / 100.0; break;
        case 47: *fp32_ops = 0 + 64 * 100 / 100.0; *mufu_ops = 0 + 32 * 100 / 100.0; *fp64_ops = 288 * 100 / 100.0; break;
        case 48: *fp32_ops = 0 + 64 * 100 / 100.0; *mufu_ops = 0 + 32 * 100 / 100.0; *fp64_ops = 288 * 100 / 100.0; break;
        case 49: *fp32_ops = 288 + 32 * 60 / 100.0; *mufu_ops = 16 + 16 * 60 / 100.0; *fp64_ops = 144 * 60 / 100.0; break;
        case 50: *fp32_ops = 288 + 32 * 80 / 100.0; *mufu_ops = 16 + 16 * 80 / 100.0; *fp64_ops = 144 * 80 / 100.0; break;
        case 51: *fp32_ops = 288 + 32 * 100 / 100.0; *mufu_ops = 16 + 16 * 100 / 100.0; *fp64_ops = 144 * 100 / 100.0; break;
        case 52: *fp32_ops = 288 + 32 * 80 / 100.0; *mufu_ops = 16 + 16 * 80 / 100.0; *fp64_ops = 144 * 80 / 100.0; break;
        case 53: *fp32_ops = 288 + 32 * 100 / 100.0; *mufu_ops = 16 + 16 * 100 / 100.0; *fp64_ops = 144 * 100 / 100.0; break;
        case 54: *fp32_ops = 288 + 32 * 120 / 100.0; *mufu_ops = 16 + 16 * 120 / 100.0; *fp64_ops = 144 * 120 / 100.0; break;
        case 55: *fp32_ops = 216 + 40 * 120 / 100.0; *mufu_ops = 12 + 20 * 120 / 100.0; *fp64_ops = 180 * 120 / 100.0; break;
        case 56: *fp32_ops = 216 + 40 * 160 / 100.0; *mufu_ops = 12 + 20 * 160 / 100.0; *fp64_ops = 180 * 160 / 100.0; break;
        case 20: case 22: *fp32_ops = 4 * (2 * 36); *mufu_ops = 4 * 4; break;
        case 23: *fp32_ops = 4 * (4 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 4 * 9; break;
        case 24: *fp32_ops = 4 * (1 * 36 + 2 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 2 * 9; break;
        case 25: *fp32_ops = 4 * (2 * 36 + 2 * 2); *mufu_ops = 4 * 6; *fp64_ops = 4 * 2 * 9; break;
        case 26: *fp32_ops = 4 * (1 * 36 + 3 * 2); *mufu_ops = 4 * 5; *fp64_ops = 4 * 3 * 9; break;
        case 27: *fp32_ops = 4 * (2 * 36 + 4 * 2); *mufu_ops = 4 * 8; *fp64_ops = 4 * 4 * 9; break;
        case 21: *fp32_ops = 4 * (4 * 2); *mufu_ops = 4 * 4; *fp64_ops = 4 * 4 * 9; break;
        case 18: *fp32_ops = 0; *mufu_ops = 0; *fp64_ops = 8 * 8; break;          // 256 FMA per warp-instr = 8 per lane
        case 19: *fp32_ops = 8 * 4 * 2; *mufu_ops = 0; *fp64_ops = 8 * 8; break;
        default: *fp32_ops = 0; *mufu_ops = 0;
    }
}

}  // namespace

// Runs one mode; rates[0..2] = fp32-pipe, MUFU and fp64-pipe lane-operations per second.
static int run_mode(int mode, int iters, int blocks, float* scratch, double* rates, float* elapsed_ms,
                    cudaStream_t stream) {
    if (mode < 0 || mode > 56 || iters <= 0 || blocks <= 0 || !scratch) return -1;
    cudaEvent_t e0, e1;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return -2;
    auto launch = [&](int n) {
        switch (mode) {
            case 0: peak_kernel<0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 1: peak_kernel<1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 2: peak_kernel<2><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 3: peak_kernel<3><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 4: peak_kernel<4><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 5: peak_kernel<5><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 6: peak_kernel<6><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 7: peak64_kernel<7><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 8: peak64_kernel<8><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 9: peak64_kernel<9><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 10: peak64_kernel<10><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 11: peak64_kernel<11><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 12: mix_kernel<1, 2, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 13: mix_kernel<1, 2, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 14: mix_kernel<2, 2, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 15: mix_kernel<1, 3, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 16: mix_kernel<2, 4, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 17: mix_kernel<2, 4, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 20: mix_kernel<2, 0, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 21: mix_kernel<0, 4, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 22: mix2_kernel<2, 0, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 23: mix2_kernel<0, 4, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 24: mix2_kernel<1, 2, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 25: mix2_kernel<2, 2, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 26: mix2_kernel<1, 3, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 27: mix2_kernel<2, 4, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 28: operand_kernel<0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 29: operand_kernel<1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 30: operand_kernel<2><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 31: operand_kernel<3><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 32: operand_kernel<4><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 33: mix2_kernel<0, 4, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 34: mix2_kernel<1, 2, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 35: mix2_kernel<2, 2, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 36: mix2_kernel<1, 3, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 37: mix2_kernel<2, 4, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 38: dp_probe_kernel<0, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 39: dp_probe_kernel<0, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 40: dp_probe_kernel<1, 0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 41: dp_probe_kernel<1, 1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 42: species_kernel<8, 40><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 43: species_kernel<8, 50><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 44: species_kernel<8, 60><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 45: species_kernel<8, 70><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 46: species_dmma_kernel<8, 0, 0><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 47: species_dmma_kernel<0, 100, 0><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 48: species_dmma_kernel<0, 100, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 49: species_dmma_kernel<8, 60, 0><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 50: species_dmma_kernel<8, 80, 0><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 51: species_dmma_kernel<8, 100, 0><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 52: species_dmma_kernel<8, 80, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 53: species_dmma_kernel<8, 100, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 54: species_dmma_kernel<8, 120, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 55: species_dmma_kernel<6, 120, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 56: species_dmma_kernel<6, 160, 1><<<blocks, 512, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 18: dmma_kernel<0><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
            case 19: dmma_kernel<1><<<blocks, 256, 0, stream>>>(scratch, n, 1.0001f, 0.5f); break;
        }
    };
    launch(iters / 8 + 1);  // warm-up
    cudaEventRecord(e0, stream);
    launch(iters);
    cudaEventRecord(e1, stream);
    cudaError_t err = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (err != cudaSuccess || cudaGetLastError() != cudaSuccess) return -3;
    double f, m, d;
    mode_ops(mode, &f, &m, &d);
    const double scale = (double)iters * (double)blocks * 256.0 / (ms * 1e-3);
    rates[0] = f * scale;
    rates[1] = m * scale;
    rates[2] = d * scale;
    if (elapsed_ms) *elapsed_ms = ms;
    return 0;
}

// `scratch` must hold blocks*256 floats.  Returns 0 or a negative error code.
extern "C" int irbfn_fp32_microbench(int mode, int iters, int blocks, float* scratch,
                                     double* fp32_lane_ops_per_s, double* mufu_lane_ops_per_s,
                                     float* elapsed_ms, void* stream_) {
    double r[3];
    const int rc = run_mode(mode, iters, blocks, scratch, r, elapsed_ms, (cudaStream_t)stream_);
    if (rc != 0) return rc;
    if (fp32_lane_ops_per_s) *fp32_lane_ops_per_s = r[0];
    if (mufu_lane_ops_per_s) *mufu_lane_ops_per_s = r[1];
    return 0;
}

// Same, all three pipes: rates[0] fp32, rates[1] MUFU, rates[2] fp64 lane-operations per second.
extern "C" int irbfn_pipe_microbench(int mode, int iters, int blocks, float* scratch, double* rates,
                                     float* elapsed_ms, void* stream_) {
    if (!rates) return -1;
    return run_mode(mode, iters, blocks, scratch, rates, elapsed_ms, (cudaStream_t)stream_);
}

#ifdef IRBFN_PEAK_MAIN
#include <stdio.h>
#include <stdlib.h>
int main(int argc, char** argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1;   // run a single mode (for ncu)
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("device %s sms=%d clock_khz=%d\n", p.name, p.multiProcessorCount, clk);
    int blocks = p.multiProcessorCount * 8;
    float* scratch;
    cudaMalloc(&scratch, sizeof(float) * blocks * 512);
    for (int rep = 0; rep < (only >= 0 ? 1 : 2); ++rep)
        for (int mode = (only >= 0 ? only : (argc > 2 ? atoi(argv[2]) : 0)); mode <= (only >= 0 ? only : 56); ++mode) {
            double r[3];
            float ms;
            int rc = irbfn_pipe_microbench(mode, only >= 0 ? 2000 : ((mode == 10 || mode == 11) ? 4000 : 20000), blocks, scratch, r, &ms, 0);
            const double u = 1.0 / p.multiProcessorCount / 1.965e9;
            printf("mode %2d rc=%d ms=%8.3f  per SM per clk @1.965GHz: fp32 %6.1f  mufu %5.2f  fp64 %5.1f\n",
                   mode, rc, ms, r[0] * u, r[1] * u, r[2] * u);
        }
    return 0;
}
#endif
