// Shared device helpers and argument structs for the IRBFN sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/irbfn_b200.h"

namespace irbfn {

constexpr float kLog2e = 1.4426950408889634f;

// Diagnostics for IRBFN_ERR_LAUNCH: the failing step and the CUDA error, kept per host thread
// (irbfn_last_error_detail()).  Defined in irbfn_api.cu.
int note_cuda_error(const char* where, int cuda_error);  // returns IRBFN_ERR_LAUNCH

// ---- packed fp32 (FFMA2 / FADD2 / FMUL2 on sm_100a: one issue slot, two fp32 lanes) ----
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 dup2(float a) { return make_float2(a, a); }

// ---- XU (MUFU) wrappers ----
__device__ __forceinline__ float ex2_approx(float v) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ float rcp_approx(float v) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

// (tanh(t) + 1) / 2 == 1 / (1 + exp(-2 t)).   k2 = 2 * delta * log2(e);  z = signed distance
// to the bound.  Saturates cleanly: ex2 -> inf gives 1/inf = 0, ex2 -> 0 gives 1.
__device__ __forceinline__ float gate_sigmoid(float k2, float z) {
    return __fdividef(1.0f, 1.0f + ex2_approx(-k2 * z));
}

// Device-side copy of IrbfnGateDesc (passed by value in kernel arguments; the arrays it points
// to are device memory owned by the caller).
struct GateDev {
    int S;
    int num_ranges;
    int num_bounds;
    const float* lower;
    const float* upper;
    const int* offset;
    const float* delta;
    const int* ranges;
};

// A thread's four consecutive query rows -> two packed pairs x2[p][j] = (x[q0+2p][j], x[q0+2p+1][j]).
// Scalar loads on purpose: each lands directly in its half of an aligned register pair.  (LDG.128 over the
// 16*D contiguous bytes was tried: the four-register destinations put the two halves of a pair in different
// quads, and ptxas then re-forms the pairs with ~115 MOVs INSIDE the hot loop, 176 -> 292 instructions.)
template <int D>
__device__ __forceinline__ void load_queries4(const float* __restrict__ x, long long q0, long long B, int Dsrc,
                                              float2 (&x2)[2][D]) {
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            x2[p][j].x = (qa < B && j < Dsrc) ? __ldg(x + qa * Dsrc + j) : 0.f;
            x2[p][j].y = (qb < B && j < Dsrc) ? __ldg(x + qb * Dsrc + j) : 0.f;
        }
    }
}

// Region gate of a thread's four queries q0..q0+3 for regions [r0, r0 + rc), written to the thread-private
// shared-memory column gcol[rr * stride]  (model.py:70-95: gamma_r = prod_j f_j[ranges[r][j]],
// f_j[i] = s(x_j - lb_i) s(ub_i - x_j)).
//   * Split dimensions with ONE interval give the same factor to every region: they are multiplied in
//     registers first and the table starts from that product; only dimensions with several intervals touch
//     the table, four intervals' factors at a time held in registers and selected per region by its
//     (warp-uniform) interval index.  For the shipped 12-region Frenet gate (8 split dimensions, 12
//     intervals, 3 dimensions with more than one) that is 36 table updates instead of 144.
//   * Every loop is a RUNTIME loop and x_j is re-read from global memory (an L1 hit: the tile's rows were
//     just loaded) instead of indexing the register copy, which would force the dimension loop to unroll
//     into ~4,000 straight-line instructions executed once per tile (instruction-fetch stalls in ncu).
//     Measured cost of the fused gate at C2: ~5 us of a 125 us tile (scripts/dev_gate_cost.py).
__device__ __forceinline__ void gate_columns(const GateDev& gd, const float* __restrict__ x, long long q0,
                                             long long B, int Dsrc, int r0, int rc, float4* gcol, int stride) {
    const int S = gd.S;
    const int rlim = min(rc, gd.num_ranges - r0);
    auto load4 = [&](int j) {
        float4 v;
        v.x = q0 + 0 < B ? __ldg(x + (q0 + 0) * Dsrc + j) : 0.f;
        v.y = q0 + 1 < B ? __ldg(x + (q0 + 1) * Dsrc + j) : 0.f;
        v.z = q0 + 2 < B ? __ldg(x + (q0 + 2) * Dsrc + j) : 0.f;
        v.w = q0 + 3 < B ? __ldg(x + (q0 + 3) * Dsrc + j) : 0.f;
        return v;
    };
    auto factor = [&](const float4& v, float k2, int i) {
        const float lb = __ldg(gd.lower + i), ub = __ldg(gd.upper + i);
        float4 f;
        f.x = gate_sigmoid(k2, v.x - lb) * gate_sigmoid(k2, ub - v.x);
        f.y = gate_sigmoid(k2, v.y - lb) * gate_sigmoid(k2, ub - v.y);
        f.z = gate_sigmoid(k2, v.z - lb) * gate_sigmoid(k2, ub - v.z);
        f.w = gate_sigmoid(k2, v.w - lb) * gate_sigmoid(k2, ub - v.w);
        return f;
    };
    float4 com = make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll 1
    for (int j = 0; j < S; ++j) {
        const int ob = __ldg(gd.offset + j), oe = __ldg(gd.offset + j + 1);
        if (oe - ob == 1) {
            const float4 f = factor(load4(j), 2.0f * kLog2e * __ldg(gd.delta + j), ob);
            com.x *= f.x; com.y *= f.y; com.z *= f.z; com.w *= f.w;
        }
    }
#pragma unroll 1
    for (int rr = 0; rr < rc; ++rr) gcol[rr * stride] = rr < rlim ? com : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int j = 0; j < S; ++j) {
        const int ob = __ldg(gd.offset + j), oe = __ldg(gd.offset + j + 1);
        if (oe - ob > 1) {
            const float k2 = 2.0f * kLog2e * __ldg(gd.delta + j);
            const float4 v = load4(j);
            const int* rg = gd.ranges + (size_t)r0 * S + j;
#pragma unroll 1
            for (int ib = ob; ib < oe; ib += 4) {
                const float4 f0 = factor(v, k2, ib);
                const float4 f1 = ib + 1 < oe ? factor(v, k2, ib + 1) : f0;
                const float4 f2 = ib + 2 < oe ? factor(v, k2, ib + 2) : f0;
                const float4 f3 = ib + 3 < oe ? factor(v, k2, ib + 3) : f0;
#pragma unroll 2
                for (int rr = 0; rr < rlim; ++rr) {
                    const int idx = __ldg(rg + rr * S) - (ib - ob);
                    if ((unsigned)idx < 4u) {
                        const float4 lo = idx & 1 ? f1 : f0, hi = idx & 1 ? f3 : f2;
                        const float4 f = idx & 2 ? hi : lo;
                        float4 g = gcol[rr * stride];
                        g.x *= f.x; g.y *= f.y; g.z *= f.z; g.w *= f.w;
                        gcol[rr * stride] = g;
                    }
                }
            }
        }
    }
}

// The same table for four ARBITRARY rows (gate-sparse forward: a permuted batch): identical arithmetic, the row
// loader is a parameter.  (The consecutive-row version above is kept verbatim from round 1: routing it through
// this template changed ptxas' schedule of the dense forward.)
template <typename Load4>
__device__ __forceinline__ void gate_columns_impl(const GateDev& gd, Load4 load4, int r0, int rc, float4* gcol,
                                                  int stride) {
    const int S = gd.S;
    const int rlim = min(rc, gd.num_ranges - r0);
    auto factor = [&](const float4& v, float k2, int i) {
        const float lb = __ldg(gd.lower + i), ub = __ldg(gd.upper + i);
        float4 f;
        f.x = gate_sigmoid(k2, v.x - lb) * gate_sigmoid(k2, ub - v.x);
        f.y = gate_sigmoid(k2, v.y - lb) * gate_sigmoid(k2, ub - v.y);
        f.z = gate_sigmoid(k2, v.z - lb) * gate_sigmoid(k2, ub - v.z);
        f.w = gate_sigmoid(k2, v.w - lb) * gate_sigmoid(k2, ub - v.w);
        return f;
    };
    float4 com = make_float4(1.f, 1.f, 1.f, 1.f);
#pragma unroll 1
    for (int j = 0; j < S; ++j) {
        const int ob = __ldg(gd.offset + j), oe = __ldg(gd.offset + j + 1);
        if (oe - ob == 1) {
            const float4 f = factor(load4(j), 2.0f * kLog2e * __ldg(gd.delta + j), ob);
            com.x *= f.x; com.y *= f.y; com.z *= f.z; com.w *= f.w;
        }
    }
#pragma unroll 1
    for (int rr = 0; rr < rc; ++rr) gcol[rr * stride] = rr < rlim ? com : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int j = 0; j < S; ++j) {
        const int ob = __ldg(gd.offset + j), oe = __ldg(gd.offset + j + 1);
        if (oe - ob > 1) {
            const float k2 = 2.0f * kLog2e * __ldg(gd.delta + j);
            const float4 v = load4(j);
            const int* rg = gd.ranges + (size_t)r0 * S + j;
#pragma unroll 1
            for (int ib = ob; ib < oe; ib += 4) {
                const float4 f0 = factor(v, k2, ib);
                const float4 f1 = ib + 1 < oe ? factor(v, k2, ib + 1) : f0;
                const float4 f2 = ib + 2 < oe ? factor(v, k2, ib + 2) : f0;
                const float4 f3 = ib + 3 < oe ? factor(v, k2, ib + 3) : f0;
#pragma unroll 2
                for (int rr = 0; rr < rlim; ++rr) {
                    const int idx = __ldg(rg + rr * S) - (ib - ob);
                    if ((unsigned)idx < 4u) {
                        const float4 lo = idx & 1 ? f1 : f0, hi = idx & 1 ? f3 : f2;
                        const float4 f = idx & 2 ? hi : lo;
                        float4 g = gcol[rr * stride];
                        g.x *= f.x; g.y *= f.y; g.z *= f.z; g.w *= f.w;
                        gcol[rr * stride] = g;
                    }
                }
            }
        }
    }
}

// four arbitrary rows (a permuted batch); qi < 0 = no query
__device__ __forceinline__ void gate_columns_rows(const GateDev& gd, const float* __restrict__ x,
                                                  const long long (&qi)[4], int Dsrc, int r0, int rc, float4* gcol,
                                                  int stride) {
    gate_columns_impl(gd, [&](int j) {
        float4 v;
        v.x = qi[0] >= 0 ? __ldg(x + qi[0] * Dsrc + j) : 0.f;
        v.y = qi[1] >= 0 ? __ldg(x + qi[1] * Dsrc + j) : 0.f;
        v.z = qi[2] >= 0 ? __ldg(x + qi[2] * Dsrc + j) : 0.f;
        v.w = qi[3] >= 0 ? __ldg(x + qi[3] * Dsrc + j) : 0.f;
        return v;
    }, r0, rc, gcol, stride);
}

// One thread per query, 256 per block: gamma (B,R) rows of the tanh gate.  Interval factors live in
// thread-private shared-memory COLUMNS fac[i * 256 + tid] (conflict-free) and the per-region factor
// indices offset[j] + ranges[r][j] in a block-shared table, built once per block instead of two dependent
// global loads per (query, region, dimension).  Dynamic shared memory: gate_rows_smem_bytes().
inline size_t gate_rows_smem_bytes(int num_bounds, int rows_per_bound, int num_ranges, int S) {
    return ((size_t)256 * num_bounds * rows_per_bound + (size_t)num_ranges * S) * sizeof(float);
}
__device__ __forceinline__ int* gate_index_table(const GateDev& gd, float* smem_after_factors, int rmax) {
    int* ridx = reinterpret_cast<int*>(smem_after_factors);
    for (int t = threadIdx.x; t < rmax * gd.S; t += blockDim.x) {
        const int r = t / gd.S, j = t - r * gd.S;
        ridx[t] = __ldg(gd.offset + j) + __ldg(gd.ranges + r * gd.S + j);
    }
    return ridx;
}
__device__ __forceinline__ void gate_rows_block(const float* __restrict__ x, const GateDev& gd,
                                                float* __restrict__ gamma, long long B, int R, int D,
                                                float* fac) {
    const int tid = threadIdx.x;
    const long long q = (long long)blockIdx.x * blockDim.x + tid;
    const int rmax = R < gd.num_ranges ? R : gd.num_ranges;
    const int* ridx = gate_index_table(gd, fac + (size_t)256 * gd.num_bounds, rmax);
    if (q < B) {
        for (int j = 0; j < gd.S; ++j) {
            const float xv = __ldg(x + q * D + j);
            const float k2 = 2.0f * kLog2e * __ldg(gd.delta + j);
            for (int i = __ldg(gd.offset + j); i < __ldg(gd.offset + j + 1); ++i)
                fac[i * 256 + tid] = gate_sigmoid(k2, xv - __ldg(gd.lower + i)) * gate_sigmoid(k2, __ldg(gd.upper + i) - xv);
        }
    }
    __syncthreads();
    if (q >= B) return;
    for (int r = 0; r < R; ++r) {
        float g = 0.f;
        if (r < rmax) {
            g = 1.f;
            for (int j = 0; j < gd.S; ++j) g *= fac[ridx[r * gd.S + j] * 256 + tid];
        }
        gamma[q * R + r] = g;
    }
}

// Forward launch arguments (one struct, passed by value).
struct FwdArgs {
    const float* x;         // (B,D)
    const float* centers;   // (R,K,D)
    const float* log_sigs;  // (R,K)
    const float* W;         // (K,ldw) ; columns [o0, o0+ocur) are used
    const float* bias;      // (O)
    const float* gate_ext;  // (B,R) or nullptr -> fused tanh gate from gd
    GateDev gd;
    float* y;               // (B,ldw) or nullptr
    float* hidden;          // (B,K) or nullptr
    long long B;
    int R, K;
    int Dsrc;               // features in memory (<= the kernel's compiled D; the rest is zero-padded)
    int ldw;                // row stride of W and y (= O)
    int o0, ocur;           // output column window handled by this launch
    int RC;                 // regions per stage
    int KC;                 // kernels per stage (multiple of the kernel tile)
    int num_tiles;          // ceil(B / queries-per-block)
    int stage_once;         // 1: a single (region, kernel) stage -> staged once per block
    // small-batch split over kernels: blockIdx.y owns kernels [y * k_per_split, (y+1) * k_per_split) of
    // every region and writes its partial readout to y + blockIdx.y * split_stride (bias from split 0)
    int nsplit;             // gridDim.y
    int k_per_split;        // multiple of the kernel tile; >= K when gridDim.y == 1
    long long split_stride; // floats between partial outputs (0 when gridDim.y == 1)
    // ---- only read by the irbfn_fwd_modes.cuh instantiations; kept BEHIND the round-1 fields so that the plain
    // ---- dense kernel's parameter offsets (and with them its ptxas schedule) stay what they were
    float* gate_out;        // (B,R) or nullptr: the fused tanh gate is also written out (training: the
                            // parameter-gradient kernel reads it instead of recomputing it)
    // gate-sparse evaluation (SPARSE instantiation only): queries are visited in the order perm[0..B), which
    // groups queries with the same set of non-negligible regions; a region whose gate is <= *gate_thr for all
    // 128 queries of a warp is skipped by that warp
    const int* perm;        // (B) or nullptr
    const float* gate_thr;  // device scalar
    unsigned long long* skip_count;   // optional: number of (warp, kernel-chunk, region) evaluations skipped / done
    unsigned* work_counter; // SPARSE with resident parameters: warps fetch their next 128-query unit here (zeroed
                            // by the caller): tiles differ 2-3x in active regions, a static split leaves a long tail
    // streamed input (STREAM instantiation only, irbfn_forward_streamed): the rows of x arrive WHILE the kernel runs;
    // ready[p] != 0 once upload piece p (tiles_per_flag tiles of 1,024 queries) has landed -- written behind the
    // piece's data on its copy stream; work_counter (the word behind the flags) hands the tiles out.
    const unsigned* ready;
    int tiles_per_flag;
    int* stream_status;     // set to 1 if a flag did not arrive within the spin budget (the result is then undefined)
};

// Arguments of the small-batch (pair-parallel) forward, irbfn_fwd_small.cu.
struct FwdSmallArgs {
    const float* x;          // (B,D)
    const float* centers;    // (R,K,D)
    const float* log_sigs;   // (R,K)
    const float* W;          // (K,O)
    const float* bias;       // (O)
    const float* gate_ext;   // (B,R) or nullptr
    GateDev gd;              // used when gate_ext == nullptr
    float* y;                // (B,O)
    long long B;
    int R, K, D, O;
};

// log-sigma -> the per-kernel scale used in the hot loop.
//   gaussian         : phi = exp(-d2 / sigma^2) = ex2(d2 * nscale),  nscale = -exp(-2 ls) * log2(e)
//   inverse quadratic: phi = 1 / (1 + d2 / sigma^2) = rcp(fma(d2, scale, 1)),  scale = exp(-2 ls)
template <int BASIS>
__device__ __forceinline__ float width_scale(float log_sig) {
    float e = expf(-2.0f * log_sig);
    return BASIS == IRBFN_BASIS_GAUSSIAN ? -e * kLog2e : e;
}

}  // namespace irbfn
