// Small-batch forward: the latency path of pred_step (src/irbfn_mpc/irbfn_planner.py:29-32 evaluates ONE query
// per control step; the notebooks 400).
//
// The throughput kernel (irbfn_fwd.cuh) is query-stationary: a CTA owns 1,024 queries and walks every
// (region, kernel) pair, so a batch of a few queries keeps a handful of CTAs busy for the whole walk, or -- split
// over kernels -- pays a second launch for the partial sums.  Here the roles are swapped: a CTA owns QC queries
// and its 256 threads share the (region, kernel) pairs, each thread blending its pairs straight into O output
// accumulators; one fixed-order block reduction closes the call.  One launch, no workspace, every parameter read
// once per CTA from L2.  The arithmetic is the throughput kernel's (direct-form distance, MUFU basis on the
// pre-scaled argument, the same gate sigmoids); only the summation order over pairs differs.
#include "irbfn_common.cuh"
#include "irbfn_plan.h"

namespace irbfn {

namespace {

constexpr int kSmallNT = 256;

// gamma_r(x) of one query: the product of model.py:88-93, factors as in gate_columns (irbfn_common.cuh)
__device__ __forceinline__ float small_gate(const GateDev& gd, const float* xs, int r) {
    if (r >= gd.num_ranges) return 0.f;
    float g = 1.f;
    for (int j = 0; j < gd.S; ++j) {
        const int i = __ldg(gd.offset + j) + __ldg(gd.ranges + (size_t)r * gd.S + j);
        const float k2 = 2.0f * kLog2e * __ldg(gd.delta + j);
        g *= gate_sigmoid(k2, xs[j] - __ldg(gd.lower + i)) * gate_sigmoid(k2, __ldg(gd.upper + i) - xs[j]);
    }
    return g;
}

template <int BASIS, int QC, int OT>
__global__ void __launch_bounds__(kSmallNT) irbfn_fwd_small_kernel(const FwdSmallArgs a) {
    extern __shared__ __align__(16) float smem[];
    float* xs = smem;                         // [QC][8]
    float* gs = xs + QC * 8;                  // [QC][R]
    float* red = gs + (size_t)QC * a.R;       // [warps][QC][OT]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int R = a.R, K = a.K, D = a.D, O = a.O;
    const long long q0 = (long long)blockIdx.x * QC;

    if (tid < QC * 8) {
        const int q = tid >> 3, j = tid & 7;
        xs[tid] = (q0 + q < a.B && j < D) ? __ldg(a.x + (q0 + q) * D + j) : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < QC * R; i += kSmallNT) {
        const int q = i / R, r = i - q * R;
        float g = 0.f;
        if (q0 + q < a.B)
            g = a.gate_ext != nullptr ? __ldg(a.gate_ext + (q0 + q) * R + r) : small_gate(a.gd, xs + q * 8, r);
        gs[i] = g;
    }
    __syncthreads();

    float xq[QC][8];
#pragma unroll
    for (int q = 0; q < QC; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) xq[q][j] = xs[q * 8 + j];

    float acc[QC][OT];
#pragma unroll
    for (int q = 0; q < QC; ++q)
#pragma unroll
        for (int o = 0; o < OT; ++o) acc[q][o] = 0.f;

    const int pairs = R * K;
    int r = tid / K, k = tid - r * K;                       // pair p = r*K + k, advanced by 256 per round
    const int dr = kSmallNT / K, dk = kSmallNT - dr * K;
    for (int p = tid; p < pairs; p += kSmallNT) {
        const float* c = a.centers + (size_t)p * D;
        float cj[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) cj[j] = j < D ? __ldg(c + j) : 0.f;
        const float sc = width_scale<BASIS>(__ldg(a.log_sigs + p));
        float wv[OT];
#pragma unroll
        for (int o = 0; o < OT; ++o) wv[o] = o < O ? __ldg(a.W + (size_t)k * O + o) : 0.f;
#pragma unroll
        for (int q = 0; q < QC; ++q) {
            float d2 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = xq[q][j] - cj[j];
                d2 = j == 0 ? d * d : fmaf(d, d, d2);
            }
            const float phi = BASIS == IRBFN_BASIS_GAUSSIAN ? ex2_approx(d2 * sc) : rcp_approx(fmaf(d2, sc, 1.0f));
            const float v = gs[q * R + r] * phi;
#pragma unroll
            for (int o = 0; o < OT; ++o) acc[q][o] = fmaf(v, wv[o], acc[q][o]);
        }
        r += dr;
        k += dk;
        if (k >= K) {
            k -= K;
            ++r;
        }
    }

    // fixed-order reduction: lanes (xor tree), then the eight warps in order, then the bias
#pragma unroll
    for (int q = 0; q < QC; ++q)
#pragma unroll
        for (int o = 0; o < OT; ++o) {
            float v = acc[q][o];
#pragma unroll
            for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
            if (lane == 0) red[(warp * QC + q) * OT + o] = v;
        }
    __syncthreads();
    if (tid < QC * OT) {
        const int q = tid / OT, o = tid - q * OT;
        if (q0 + q < a.B && o < O) {
            float v = 0.f;
#pragma unroll
            for (int w = 0; w < kSmallNT / 32; ++w) v += red[(w * QC + q) * OT + o];
            a.y[(q0 + q) * O + o] = v + __ldg(a.bias + o);
        }
    }
}

template <int BASIS, int QC, int OT>
int launch_inst(const FwdSmallArgs& a, cudaStream_t st) {
    const long long ctas = (a.B + QC - 1) / QC;
    const size_t smem = ((size_t)QC * 8 + (size_t)QC * a.R + (size_t)(kSmallNT / 32) * QC * OT) * sizeof(float);
    irbfn_fwd_small_kernel<BASIS, QC, OT><<<(unsigned)ctas, kSmallNT, smem, st>>>(a);
    return (int)cudaGetLastError();
}

template <int BASIS, int QC>
int launch_ot(const FwdSmallArgs& a, cudaStream_t st) {
    switch (fwd_ot_for(a.O)) {
        case 2: return launch_inst<BASIS, QC, 2>(a, st);
        case 4: return launch_inst<BASIS, QC, 4>(a, st);
        case 8: return launch_inst<BASIS, QC, 8>(a, st);
        case 10: return launch_inst<BASIS, QC, 10>(a, st);
        default: return launch_inst<BASIS, QC, 16>(a, st);
    }
}

}  // namespace

int launch_fwd_small(const FwdSmallArgs& a, int basis, cudaStream_t st) {
    // four queries per CTA once one query per CTA would need more than one CTA per SM (a CTA re-reads every parameter)
    const bool q4 = a.B > kFwdSmallQ1MaxB;
    if (basis == IRBFN_BASIS_GAUSSIAN)
        return q4 ? launch_ot<IRBFN_BASIS_GAUSSIAN, 4>(a, st) : launch_ot<IRBFN_BASIS_GAUSSIAN, 1>(a, st);
    return q4 ? launch_ot<IRBFN_BASIS_INVERSE_QUADRATIC, 4>(a, st) : launch_ot<IRBFN_BASIS_INVERSE_QUADRATIC, 1>(a, st);
}

}  // namespace irbfn
