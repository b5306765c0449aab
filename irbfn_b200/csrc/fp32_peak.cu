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

// fp32-pipe lane-operations (FADD/FMUL/FFMA lanes) and MUFU lane-operations per thread per iteration
void mode_ops(int mode, double* fp32_ops, double* mufu_ops) {
    switch (mode) {
        case 0: *fp32_ops = 8 * 8 * 2; *mufu_ops = 0; break;
        case 1: *fp32_ops = 8 * 8 * 2; *mufu_ops = 0; break;
        case 2: *fp32_ops = 4 * 8 * 4; *mufu_ops = 0; break;
        case 3: *fp32_ops = 0; *mufu_ops = 4 * 8 * 2; break;
        case 4: *fp32_ops = 8 * (8 * 4 + 2 + 2); *mufu_ops = 8 * 2; break;
        case 5: *fp32_ops = 8 * (8 * 4 + 2 + 2); *mufu_ops = 8 * 2; break;
        case 6: *fp32_ops = 8 * 9 * 2; *mufu_ops = 8 * 2; break;
        default: *fp32_ops = 0; *mufu_ops = 0;
    }
}

}  // namespace

// Runs one mode and reports fp32-pipe lane-ops/s and MUFU lane-ops/s.
// `scratch` must hold blocks*256 floats.  Returns 0 or a negative error code.
extern "C" int irbfn_fp32_microbench(int mode, int iters, int blocks, float* scratch,
                                     double* fp32_lane_ops_per_s, double* mufu_lane_ops_per_s,
                                     float* elapsed_ms, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (mode < 0 || mode > 6 || iters <= 0 || blocks <= 0 || !scratch) return -1;
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
    double f, m;
    mode_ops(mode, &f, &m);
    double threads = (double)blocks * 256.0;
    if (fp32_lane_ops_per_s) *fp32_lane_ops_per_s = f * iters * threads / (ms * 1e-3);
    if (mufu_lane_ops_per_s) *mufu_lane_ops_per_s = m * iters * threads / (ms * 1e-3);
    if (elapsed_ms) *elapsed_ms = ms;
    return 0;
}

#ifdef IRBFN_PEAK_MAIN
#include <stdio.h>
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("device %s sms=%d clock_khz=%d\n", p.name, p.multiProcessorCount, clk);
    int blocks = p.multiProcessorCount * 8;
    float* scratch;
    cudaMalloc(&scratch, sizeof(float) * blocks * 256);
    for (int rep = 0; rep < 2; ++rep)
        for (int mode = 0; mode <= 6; ++mode) {
            double f, m;
            float ms;
            int rc = irbfn_fp32_microbench(mode, 20000, blocks, scratch, &f, &m, &ms, 0);
            printf("mode %d rc=%d ms=%.3f fp32_lane_Gops=%.1f (per SM per clk @1.965GHz: %.1f) mufu_lane_Gops=%.1f (per SM per clk: %.2f)\n",
                   mode, rc, ms, f * 1e-9, f / p.multiProcessorCount / 1.965e9, m * 1e-9,
                   m / p.multiProcessorCount / 1.965e9);
        }
    return 0;
}
#endif
