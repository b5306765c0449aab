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
