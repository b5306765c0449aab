// Gate-sparse, region-binned forward: skip the local networks whose gate cannot change the
// result, and visit the queries in an order that makes the skipping uniform across a warp.
//
// Reference arithmetic being evaluated: WCRBFNet.__call__, src/irbfn_mpc/model.py:169-198 -- unchanged; this
// only drops terms gamma_r * phi_rk * W_ko whose total is provably below the caller's absolute tolerance:
//     |y_o - y_o(dense)| <= sum over dropped r of gamma_r * sum_k |W_ko|      (phi <= 1 for both bases)
// so with  thr = tol_abs / (R * max_o sum_k |W_ko|)  dropping every region with gamma_r <= thr keeps every
// output within tol_abs of the dense result (plus ordinary fp32 rounding).
//
// Passes (all on the caller's stream, all in the caller's workspace):
//   1. wabs_thr_kernel     thr from W (one block)
//   2. sparse_key_kernel   per query: signature = one bit per gate interval, f_j[i](x_j) > thr.  A region can only be
//                          non-negligible if all its interval factors are (the others are <= 1), so queries with equal
//                          signatures have (nearly) equal active sets.  Histogram of the signatures by atomics.
//   3. sparse_scan_kernel  exclusive scan of the histogram (one block)
//   4. sparse_scatter_kernel   perm = queries ordered by signature (counting sort; the order inside a bin is
//                          arbitrary, which cannot change any query's result)
//   5. irbfn_fwd_kernel<.., SPARSE>  (irbfn_fwd.cuh): x gathered / y scattered through perm; per warp, the regions
//                          whose exact gate is <= thr for all of its 128 queries are skipped; with parameters that
//                          are staged in chunks only the CTA's active regions are staged.
#include <cuda_runtime.h>

#include "irbfn_common.cuh"
#include "irbfn_fwd.cuh"   // FwdShape
#include "irbfn_plan.h"

namespace irbfn {
int launch_fwdm_d4(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);   // irbfn_fwd_modes.cuh
int launch_fwdm_d5(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d6(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d7(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d8(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int device_info(int* sms, int* smem_optin);                                       // irbfn_api.cu
int check_gate(const float* gate, const IrbfnGateDesc* gd, int D, GateDev* out);  // irbfn_api.cu

namespace {

constexpr int kKeyBitsMax = 20;   // signature bits used for binning (more intervals are folded in)

__global__ void __launch_bounds__(256) wabs_thr_kernel(const float* __restrict__ W, int K, int O, int R,
                                                       float tol_abs, float* __restrict__ thr) {
    __shared__ float red[256];
    float best = 0.f;
    for (int o = 0; o < O; ++o) {
        float s = 0.f;
        for (int k = threadIdx.x; k < K; k += 256) s += fabsf(__ldg(W + (size_t)k * O + o));
        red[threadIdx.x] = s;
        __syncthreads();
        for (int h = 128; h > 0; h >>= 1) {
            if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
            __syncthreads();
        }
        best = fmaxf(best, red[0]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        // best == 0: the readout is zero, every region is negligible (thr = +inf keeps y = bias)
        *thr = tol_abs / ((float)R * best);
    }
}

__device__ __forceinline__ unsigned signature(const float* __restrict__ xr, const GateDev& gd, float thr, int key_bits) {
    unsigned long long bits = 0ull;
    for (int j = 0; j < gd.S; ++j) {
        const float xv = __ldg(xr + j);
        const float k2 = 2.0f * kLog2e * __ldg(gd.delta + j);
        for (int i = __ldg(gd.offset + j); i < __ldg(gd.offset + j + 1); ++i) {
            const float f = gate_sigmoid(k2, xv - __ldg(gd.lower + i)) * gate_sigmoid(k2, __ldg(gd.upper + i) - xv);
            if (f > thr) bits |= 1ull << (i & 63);
        }
    }
    unsigned key = 0;
    for (int sh = 0; sh < 64; sh += key_bits) key ^= (unsigned)((bits >> sh) & ((1ull << key_bits) - 1ull));
    return key;
}

// Lattice sweeps put most queries in a handful of bins: one global atomic per query would serialise on those
// addresses.  Lanes of a warp with equal keys elect a leader (match.any) that adds their count once.
__device__ __forceinline__ unsigned warp_agg_add(unsigned* base, unsigned key, bool valid, int* rank_out) {
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    unsigned pos = 0;
    if (valid) {
        const unsigned peers = __match_any_sync(act, key);
        const int lane = threadIdx.x & 31;
        const int leader = __ffs(peers) - 1;
        unsigned b = 0;
        if (lane == leader) b = atomicAdd(base + key, (unsigned)__popc(peers));
        b = __shfl_sync(peers, b, leader);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        if (rank_out) *rank_out = rank;
        pos = b + (unsigned)rank;
    }
    return pos;
}

__global__ void __launch_bounds__(256) sparse_key_kernel(const float* __restrict__ x, GateDev gd,
                                                         const float* __restrict__ thr, unsigned* __restrict__ keys,
                                                         unsigned* __restrict__ hist, long long B, int D, int key_bits) {
    const float t = __ldg(thr);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q0 = (long long)blockIdx.x * blockDim.x; q0 < B; q0 += stride) {   // warp-uniform trip count
        const long long q = q0 + threadIdx.x;
        const bool valid = q < B;
        unsigned key = 0;
        if (valid) {
            key = signature(x + q * D, gd, t, key_bits);
            keys[q] = key;
        }
        warp_agg_add(hist, key, valid, nullptr);
    }
}

// exclusive scan of hist[0 .. nb) in place, one block of 1,024 threads: thread t owns a contiguous slab
__global__ void __launch_bounds__(1024) sparse_scan_kernel(unsigned* __restrict__ hist, int nb) {
    __shared__ unsigned part[1024];
    const int per = (nb + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(nb, lo + per);
    unsigned s = 0;
    for (int i = lo; i < hi; ++i) s += hist[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        const unsigned v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = threadIdx.x > 0 ? part[threadIdx.x - 1] : 0u;
    for (int i = lo; i < hi; ++i) {
        const unsigned c = hist[i];
        hist[i] = run;
        run += c;
    }
}

__global__ void __launch_bounds__(256) sparse_scatter_kernel(const unsigned* __restrict__ keys, unsigned* __restrict__ cursor,
                                                             int* __restrict__ perm, long long B) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q0 = (long long)blockIdx.x * blockDim.x; q0 < B; q0 += stride) {
        const long long q = q0 + threadIdx.x;
        const bool valid = q < B;
        const unsigned key = valid ? keys[q] : 0u;
        const unsigned pos = warp_agg_add(cursor, key, valid, nullptr);
        if (valid) perm[pos] = (int)q;
    }
}

struct SparseLayout {
    size_t thr_off, skip_off, work_off, hist_off, keys_off, perm_off, total;
    int key_bits, nb;
};

inline size_t a256(size_t v) { return (v + 255) & ~(size_t)255; }

SparseLayout sparse_layout(int64_t B, int num_bounds) {
    SparseLayout L;
    L.key_bits = num_bounds < kKeyBitsMax ? (num_bounds < 1 ? 1 : num_bounds) : kKeyBitsMax;
    L.nb = 1 << L.key_bits;
    size_t off = 0;
    L.thr_off = off; off = a256(off + sizeof(float));
    L.skip_off = off; off = a256(off + 2 * sizeof(unsigned long long));
    L.work_off = off; off = a256(off + 8 * sizeof(unsigned));          // one work counter per output window
    L.hist_off = off; off = a256(off + (size_t)L.nb * sizeof(unsigned));
    L.keys_off = off; off = a256(off + (size_t)B * sizeof(unsigned));
    L.perm_off = off; off = a256(off + (size_t)B * sizeof(int));
    L.total = off;
    return L;
}

}  // namespace
}  // namespace irbfn

using namespace irbfn;

extern "C" size_t irbfn_forward_sparse_workspace_bytes(int64_t B, int num_bounds) {
    if (B <= 0 || num_bounds < 0) return 0;
    return sparse_layout(B, num_bounds).total;
}

extern "C" int irbfn_forward_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                                    const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                                    int D, int O, int basis, float tol_abs, uint64_t* counters, void* workspace,
                                    size_t workspace_bytes, irbfn_stream_t stream) {
    if (B < 0 || R <= 0 || K <= 0 || D <= 0 || O <= 0) return IRBFN_ERR_BAD_DIMS;
    if (B > 0x7fffffffll) return IRBFN_ERR_BAD_DIMS;                   // perm holds int32 row numbers
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (basis != IRBFN_BASIS_GAUSSIAN && basis != IRBFN_BASIS_INVERSE_QUADRATIC) return IRBFN_ERR_BAD_BASIS;
    if (!centers || !log_sigs || !W || !bias || !y) return IRBFN_ERR_NULL_POINTER;
    if (!(tol_abs >= 0.f)) return IRBFN_ERR_BAD_DIMS;
    if (gd == nullptr) return IRBFN_ERR_BAD_GATE;                      // the tanh gate only
    GateDev gdev;
    int rc = check_gate(nullptr, gd, D, &gdev);
    if (rc != IRBFN_OK) return rc;
    if (B == 0) return IRBFN_OK;
    if (!x) return IRBFN_ERR_NULL_POINTER;
    int sms, smem;
    rc = device_info(&sms, &smem);
    if (rc != IRBFN_OK) return rc;
    const SparseLayout L = sparse_layout(B, gdev.num_bounds);
    if (workspace == nullptr || workspace_bytes < L.total) return IRBFN_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    float* thr = (float*)(ws + L.thr_off);
    unsigned long long* skip = (unsigned long long*)(ws + L.skip_off);
    unsigned* hist = (unsigned*)(ws + L.hist_off);
    unsigned* keys = (unsigned*)(ws + L.keys_off);
    int* perm = (int*)(ws + L.perm_off);

    if (cudaMemsetAsync(ws + L.skip_off, 0, L.keys_off - L.skip_off, st) != cudaSuccess)      // counters + histogram
        return note_cuda_error("cudaMemsetAsync", (int)cudaGetLastError());
    wabs_thr_kernel<<<1, 256, 0, st>>>(W, K, O, R, tol_abs, thr);
    const int nblk = (int)((B + 255) / 256 < (long long)sms * 16 ? (B + 255) / 256 : (long long)sms * 16);
    sparse_key_kernel<<<nblk, 256, 0, st>>>(x, gdev, thr, keys, hist, B, D, L.key_bits);
    sparse_scan_kernel<<<1, 1024, 0, st>>>(hist, L.nb);
    sparse_scatter_kernel<<<nblk, 256, 0, st>>>(keys, hist, perm, B);
    { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("sparse binning", (int)ce); }

    const int DT = D < 4 ? 4 : D;
    int o0 = 0;
    do {
        const int ocur = O - o0 > 16 ? 16 : O - o0;
        const int OT = fwd_ot_for(ocur);
        FwdPlan plan;
        if (!plan_forward(R, K, DT, OT, smem, &plan)) return IRBFN_ERR_BAD_DIMS;
        if (plan.RC > 64) {
            // more regions per stage than the mask holds: cut the region chunk (the stage then cycles)
            FwdPlan p2;
            bool ok = false;
            for (int RC = 64; RC >= 1 && !ok; RC >>= 1) {
                const size_t fixed = (size_t)RC * plan.NT * kFwdQ * sizeof(float);
                if (fixed >= (size_t)smem) continue;
                const size_t per_k = ((size_t)RC * fwd_tile_floats(DT) / kFwdKT + OT) * sizeof(float);
                long long KC = (long long)(((size_t)smem - fixed) / per_k);
                KC = (KC / kFwdKT) * kFwdKT;
                const int Kp = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
                if (KC > Kp) KC = Kp;
                if (KC >= kFwdKT) { p2 = FwdPlan{RC, (int)KC, 0, fwd_smem_bytes(RC, (int)KC, DT, OT, plan.NT), plan.NT}; ok = true; }
            }
            if (!ok) return IRBFN_ERR_BAD_DIMS;
            plan = p2;
        }
        FwdArgs a{};
        a.nsplit = 1;
        a.k_per_split = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
        a.split_stride = 0;
        a.x = x; a.centers = centers; a.log_sigs = log_sigs; a.W = W; a.bias = bias;
        a.gate_ext = nullptr; a.gd = gdev; a.y = y; a.hidden = nullptr; a.gate_out = nullptr;
        a.perm = perm; a.gate_thr = thr; a.skip_count = o0 == 0 ? skip : nullptr;
        a.work_counter = (o0 / 16 < 8) ? (unsigned*)(ws + L.work_off) + o0 / 16 : nullptr;
        a.B = B; a.R = R; a.K = K; a.Dsrc = D; a.ldw = O; a.o0 = o0; a.ocur = ocur;
        a.RC = plan.RC; a.KC = plan.KC; a.stage_once = plan.stage_once;
        const long long tq = (long long)plan.NT * FwdShape::Q;
        const long long tiles = (B + tq - 1) / tq;
        a.num_tiles = (int)tiles;
        const int grid = (int)(tiles < sms ? tiles : sms);
        int e;
        switch (DT) {
            case 4: e = launch_fwdm_d4(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            case 5: e = launch_fwdm_d5(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            case 6: e = launch_fwdm_d6(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            case 7: e = launch_fwdm_d7(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            default: e = launch_fwdm_d8(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
        }
        if (e != 0) return note_cuda_error("irbfn_fwd_kernel<sparse>", e);
        o0 += ocur;
    } while (o0 < O);
    if (counters != nullptr) {
        if (cudaMemcpyAsync(counters, skip, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            return note_cuda_error("cudaMemcpyAsync", (int)cudaGetLastError());
    }
    return IRBFN_OK;
}
