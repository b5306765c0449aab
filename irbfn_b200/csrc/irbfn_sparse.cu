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

// Where an interval factor exceeds the threshold: f_i(x) = s(x - lb_i) s(ub_i - x) rises to its maximum at the
// interval's middle and falls again, so {x : f_i(x) > thr} is an open interval (edge[2i], edge[2i+1]), found once per
// call by bisection with the SAME arithmetic the gate uses.  The signature pass then needs two compares per interval
// instead of four MUFUs (it was 42 us per 2^20 queries, XU-bound; the signature only groups queries -- the exact
// per-warp test in the forward decides what is skipped, so an edge that is off by an ulp cannot change a result).
__global__ void __launch_bounds__(64) sparse_edges_kernel(GateDev gd, const float* __restrict__ thr_p,
                                                          float* __restrict__ edge) {
    const float thr = __ldg(thr_p);
    for (int i = threadIdx.x; i < gd.num_bounds; i += blockDim.x) {
        int j = 0;
        while (j + 1 < gd.S && __ldg(gd.offset + j + 1) <= i) ++j;
        const float delta = __ldg(gd.delta + j);
        const float k2 = 2.0f * kLog2e * delta;
        const float lb = __ldg(gd.lower + i), ub = __ldg(gd.upper + i);
        auto f = [&](float xv) { return gate_sigmoid(k2, xv - lb) * gate_sigmoid(k2, ub - xv); };
        const float mid = 0.5f * (lb + ub);
        float a = INFINITY, b = -INFINITY;                       // empty: the factor never exceeds thr
        if (f(mid) > thr) {
            const float reach = fabsf(ub - lb) + 1.0f + 200.0f / fmaxf(fabsf(delta), 1e-6f);   // f underflows to 0 out there
            float lo = mid - reach, hi = mid;
            if (f(lo) > thr) lo = -INFINITY;
            else for (int it = 0; it < 64; ++it) { const float m = 0.5f * (lo + hi); if (f(m) > thr) hi = m; else lo = m; }
            a = lo;
            lo = mid, hi = mid + reach;
            if (f(hi) > thr) hi = INFINITY;
            else for (int it = 0; it < 64; ++it) { const float m = 0.5f * (lo + hi); if (f(m) > thr) lo = m; else hi = m; }
            b = hi;
        }
        edge[2 * i] = a;
        edge[2 * i + 1] = b;
    }
}

__device__ __forceinline__ unsigned signature(const float* __restrict__ xr, const GateDev& gd,
                                              const float* __restrict__ edge, int key_bits) {
    unsigned long long bits = 0ull;
    for (int j = 0; j < gd.S; ++j) {
        const float xv = __ldg(xr + j);
        for (int i = __ldg(gd.offset + j); i < __ldg(gd.offset + j + 1); ++i)
            if (xv > edge[2 * i] && xv < edge[2 * i + 1]) bits |= 1ull << (i & 63);
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
                                                         const float* __restrict__ edge_g, unsigned* __restrict__ keys,
                                                         unsigned* __restrict__ hist, long long B, int D, int key_bits) {
    extern __shared__ float edge[];                       // [num_bounds][2]
    for (int i = threadIdx.x; i < 2 * gd.num_bounds; i += blockDim.x) edge[i] = __ldg(edge_g + i);
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q0 = (long long)blockIdx.x * blockDim.x; q0 < B; q0 += stride) {   // warp-uniform trip count
        const long long q = q0 + threadIdx.x;
        const bool valid = q < B;
        unsigned key = 0;
        if (valid) {
            key = signature(x + q * D, gd, edge, key_bits);
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

// Block-aggregated variants for key spaces that fit shared memory (nb <= kBlockBins; the shipped gates have 12-13
// intervals): a block owns a contiguous range of queries, counts its keys in a shared-memory histogram and touches each
// global bin ONCE.  Lattice sweeps put most queries into a handful of bins; the per-warp atomics of the kernels above
// then serialise on those addresses (40 + 35 us per 2^20 queries).
constexpr int kBlockBins = 8192;

__global__ void __launch_bounds__(256) sparse_key_block_kernel(const float* __restrict__ x, GateDev gd,
                                                               const float* __restrict__ edge_g, unsigned* __restrict__ keys,
                                                               unsigned* __restrict__ hist, long long B, int D,
                                                               int key_bits, long long per_block) {
    extern __shared__ unsigned sh[];                      // [nb] histogram, then [num_bounds][2] edges
    const int nb = 1 << key_bits;
    float* edge = reinterpret_cast<float*>(sh + nb);
    for (int i = threadIdx.x; i < nb; i += blockDim.x) sh[i] = 0u;
    for (int i = threadIdx.x; i < 2 * gd.num_bounds; i += blockDim.x) edge[i] = __ldg(edge_g + i);
    __syncthreads();
    const long long lo = (long long)blockIdx.x * per_block, hi = min(B, lo + per_block);
    for (long long q0 = lo; q0 < hi; q0 += blockDim.x) {            // warp-uniform trip count
        const long long q = q0 + threadIdx.x;
        const bool valid = q < hi;
        unsigned key = 0;
        if (valid) {
            key = signature(x + q * D, gd, edge, key_bits);
            keys[q] = key;
        }
        warp_agg_add(sh, key, valid, nullptr);                      // lanes with equal keys add once (no 32-way conflict)
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nb; i += blockDim.x) {
        const unsigned c = sh[i];
        if (c) atomicAdd(hist + i, c);
    }
}

__global__ void __launch_bounds__(256) sparse_scatter_block_kernel(const unsigned* __restrict__ keys,
                                                                   unsigned* __restrict__ cursor, int* __restrict__ perm,
                                                                   long long B, int nb, long long per_block) {
    extern __shared__ unsigned sh[];                      // [nb] counts -> running positions
    for (int i = threadIdx.x; i < nb; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    const long long lo = (long long)blockIdx.x * per_block, hi = min(B, lo + per_block);
    for (long long q = lo + threadIdx.x; q < hi; q += blockDim.x) atomicAdd(&sh[keys[q]], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < nb; i += blockDim.x) {
        const unsigned c = sh[i];
        if (c) sh[i] = atomicAdd(cursor + i, c);          // this block's slice of bin i starts here
    }
    __syncthreads();
    for (long long q = lo + threadIdx.x; q < hi; q += blockDim.x) perm[atomicAdd(&sh[keys[q]], 1u)] = (int)q;
}

struct SparseLayout {
    size_t thr_off, edge_off, skip_off, work_off, hist_off, keys_off, perm_off, total;
    int key_bits, nb;
};

inline size_t a256(size_t v) { return (v + 255) & ~(size_t)255; }

SparseLayout sparse_layout(int64_t B, int num_bounds) {
    SparseLayout L;
    L.key_bits = num_bounds < kKeyBitsMax ? (num_bounds < 1 ? 1 : num_bounds) : kKeyBitsMax;
    L.nb = 1 << L.key_bits;
    size_t off = 0;
    L.thr_off = off; off = a256(off + sizeof(float));
    L.edge_off = off; off = a256(off + 2 * (size_t)(num_bounds < 1 ? 1 : num_bounds) * sizeof(float));
    L.skip_off = off; off = a256(off + 2 * sizeof(unsigned long long));
    L.work_off = off; off = a256(off + 8 * sizeof(unsigned));          // one work counter per output window
    L.hist_off = off; off = a256(off + (size_t)L.nb * sizeof(unsigned));
    L.keys_off = off; off = a256(off + (size_t)B * sizeof(unsigned));
    L.perm_off = off; off = a256(off + (size_t)B * sizeof(int));
    L.total = off;
    return L;
}

__global__ void set_thr_kernel(float* thr, float v) { *thr = v; }

}  // namespace

// Binning for the other gate-sparse kernels (irbfn_backward_sparse's query-side pass): perm = the queries grouped by
// their set of gate intervals whose factor exceeds thr.  Workspace of sparse_bin_workspace_bytes(B, num_bounds) bytes;
// returns the device pointer of perm (inside the workspace) through perm_out.  Returns 0 or a cudaError_t.
size_t sparse_bin_workspace_bytes(long long B, int num_bounds) { return sparse_layout(B, num_bounds).total; }

int sparse_bin_queries(const float* x, const GateDev& gdev, float thr_value, long long B, int D, int sms, void* workspace,
                       const int** perm_out, cudaStream_t st, const float** edge_out) {
    const SparseLayout L = sparse_layout(B, gdev.num_bounds);
    char* ws = (char*)workspace;
    float* thr = (float*)(ws + L.thr_off);
    float* edge = (float*)(ws + L.edge_off);
    unsigned* hist = (unsigned*)(ws + L.hist_off);
    unsigned* keys = (unsigned*)(ws + L.keys_off);
    int* perm = (int*)(ws + L.perm_off);
    if (cudaMemsetAsync(ws + L.hist_off, 0, (size_t)L.nb * sizeof(unsigned), st) != cudaSuccess) return (int)cudaGetLastError();
    set_thr_kernel<<<1, 1, 0, st>>>(thr, thr_value);
    sparse_edges_kernel<<<1, 64, 0, st>>>(gdev, thr, edge);
    const size_t edge_bytes = 2 * (size_t)gdev.num_bounds * sizeof(float);
    if (L.nb <= kBlockBins && (size_t)L.nb * sizeof(unsigned) + edge_bytes <= 48 * 1024) {
        long long nblk2 = (B + 2047) / 2048;
        if (nblk2 > 4ll * sms) nblk2 = 4ll * sms;
        const long long per = (B + nblk2 - 1) / nblk2;
        sparse_key_block_kernel<<<(unsigned)nblk2, 256, (size_t)L.nb * sizeof(unsigned) + edge_bytes, st>>>(
            x, gdev, edge, keys, hist, B, D, L.key_bits, per);
        sparse_scan_kernel<<<1, 1024, 0, st>>>(hist, L.nb);
        sparse_scatter_block_kernel<<<(unsigned)nblk2, 256, (size_t)L.nb * sizeof(unsigned), st>>>(keys, hist, perm, B, L.nb, per);
    } else {
        if (edge_bytes > 48 * 1024) return (int)cudaErrorInvalidValue;
        const int nblk = (int)((B + 255) / 256 < (long long)sms * 16 ? (B + 255) / 256 : (long long)sms * 16);
        sparse_key_kernel<<<nblk, 256, edge_bytes, st>>>(x, gdev, edge, keys, hist, B, D, L.key_bits);
        sparse_scan_kernel<<<1, 1024, 0, st>>>(hist, L.nb);
        sparse_scatter_kernel<<<nblk, 256, 0, st>>>(keys, hist, perm, B);
    }
    *perm_out = perm;
    if (edge_out != nullptr) *edge_out = edge;
    return (int)cudaGetLastError();
}

}  // namespace irbfn

using namespace irbfn;

extern "C" size_t irbfn_forward_sparse_workspace_bytes(int64_t B, int num_bounds) {
    if (B <= 0 || num_bounds < 0) return 0;
    return sparse_layout(B, num_bounds).total;
}

// tol_is_gate = 0: tol is the absolute output tolerance, the gate threshold is derived from W (wabs_thr_kernel).
// tol_is_gate = 1: tol IS the gate threshold (irbfn_forward_sparse_gate: the training step, which hands the same
// threshold to irbfn_backward_sparse).
static int forward_sparse_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                               const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                               int D, int O, int basis, float tol_abs, int tol_is_gate, uint64_t* counters,
                               void* workspace, size_t workspace_bytes, irbfn_stream_t stream);

extern "C" int irbfn_forward_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                                    const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                                    int D, int O, int basis, float tol_abs, uint64_t* counters, void* workspace,
                                    size_t workspace_bytes, irbfn_stream_t stream) {
    return forward_sparse_impl(x, centers, log_sigs, W, bias, gd, y, B, R, K, D, O, basis, tol_abs, 0, counters,
                               workspace, workspace_bytes, stream);
}

extern "C" int irbfn_forward_sparse_gate(const float* x, const float* centers, const float* log_sigs, const float* W,
                                         const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R,
                                         int K, int D, int O, int basis, float gate_tol, uint64_t* counters,
                                         void* workspace, size_t workspace_bytes, irbfn_stream_t stream) {
    return forward_sparse_impl(x, centers, log_sigs, W, bias, gd, y, B, R, K, D, O, basis, gate_tol, 1, counters,
                               workspace, workspace_bytes, stream);
}

static int forward_sparse_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                               const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                               int D, int O, int basis, float tol_abs, int tol_is_gate, uint64_t* counters,
                               void* workspace, size_t workspace_bytes, irbfn_stream_t stream) {
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
    if (tol_is_gate) set_thr_kernel<<<1, 1, 0, st>>>(thr, tol_abs);
    else wabs_thr_kernel<<<1, 256, 0, st>>>(W, K, O, R, tol_abs, thr);
    const int nblk = (int)((B + 255) / 256 < (long long)sms * 16 ? (B + 255) / 256 : (long long)sms * 16);
    float* edge = (float*)(ws + L.edge_off);
    sparse_edges_kernel<<<1, 64, 0, st>>>(gdev, thr, edge);
    const size_t edge_bytes = 2 * (size_t)gdev.num_bounds * sizeof(float);
    if (L.nb <= kBlockBins && (size_t)L.nb * sizeof(unsigned) + edge_bytes <= 48 * 1024) {
        // a contiguous range of at least 2,048 queries per block, at most four blocks per SM
        long long nblk2 = (B + 2047) / 2048;
        if (nblk2 > 4ll * sms) nblk2 = 4ll * sms;
        const long long per = (B + nblk2 - 1) / nblk2;
        sparse_key_block_kernel<<<(unsigned)nblk2, 256, (size_t)L.nb * sizeof(unsigned) + edge_bytes, st>>>(
            x, gdev, edge, keys, hist, B, D, L.key_bits, per);
        sparse_scan_kernel<<<1, 1024, 0, st>>>(hist, L.nb);
        sparse_scatter_block_kernel<<<(unsigned)nblk2, 256, (size_t)L.nb * sizeof(unsigned), st>>>(keys, hist, perm, B, L.nb, per);
    } else {
        sparse_key_kernel<<<nblk, 256, edge_bytes, st>>>(x, gdev, edge, keys, hist, B, D, L.key_bits);
        sparse_scan_kernel<<<1, 1024, 0, st>>>(hist, L.nb);
        sparse_scatter_kernel<<<nblk, 256, 0, st>>>(keys, hist, perm, B);
    }
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
