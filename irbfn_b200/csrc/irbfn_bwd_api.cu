// C ABI, backward half (include/irbfn_b200.h): validation, workspace layout, launch planning,
// the split/region reduction of the parameter-gradient partials and the tanh-gate backward.
#include <cuda_runtime.h>

#include "irbfn_bwd_p.cuh"
#include "irbfn_bwd_pw.cuh"
#include "irbfn_bwd_x.cuh"
#include "irbfn_common.cuh"
#include "irbfn_plan.h"

namespace irbfn {
int launch_bwd_p_d4(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_p_d5(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_p_d6(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_p_d7(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_p_d8(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_ps_d4(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);   // gate-sparse (SPARSE = 1); -1: none
int launch_bwd_ps_d5(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_ps_d6(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_ps_d7(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_ps_d8(const BwdPArgs&, int, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_xs_d4(const BwdXArgs&, int, int, int, size_t, cudaStream_t);   // gate-sparse query-side backward
int launch_bwd_xs_d5(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_xs_d6(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_xs_d7(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_xs_d8(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
size_t sparse_bin_workspace_bytes(long long B, int num_bounds);                // irbfn_sparse.cu
int sparse_bin_queries(const float*, const GateDev&, float, long long, int, int, void*, const int**, cudaStream_t,
                       const float** edge_out = nullptr);
int launch_bwd_pw_d4(const BwdPWArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_pw_d5(const BwdPWArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_pw_d6(const BwdPWArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_pw_d7(const BwdPWArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_pw_d8(const BwdPWArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_bwd_x_d4(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_x_d5(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_x_d6(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_x_d7(const BwdXArgs&, int, int, int, size_t, cudaStream_t);
int launch_bwd_x_d8(const BwdXArgs&, int, int, int, size_t, cudaStream_t);

int device_info(int* sms, int* smem_optin);                                   // irbfn_api.cu
int check_gate(const float* gate, const IrbfnGateDesc* gd, int D, GateDev* out);  // irbfn_api.cu

// Nominal device used to size the workspace without a GPU present (B200: 148 SMs, 227 KB).
constexpr int kNominalSms = 148;
constexpr int kNominalSmem = 232448;

// ---------------------------------------------------------------------------------------
// Reduce the per-split partial sums (fixed order, double accumulation) into the gradients.
//   dC[f][j]  = 2 e_f * sum_s part[s][f][j]           e_f = exp(-2 log_sig_f)
//   dLS[f]    = 2 e_f * sum_s part[s][f][D]
//   dW[k][o]  = sum_r sum_s part[s][r*K+k][D+1+o]
//   db[o]     = sum_b part_db[b][o]                   (column sums of dy, db_partial_kernel)
// ---------------------------------------------------------------------------------------
struct ReduceArgs {
    const float* part;
    const float* part_db;
    const float* log_sigs;
    float* dC;
    float* dLS;
    float* dW;
    float* db;
    int splits, db_blocks, R, K, D, Dsrc, PW, O, o0, ocur;
};

// 32 consecutive output elements per block, one per lane, so that every read of a split's partial
// sums is a coalesced row segment; the block's 32 warps take the terms t = w, w+32, ... in order, eight
// independent loads in flight at a time (double accumulation), and warp 0 adds the 32 phase sums in
// phase order.  The order depends only on the launch geometry: deterministic.  (The first version gave
// each element to one warp with the splits over its lanes: every lane read a lone 4-byte word of a
// different split, 8x the traffic, and the R * splits terms of a dW element were a chain of dependent
// loads that set the kernel's time: 35 us at C3.)
constexpr int kReducePhases = 32;
template <typename PtrOf>
__device__ __forceinline__ double phase_sum(unsigned terms, int w, PtrOf ptr_of) {   // 32-bit term index: a 64-bit
    double s = 0.0;                                                                   // divide per term cost 25 us here
    unsigned t = (unsigned)w;
    for (; t + 7 * kReducePhases < terms; t += 8 * kReducePhases) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = *ptr_of(t + i * kReducePhases);
#pragma unroll
        for (int i = 0; i < 8; ++i) s += (double)v[i];
    }
    for (; t < terms; t += kReducePhases) s += (double)*ptr_of(t);
    return s;
}

__global__ void __launch_bounds__(32 * kReducePhases) reduce_p_kernel(const ReduceArgs a) {
    __shared__ double red[kReducePhases][32];
    const long long RK = (long long)a.R * a.K;
    const int DC = a.D + 1;                                       // dC columns (D, zero-padded) + the dLS column
    const long long nCL = (a.dC || a.dLS) ? RK * DC : 0;
    const long long nW = a.dW ? (long long)a.K * a.ocur : 0;
    const long long nB = a.db ? a.O : 0;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const long long e = (long long)blockIdx.x * 32 + lane;
    const size_t sstride = (size_t)RK * a.PW;
    double s = 0.0;
    if (e < nCL) {
        const long long f = e / DC;
        const int c = (int)(e - f * DC);
        const float* p = a.part + (size_t)f * a.PW + c;
        s = phase_sum((unsigned)a.splits, w, [&](unsigned t) { return p + (size_t)t * sstride; });
    } else if (e < nCL + nW) {
        const long long i = e - nCL;
        const int k = (int)(i / a.ocur), o = (int)(i - (long long)k * a.ocur);
        const float* p = a.part + (size_t)k * a.PW + a.D + 1 + o;
        const size_t rstride = (size_t)a.K * a.PW;
        const unsigned splits = (unsigned)a.splits;
        s = phase_sum((unsigned)a.R * splits, w, [&](unsigned t) {        // R * splits < 2^31: checked by the host
            const unsigned r = t / splits, sp = t - r * splits;
            return p + (size_t)sp * sstride + (size_t)r * rstride;
        });
    } else if (e < nCL + nW + nB) {
        const float* p = a.part_db + (int)(e - nCL - nW);
        const int O = a.O;
        s = phase_sum((unsigned)a.db_blocks, w, [&](unsigned t) { return p + (size_t)t * O; });
    }
    red[w][lane] = s;
    __syncthreads();
    if (w != 0) return;
    s = red[0][lane];
#pragma unroll
    for (int ph = 1; ph < kReducePhases; ++ph) s += red[ph][lane];
    if (e < nCL) {
        const long long f = e / DC;
        const int c = (int)(e - f * DC);
        const double sc = 2.0 * exp(-2.0 * (double)a.log_sigs[f]);
        if (c < a.Dsrc) { if (a.dC) a.dC[f * a.Dsrc + c] = (float)(sc * s); }
        else if (c == a.D) { if (a.dLS) a.dLS[f] = (float)(sc * s); }
    } else if (e < nCL + nW) {
        const long long i = e - nCL;
        const int k = (int)(i / a.ocur), o = (int)(i - (long long)k * a.ocur);
        a.dW[(size_t)k * a.O + a.o0 + o] = (float)s;
    } else if (e < nCL + nW + nB) {
        a.db[(int)(e - nCL - nW)] = (float)s;
    }
}

// Column sums of dy in two deterministic levels: block b sums a contiguous slab of rows
// (threads = row phases x columns, combined in phase order), reduce_p_kernel sums over blocks.
constexpr int kDbBlocks = 148;
__global__ void __launch_bounds__(256) db_partial_kernel(const float* __restrict__ dy,
                                                         float* __restrict__ part_db, long long B, int O) {
    __shared__ float red[256];
    const int tid = threadIdx.x;
    const long long rows = (B + gridDim.x - 1) / gridDim.x;
    const long long r0 = (long long)blockIdx.x * rows;
    const long long r1 = r0 + rows < B ? r0 + rows : B;
    for (int ob = 0; ob < O; ob += 256) {
        const int OC = O - ob < 256 ? O - ob : 256;
        const int RP = 256 / OC;
        const int o = tid % OC, ph = tid / OC;
        float acc = 0.f;
        if (ph < RP)
            for (long long r = r0 + ph; r < r1; r += RP) acc += __ldg(dy + r * O + ob + o);
        red[tid] = acc;
        __syncthreads();
        if (tid < OC) {
            float s = 0.f;
            for (int p = 0; p < RP; ++p) s += red[p * OC + tid];
            part_db[(size_t)blockIdx.x * O + ob + tid] = s;
        }
        __syncthreads();
    }
}

// gamma (B,R) for kernel P (same arithmetic as region_gate_kernel in irbfn_api.cu)
__global__ void __launch_bounds__(256) gate_fill_kernel(const float* __restrict__ x, GateDev gd,
                                                        float* __restrict__ gamma, long long B, int R, int D) {
    extern __shared__ float fac[];
    gate_rows_block(x, gd, gamma, B, R, D, fac);
}

// dx[q, j] += sum_r dgamma[q,r] * d gamma_r / d x_j      (tanh gate, model.py:70-95)
// With s(z) = 1/(1+exp(-2 delta z)):  f = s(x-lb) s(ub-x),  f' = 2 delta f (s(ub-x) - s(x-lb)).
__global__ void __launch_bounds__(256) gate_bwd_kernel(const float* __restrict__ x, GateDev gd,
                                                       const float* __restrict__ dgamma,
                                                       float* __restrict__ dx, long long B, int R, int D) {
    extern __shared__ float fac[];  // factor columns f[i][256], derivative columns fp[i][256], region index table
    const int tid = threadIdx.x;
    const long long q = (long long)blockIdx.x * blockDim.x + tid;
    const int nb = gd.num_bounds;
    float* f = fac;
    float* fp = fac + (size_t)256 * nb;
    const int rmax = R < gd.num_ranges ? R : gd.num_ranges;
    const int* ridx = gate_index_table(gd, fac + (size_t)512 * nb, rmax);
    if (q < B) {
        for (int j = 0; j < gd.S; ++j) {
            const float xv = __ldg(x + q * D + j);
            const float dl = __ldg(gd.delta + j);
            const float k2 = 2.0f * kLog2e * dl;
            for (int i = __ldg(gd.offset + j); i < __ldg(gd.offset + j + 1); ++i) {
                const float sl = gate_sigmoid(k2, xv - __ldg(gd.lower + i));
                const float su = gate_sigmoid(k2, __ldg(gd.upper + i) - xv);
                f[i * 256 + tid] = sl * su;
                fp[i * 256 + tid] = 2.0f * dl * sl * su * (su - sl);
            }
        }
    }
    __syncthreads();
    if (q >= B) return;
    float acc[IRBFN_MAX_S];
#pragma unroll
    for (int j = 0; j < IRBFN_MAX_S; ++j) acc[j] = 0.f;
    for (int r = 0; r < rmax; ++r) {
        const float dg = __ldg(dgamma + q * R + r);
        float fv[IRBFN_MAX_S], fd[IRBFN_MAX_S];
#pragma unroll
        for (int j = 0; j < IRBFN_MAX_S; ++j) {
            if (j < gd.S) {
                const int i = ridx[r * gd.S + j];
                fv[j] = f[i * 256 + tid];
                fd[j] = fp[i * 256 + tid];
            } else {
                fv[j] = 1.f;
                fd[j] = 0.f;
            }
        }
        float pre[IRBFN_MAX_S];
        float run = 1.f;
#pragma unroll
        for (int j = 0; j < IRBFN_MAX_S; ++j) { pre[j] = run; run *= fv[j]; }
        run = dg;
#pragma unroll
        for (int j = IRBFN_MAX_S - 1; j >= 0; --j) { acc[j] = fmaf(fd[j] * pre[j], run, acc[j]); run *= fv[j]; }
    }
#pragma unroll
    for (int j = 0; j < IRBFN_MAX_S; ++j)
        if (j < gd.S) dx[q * D + j] += acc[j];
}

// dh (B,K) = dy (B,O) W^T (+ dhidden): only used when O exceeds the compiled output window.
__global__ void __launch_bounds__(256) dh_kernel(const float* __restrict__ dy, const float* __restrict__ W,
                                                 const float* __restrict__ dhidden, float* __restrict__ dh,
                                                 long long B, int K, int O) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * K) return;
    const long long q = i / K;
    const int k = (int)(i - q * K);
    float acc = dhidden ? dhidden[i] : 0.f;
    for (int o = 0; o < O; ++o) acc = fmaf(__ldg(dy + q * O + o), __ldg(W + (size_t)k * O + o), acc);
    dh[i] = acc;
}

constexpr int kSparseBoundsMax = 1024;   // gate intervals the sparse query-side pass is sized for (else it runs dense)

struct BwdLayout {
    size_t gate_off, part_off, partdb_off, dgate_off, dh_off, sparse_off, total;
    BwdPPlan pplan;
    size_t max_part_floats;
    int need_p, need_x;
};

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// One function decides the workspace layout for both irbfn_backward_workspace_bytes and
// irbfn_backward.  tanh_gate: 1 when the gate comes from a descriptor.  The layout is computed
// for the nominal device when none is present so that the size query works on a CPU-only box.
static bool bwd_layout(int64_t B, int R, int K, int D, int O, unsigned mask, int tanh_gate,
                       int has_dhid, int sms, int smem, BwdLayout* L) {
    *L = BwdLayout{};
    L->need_p = (mask & IRBFN_GRAD_PARAMS) != 0;
    L->need_x = (mask & (IRBFN_GRAD_X | IRBFN_GRAD_GATE)) != 0;
    const int DT = D < 4 ? 4 : D;
    size_t off = 0;
    if (L->need_p) {
        if (tanh_gate) { L->gate_off = off; off = align256(off + (size_t)B * R * sizeof(float)); }
        // one launch per window of <= 16 output columns; size for the largest of them
        size_t max_part = 0;
        int max_splits = 0, o0 = 0;
        do {
            const int ocur = O - o0 > 16 ? 16 : O - o0;
            BwdPPlan pp;
            if (!plan_bwd_p(B, R, K, DT, D, O, fwd_ot_for(ocur), has_dhid, sms, smem, &pp)) return false;
            if (o0 == 0) L->pplan = pp;
            if (pp.part_floats > max_part) max_part = pp.part_floats;
            BwdPWPlan pw;   // the region-lane kernel may take this window instead: size for either
            if (bwdpw_applies(R, O, has_dhid) && plan_bwd_pw(B, R, K, DT, O, fwd_ot_for(ocur), sms, smem, &pw)) {
                if (pw.part_floats > max_part) max_part = pw.part_floats;
                if (pw.splits > max_splits) max_splits = pw.splits;
            }
            if (pp.splits > max_splits) max_splits = pp.splits;
            o0 += ocur;
        } while (o0 < O);
        L->max_part_floats = max_part;
        L->part_off = off;
        off = align256(off + max_part * sizeof(float));
        L->partdb_off = off;
        const int db_rows = max_splits > kDbBlocks ? max_splits : kDbBlocks;
        off = align256(off + (size_t)db_rows * (O > 0 ? O : 1) * sizeof(float));
    }
    if (L->need_x) {
        if (tanh_gate && (mask & IRBFN_GRAD_X)) { L->dgate_off = off; off = align256(off + (size_t)B * R * sizeof(float)); }
        if (O > 16) { L->dh_off = off; off = align256(off + (size_t)B * K * sizeof(float)); }
        // binning scratch of the gate-sparse query-side pass (irbfn_backward_sparse): 8 bytes per query + histogram
        if (tanh_gate && (mask & IRBFN_GRAD_X)) { L->sparse_off = off; off = align256(off + sparse_bin_workspace_bytes(B, kSparseBoundsMax)); }
    }
    L->total = off;
    return true;
}

}  // namespace irbfn

using namespace irbfn;

extern "C" {

size_t irbfn_backward_workspace_bytes(int64_t B, int R, int K, int D, int O, unsigned grad_mask) {
    if (B <= 0 || R <= 0 || K <= 0 || D <= 0 || D > IRBFN_MAX_D || O < 0) return 0;
    int sms = kNominalSms, smem = kNominalSmem;
    device_info(&sms, &smem);  // keeps the nominal values when no device is present
    BwdLayout a, b;
    // the caller may pass either gate form and may or may not pass dhidden: take the larger layout
    if (!bwd_layout(B, R, K, D, O, grad_mask, 1, 1, sms, smem, &a)) return 0;
    if (!bwd_layout(B, R, K, D, O, grad_mask, 1, 0, sms, smem, &b)) return a.total;
    return a.total > b.total ? a.total : b.total;
}

// gate_thr < 0: dense.  gate_thr >= 0: parameter-gradient terms of (query, region) pairs with gamma <= gate_thr are
// dropped where a gate-sparse instantiation exists (irbfn_backward_sparse).
static int backward_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                         const float* gate, const IrbfnGateDesc* gd, const float* dy,
                         const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                         float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                         unsigned grad_mask, float gate_thr, void* workspace, size_t workspace_bytes,
                         irbfn_stream_t stream);

int irbfn_backward(const float* x, const float* centers, const float* log_sigs, const float* W,
                   const float* gate, const IrbfnGateDesc* gd, const float* dy,
                   const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                   float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                   unsigned grad_mask, void* workspace, size_t workspace_bytes,
                   irbfn_stream_t stream) {
    return backward_impl(x, centers, log_sigs, W, gate, gd, dy, dhidden, dcenters, dlog_sigs, dW, dbias, dx, dgate, B,
                         R, K, D, O, basis, grad_mask, -1.0f, workspace, workspace_bytes, stream);
}

int irbfn_backward_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                          const float* gate, const IrbfnGateDesc* gd, const float* dy,
                          const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                          float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                          unsigned grad_mask, float gate_tol, void* workspace, size_t workspace_bytes,
                          irbfn_stream_t stream) {
    if (!(gate_tol >= 0.0f)) return IRBFN_ERR_BAD_DIMS;   // also rejects NaN
    return backward_impl(x, centers, log_sigs, W, gate, gd, dy, dhidden, dcenters, dlog_sigs, dW, dbias, dx, dgate, B,
                         R, K, D, O, basis, grad_mask, gate_tol, workspace, workspace_bytes, stream);
}

static int backward_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                         const float* gate, const IrbfnGateDesc* gd, const float* dy,
                         const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                         float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                         unsigned grad_mask, float gate_thr, void* workspace, size_t workspace_bytes,
                         irbfn_stream_t stream) {
    if (B < 0 || R <= 0 || K <= 0 || D <= 0 || O < 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (basis != IRBFN_BASIS_GAUSSIAN && basis != IRBFN_BASIS_INVERSE_QUADRATIC)
        return IRBFN_ERR_BAD_BASIS;
    const unsigned known = IRBFN_GRAD_PARAMS | IRBFN_GRAD_X | IRBFN_GRAD_GATE;
    if (grad_mask == 0 || (grad_mask & ~known)) return IRBFN_ERR_BAD_MASK;
    if ((grad_mask & IRBFN_GRAD_CENTERS) && !dcenters) return IRBFN_ERR_BAD_MASK;
    if ((grad_mask & IRBFN_GRAD_LOG_SIGS) && !dlog_sigs) return IRBFN_ERR_BAD_MASK;
    if ((grad_mask & IRBFN_GRAD_KERNEL) && O > 0 && (!dW || !dbias)) return IRBFN_ERR_BAD_MASK;
    if ((grad_mask & IRBFN_GRAD_X) && !dx) return IRBFN_ERR_BAD_MASK;
    if ((grad_mask & IRBFN_GRAD_GATE) && (!dgate || gate == nullptr)) return IRBFN_ERR_BAD_MASK;
    if (dy == nullptr && dhidden == nullptr) return IRBFN_ERR_NULL_POINTER;
    if (dy != nullptr && (O <= 0 || W == nullptr)) return IRBFN_ERR_NULL_POINTER;
    if (!centers || !log_sigs) return IRBFN_ERR_NULL_POINTER;
    GateDev gdev;
    int rc = check_gate(gate, gd, D, &gdev);
    if (rc != IRBFN_OK) return rc;
    if (B == 0) return IRBFN_OK;  // NOTE: parameter gradients of an empty batch are left untouched
    if (!x) return IRBFN_ERR_NULL_POINTER;
    const int Oeff = dy != nullptr ? O : 0;

    int sms, smem;
    rc = device_info(&sms, &smem);
    if (rc != IRBFN_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    BwdLayout L;
    if (!bwd_layout(B, R, K, D, Oeff, grad_mask, gd != nullptr, dhidden != nullptr, sms, smem, &L))
        return IRBFN_ERR_BAD_DIMS;
    if (L.total > 0 && (workspace == nullptr || workspace_bytes < L.total)) return IRBFN_ERR_WORKSPACE;
    char* ws = (char*)workspace;
    const int DT = D < 4 ? 4 : D;

    // ---------------- parameter gradients ----------------
    if (L.need_p) {
        const float* gate_p = gate;
        if (gd != nullptr) {
            float* gbuf = (float*)(ws + L.gate_off);
            const size_t gsm = gate_rows_smem_bytes(gdev.num_bounds, 1, gdev.num_ranges, gdev.S);
            if (gsm > (size_t)smem) return IRBFN_ERR_BAD_GATE;
            if (gsm > 48 * 1024 &&
                cudaFuncSetAttribute(gate_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm) != cudaSuccess)
                return IRBFN_ERR_LAUNCH;
            gate_fill_kernel<<<(unsigned)((B + 255) / 256), 256, gsm, st>>>(x, gdev, gbuf, B, R, D);
            { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("gate_fill_kernel", (int)ce); }
            gate_p = gbuf;
        }
        int o0 = 0;
        bool first = true;
        do {
            const int ocur = Oeff - o0 > 16 ? 16 : Oeff - o0;
            const int OT = fwd_ot_for(ocur);
            BwdPPlan pp;
            BwdPWPlan pw;
            // the bias gradient's column sums of dy ride along in the chunk-0 CTAs (one thread per column and
            // row phase); outputs wider than a CTA has threads fall back to the separate pass
            const bool in_kernel_db = (grad_mask & IRBFN_GRAD_KERNEL) && Oeff > 0 && Oeff <= 128;
            // gate-sparse calls take the thread-private kernel (it skips per (warp, query); the region-lane kernel
            // cannot: its lanes ARE the regions) whenever that kernel forms dh in the thread
            const bool sparse = gate_thr >= 0.0f && R > 1;
            const bool sparse_p = sparse && plan_bwd_p(B, R, K, DT, D, Oeff, OT, dhidden != nullptr, sms, smem, &pp, 1) &&
                                  pp.dht && pp.part_floats <= L.max_part_floats;
            const bool wide = !sparse_p && bwdpw_applies(R, Oeff, dhidden != nullptr) &&
                              plan_bwd_pw(B, R, K, DT, Oeff, OT, sms, smem, &pw);
            if (wide) {
                // several regions share a wide readout: lanes = regions, dh from a per-tile table, dW from the
                // region-summed activation (irbfn_bwd_pw.cuh)
                if (pw.part_floats > L.max_part_floats) return IRBFN_ERR_WORKSPACE;
                if ((long long)R * pw.splits >= (1ll << 31)) return IRBFN_ERR_BAD_DIMS;
                BwdPWArgs w{};
                w.x = x; w.centers = centers; w.log_sigs = log_sigs; w.W = W; w.gate = gate_p; w.dy = dy;
                w.part = (float*)(ws + L.part_off);
                w.part_db = (first && in_kernel_db) ? (float*)(ws + L.partdb_off) : nullptr;
                w.B = B; w.R = R; w.K = K; w.Dsrc = D; w.O = Oeff;
                w.G = pw.G; w.NRG = pw.NRG; w.units = pw.units; w.units_per_cta = pw.units_per_cta;
                w.nchunks = pw.nchunks; w.splits = pw.splits; w.TQ = pw.TQ; w.Kspan = pw.Kspan;
                const int grid = pw.nchunks * pw.splits;
                int e;
                switch (DT) {
                    case 4: e = launch_bwd_pw_d4(w, OT, basis, grid, pw.NT, pw.smem_bytes, st); break;
                    case 5: e = launch_bwd_pw_d5(w, OT, basis, grid, pw.NT, pw.smem_bytes, st); break;
                    case 6: e = launch_bwd_pw_d6(w, OT, basis, grid, pw.NT, pw.smem_bytes, st); break;
                    case 7: e = launch_bwd_pw_d7(w, OT, basis, grid, pw.NT, pw.smem_bytes, st); break;
                    default: e = launch_bwd_pw_d8(w, OT, basis, grid, pw.NT, pw.smem_bytes, st); break;
                }
                if (e != 0) return note_cuda_error("irbfn_bwd_pw_kernel", e);
                pp = BwdPPlan{};
                pp.splits = pw.splits; pp.PW = pw.PW;
            } else {
            if (!sparse_p && !plan_bwd_p(B, R, K, DT, D, Oeff, OT, dhidden != nullptr, sms, smem, &pp)) return IRBFN_ERR_BAD_DIMS;
            if (pp.part_floats > L.max_part_floats) return IRBFN_ERR_WORKSPACE;
            if ((long long)R * pp.splits >= (1ll << 31)) return IRBFN_ERR_BAD_DIMS;   // reduce_p_kernel's term index
            BwdPArgs a{};
            a.x = x; a.centers = centers; a.log_sigs = log_sigs; a.W = W; a.gate = gate_p;
            a.dy = dy; a.dhidden = dhidden;
            a.part = (float*)(ws + L.part_off);
            a.part_db = (first && in_kernel_db) ? (float*)(ws + L.partdb_off) : nullptr;
            a.B = B; a.R = R; a.K = K; a.Dsrc = D; a.O = Oeff; a.o0 = o0; a.ocur = ocur;
            a.nchunks = pp.nchunks; a.splits = pp.splits; a.pairs_per_chunk = pp.pairs_per_chunk; a.ppw = pp.ppw;
            a.TQ = pp.TQ; a.Kspan = pp.Kspan; a.Rspan = pp.Rspan;
            a.write_cls = first ? 1 : 0;
            a.gate_thr = gate_thr;
            const int grid = pp.nchunks * pp.splits;
            int e = -1;
            if (sparse) {   // gate-sparse instantiation, where one exists
                switch (DT) {
                    case 4: e = launch_bwd_ps_d4(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                    case 5: e = launch_bwd_ps_d5(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                    case 6: e = launch_bwd_ps_d6(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                    case 7: e = launch_bwd_ps_d7(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                    default: e = launch_bwd_ps_d8(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                }
            }
            if (e < 0)
            switch (DT) {
                case 4: e = launch_bwd_p_d4(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                case 5: e = launch_bwd_p_d5(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                case 6: e = launch_bwd_p_d6(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                case 7: e = launch_bwd_p_d7(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
                default: e = launch_bwd_p_d8(a, OT, basis, pp.dht, grid, pp.NT, pp.smem_bytes, st); break;
            }
            if (e != 0) return note_cuda_error("irbfn_bwd_p_kernel", e);
            }
            ReduceArgs r{};
            r.part = (float*)(ws + L.part_off); r.part_db = (float*)(ws + L.partdb_off); r.log_sigs = log_sigs;
            r.dC = (first && (grad_mask & IRBFN_GRAD_CENTERS)) ? dcenters : nullptr;
            r.dLS = (first && (grad_mask & IRBFN_GRAD_LOG_SIGS)) ? dlog_sigs : nullptr;
            r.dW = ((grad_mask & IRBFN_GRAD_KERNEL) && ocur > 0) ? dW : nullptr;
            r.db = (first && (grad_mask & IRBFN_GRAD_KERNEL) && Oeff > 0) ? dbias : nullptr;
            r.splits = pp.splits; r.db_blocks = in_kernel_db ? pp.splits : kDbBlocks; r.R = R; r.K = K; r.D = DT; r.Dsrc = D; r.PW = pp.PW;
            r.O = Oeff; r.o0 = o0; r.ocur = ocur;
            const long long n = ((r.dC || r.dLS) ? (long long)R * K * (DT + 1) : 0) +
                                (r.dW ? (long long)K * ocur : 0) + (r.db ? Oeff : 0);
            if (r.db != nullptr && !in_kernel_db) {
                db_partial_kernel<<<kDbBlocks, 256, 0, st>>>(dy, (float*)(ws + L.partdb_off), B, Oeff);
                { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("db_partial_kernel", (int)ce); }
            }
            if (n > 0) {
                reduce_p_kernel<<<(unsigned)((n + 31) / 32), 32 * kReducePhases, 0, st>>>(r);   // 32 elements per block
                { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("reduce_p_kernel", (int)ce); }
            }
            o0 += ocur;
            first = false;
        } while (o0 < Oeff);
    }

    // ---------------- query-side gradients ----------------
    if (L.need_x) {
        const float* dy_x = dy;
        const float* dhid_x = dhidden;
        int Ox = Oeff;
        if (Oeff > 16) {  // wider than the compiled window: materialise dh once
            float* dh = (float*)(ws + L.dh_off);
            dh_kernel<<<(unsigned)(((long long)B * K + 255) / 256), 256, 0, st>>>(dy, W, dhidden, dh, B, K, Oeff);
            { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("dh_kernel", (int)ce); }
            dy_x = nullptr;
            dhid_x = dh;
            Ox = 0;
        }
        const int OT = fwd_ot_for(Ox);
        FwdPlan plan;
        if (!plan_bwd_x(R, K, DT, OT, smem, &plan)) return IRBFN_ERR_BAD_DIMS;
        BwdXArgs a{};
        a.x = x; a.centers = centers; a.log_sigs = log_sigs; a.W = W; a.gate_ext = gate; a.gd = gdev;
        a.dy = dy_x; a.dhidden = dhid_x;
        a.dx = (grad_mask & IRBFN_GRAD_X) ? dx : nullptr;
        // tanh gate with few split dims: the kernel folds d gamma into dx itself; otherwise d gamma goes
        // to the dgate output (external gate) or to scratch for gate_bwd_kernel
        const bool fuse = gd != nullptr && (grad_mask & IRBFN_GRAD_X) && gdev.S <= BwdXShape::MAX_FUSED_S;
        float* dgbuf = nullptr;
        if (gd != nullptr && (grad_mask & IRBFN_GRAD_X) && !fuse) dgbuf = (float*)(ws + L.dgate_off);
        if (grad_mask & IRBFN_GRAD_GATE) dgbuf = dgate;
        a.dgate = dgbuf;
        a.fuse_gate = fuse ? 1 : 0;
        a.B = B; a.R = R; a.K = K; a.Dsrc = D; a.O = Ox;
        a.RC = plan.RC; a.KC = plan.KC; a.stage_once = plan.stage_once;
        const long long tiles = (B + BwdXShape::TQ - 1) / BwdXShape::TQ;
        a.num_tiles = (int)tiles;
        const int grid = (int)(tiles < sms ? tiles : sms);
        int e = -1;
        if (gate_thr >= 0.0f && gd != nullptr && (grad_mask & IRBFN_GRAD_X) && R > 1 && plan.RC <= 64 &&
            gdev.num_bounds <= kSparseBoundsMax && B <= 0x7fffffffll) {
            // gate-sparse: bin the queries by their set of non-negligible gate intervals, then the SPARSE instantiation
            const int* perm = nullptr;
            const int be = sparse_bin_queries(x, gdev, gate_thr, B, D, sms, ws + L.sparse_off, &perm, st);
            if (be != 0) return note_cuda_error("sparse binning (backward)", be);
            a.perm = perm;
            a.gate_thr = gate_thr;
            switch (DT) {
                case 4: e = launch_bwd_xs_d4(a, OT, basis, grid, plan.smem_bytes, st); break;
                case 5: e = launch_bwd_xs_d5(a, OT, basis, grid, plan.smem_bytes, st); break;
                case 6: e = launch_bwd_xs_d6(a, OT, basis, grid, plan.smem_bytes, st); break;
                case 7: e = launch_bwd_xs_d7(a, OT, basis, grid, plan.smem_bytes, st); break;
                default: e = launch_bwd_xs_d8(a, OT, basis, grid, plan.smem_bytes, st); break;
            }
        } else
        switch (DT) {
            case 4: e = launch_bwd_x_d4(a, OT, basis, grid, plan.smem_bytes, st); break;
            case 5: e = launch_bwd_x_d5(a, OT, basis, grid, plan.smem_bytes, st); break;
            case 6: e = launch_bwd_x_d6(a, OT, basis, grid, plan.smem_bytes, st); break;
            case 7: e = launch_bwd_x_d7(a, OT, basis, grid, plan.smem_bytes, st); break;
            default: e = launch_bwd_x_d8(a, OT, basis, grid, plan.smem_bytes, st); break;
        }
        if (e != 0) return note_cuda_error("irbfn_bwd_x_kernel", e);
        if (gd != nullptr && (grad_mask & IRBFN_GRAD_X) && !fuse) {
            const size_t gsm = gate_rows_smem_bytes(gdev.num_bounds, 2, gdev.num_ranges, gdev.S);
            if (gsm > (size_t)smem) return IRBFN_ERR_BAD_GATE;
            if (gsm > 48 * 1024 &&
                cudaFuncSetAttribute(gate_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm) != cudaSuccess)
                return IRBFN_ERR_LAUNCH;
            gate_bwd_kernel<<<(unsigned)((B + 255) / 256), 256, gsm, st>>>(x, gdev, dgbuf, dx, B, R, D);
            { const cudaError_t ce = cudaGetLastError(); if (ce != cudaSuccess) return note_cuda_error("gate_bwd_kernel", (int)ce); }
        }
    }
    return IRBFN_OK;
}

}  // extern "C"
