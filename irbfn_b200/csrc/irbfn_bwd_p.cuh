// IRBFN backward, parameter gradients (sm_100a): dcenters, dlog_sigs, dkernel, dbias.
//
// Reference: jax.value_and_grad(loss_fn)(params) through WCRBFNet.apply
//   scripts/train_nmpc.py:297,416 ; scripts/train_nmpc_frenet.py:388,416,448
//   forward being differentiated: model.py:187-196, flax_rbf.py:275-283
//
// Design ("backward kernel P"): the roles of the forward are swapped.
//   * parameter-stationary: a thread owns NKP packed PAIRS of kernels (f, f+1) of the flat
//     (region,kernel) index; their centres, width scales and all of their gradient accumulators
//     (D + 1 + OT packed floats per pair) live in registers for the whole launch, so the
//     cross-query reduction of dC / dlog_sig / dW costs nothing per query: no shuffles, no atomics.
//   * queries stream through shared memory in tiles of TQ: x duplicated as (x_j, x_j) so a
//     broadcast LDS.128 yields two ready packed operands; dy duplicated the same way; the gate
//     slab gamma[q, r_lo..r_hi] and dh[q,k] = dy[q,:] W[k,:]^T (+ dhidden) are computed once per
//     tile.  The next tile's raw rows arrive by cp.async while the current one is consumed.
//   * phi is recomputed from x and the parameters (nothing saved by the forward).
//   * grid = splits x chunks: a chunk is a contiguous range of kernel pairs, a split a strided
//     set of query tiles.  Every CTA writes its raw partial sums to the workspace once; a second
//     small kernel reduces over splits (and over regions for dW) in a fixed order, in double, and
//     applies the 2 exp(-2 log_sig) factor.  Fixed partition + fixed order = deterministic.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct BwdPArgs {
    const float* x;         // (B,Dsrc)
    const float* centers;   // (R,K,Dsrc)
    const float* log_sigs;  // (R,K)
    const float* W;         // (K,O) or nullptr (O == 0)
    const float* gate;      // (B,R)  -- always materialised for this kernel
    const float* dy;        // (B,O) or nullptr
    const float* dhidden;   // (B,K) or nullptr
    float* part;            // [splits][R*K][PW]   PW = D + 1 + OT
    long long B;
    int R, K, Dsrc, O;
    int o0, ocur;           // dW column window of this launch
    int nchunks, splits;
    int pairs_per_chunk;    // Pc
    int TQ;                 // queries per tile
    int Kspan;              // columns of the dh table (even)
    int Rspan;              // columns of the gate table, + 1 zero column at index Rspan
    int write_cls;          // 1: write dC / dLS partials (first output window only)
};

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
    unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int D, int OT, int BASIS, int NKP>
__global__ void __launch_bounds__(256, 1) irbfn_bwd_p_kernel(const BwdPArgs a) {
    constexpr int OP = OT / 2;                  // dy columns come duplicated: (dy_o, dy_o)
    constexpr int OTS = OT > 0 ? OT : 1;
    constexpr int PW = D + 1 + OT;
    constexpr int XS = ((2 * D + 3) / 4) * 4;   // row stride of xs: keeps LDS.128 aligned for odd D
    const int NT = blockDim.x;
    const int tid = threadIdx.x;
    const int chunk = blockIdx.x % a.nchunks;
    const int split = blockIdx.x / a.nchunks;
    const int RK = a.R * a.K;
    const int TQ = a.TQ, O = a.O, Dsrc = a.Dsrc, Kspan = a.Kspan, Rspan = a.Rspan, K = a.K;
    const int GS = Rspan + 1;                   // gate table row stride (last column is zero)

    const int f_lo = 2 * chunk * a.pairs_per_chunk;
    const int f_len = 2 * a.pairs_per_chunk;
    const int f_hi = min(f_lo + f_len, RK);     // exclusive
    const int r_lo = f_lo / K;
    const bool k_direct = f_len >= K;           // dh table indexed by k itself (else by f - f_lo)
    const int raw_dh = a.dhidden != nullptr ? Kspan : 0;

    extern __shared__ __align__(16) float smem[];
    const int raw_stride = TQ * (Dsrc + O + Rspan + raw_dh);
    float* raw0 = smem;                                   // 2 raw stages
    float* xs = smem + 2 * (size_t)raw_stride;            // [TQ][XS]  (x_j, x_j) pairs
    xs += ((4 - ((xs - smem) & 3)) & 3);
    float* dys = xs + (size_t)TQ * XS;                    // [TQ][2*OT]
    float* dhs = dys + (size_t)TQ * 2 * OTS;              // [TQ][Kspan]
    float* gs = dhs + (size_t)TQ * Kspan;                 // [TQ][GS]

    // ---- this thread's kernel pairs: parameters -> registers ----
    float2 nc2[NKP][D];   // (-c_j[f0], -c_j[f1])
    float2 sc2[NKP];      // width scale of (f0, f1)
    int kidx[NKP][2], ridx[NKP][2];
    bool contig[NKP];
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        const int pl = tid + p * NT;  // pair slot within the chunk
        sc2[p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int f = f_lo + 2 * pl + h;
            const bool valid = pl < a.pairs_per_chunk && f < f_hi;
            const int fc = valid ? f : f_lo;
            const int r = fc / K, k = fc - r * K;
            kidx[p][h] = k_direct ? k : (fc - f_lo);
            ridx[p][h] = valid ? (r - r_lo) : Rspan;  // zero column for padding kernels
            float s = valid ? width_scale<BASIS>(__ldg(a.log_sigs + fc)) : 0.f;
            if (h == 0) sc2[p].x = s; else sc2[p].y = s;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float c = (valid && j < Dsrc) ? -__ldg(a.centers + (size_t)fc * Dsrc + j) : 0.f;
                if (h == 0) nc2[p][j].x = c; else nc2[p][j].y = c;
            }
        }
        contig[p] = (kidx[p][1] == kidx[p][0] + 1) && ((kidx[p][0] & 1) == 0);
    }

    float2 aC[NKP][D], aL[NKP], aW[NKP][OP > 0 ? OP * 2 : 1];
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        aL[p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < D; ++j) aC[p][j] = make_float2(0.f, 0.f);
#pragma unroll
        for (int o = 0; o < OT; ++o) aW[p][o] = make_float2(0.f, 0.f);
    }

    const long long ntiles = (a.B + TQ - 1) / TQ;

    auto issue_tile = [&](long long tile, int stage) {
        float* raw = raw0 + (size_t)stage * raw_stride;
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        float* rx = raw;
        float* rdy = rx + (size_t)TQ * Dsrc;
        float* rg = rdy + (size_t)TQ * O;
        float* rdh = rg + (size_t)TQ * Rspan;
        for (int i = tid; i < tq * Dsrc; i += NT) cp_async4(rx + i, a.x + q0 * Dsrc + i);
        if (a.dy != nullptr)
            for (int i = tid; i < tq * O; i += NT) cp_async4(rdy + i, a.dy + q0 * O + i);
        for (int i = tid; i < tq * Rspan; i += NT) {
            const int q = i / Rspan, rr = i - q * Rspan;
            const int r = min(r_lo + rr, a.R - 1);
            cp_async4(rg + i, a.gate + (q0 + q) * a.R + r);
        }
        if (raw_dh)
            for (int i = tid; i < tq * Kspan; i += NT) {
                const int q = i / Kspan, kk = i - q * Kspan;
                int k = k_direct ? kk : (f_lo + kk) % K;
                k = min(k, K - 1);
                cp_async4(rdh + i, a.dhidden + (q0 + q) * K + k);
            }
        cp_async_commit();
    };

    long long tile = split;
    int stage = 0;
    if (tile < ntiles) issue_tile(tile, 0);

    for (; tile < ntiles; tile += a.splits, stage ^= 1) {
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        cp_async_wait_all();
        __syncthreads();  // raw[stage] landed; everyone is done computing on the previous tile

        // ---- transform raw rows into the compute layout ----
        {
            const float* raw = raw0 + (size_t)stage * raw_stride;
            const float* rx = raw;
            const float* rdy = rx + (size_t)TQ * Dsrc;
            const float* rg = rdy + (size_t)TQ * O;
            const float* rdh = rg + (size_t)TQ * Rspan;
            for (int i = tid; i < tq * D; i += NT) {
                const int q = i / D, j = i - q * D;
                const float v = j < Dsrc ? rx[q * Dsrc + j] : 0.f;
                *reinterpret_cast<float2*>(xs + (size_t)q * XS + 2 * j) = make_float2(v, v);
            }
            if (OT > 0)
                for (int i = tid; i < tq * OT; i += NT) {
                    const int q = i / OT, o = i - q * OT;
                    const float v = o < a.ocur ? rdy[q * O + a.o0 + o] : 0.f;
                    reinterpret_cast<float2*>(dys)[q * OT + o] = make_float2(v, v);
                }
            for (int i = tid; i < tq * GS; i += NT) {
                const int q = i / GS, rr = i - q * GS;
                gs[i] = rr < Rspan ? rg[q * Rspan + rr] : 0.f;
            }
            for (int i = tid; i < tq * Kspan; i += NT) {
                const int q = i / Kspan, kk = i - q * Kspan;
                int k = k_direct ? kk : (f_lo + kk) % K;
                float acc = raw_dh ? rdh[i] : 0.f;
                if (k < K && a.dy != nullptr)
                    for (int o = 0; o < O; ++o) acc = fmaf(rdy[q * O + o], __ldg(a.W + (size_t)k * O + o), acc);
                dhs[i] = acc;
            }
        }
        __syncthreads();
        if (tile + a.splits < ntiles) issue_tile(tile + a.splits, stage ^ 1);

        // ---- hot loop over the queries of the tile ----
        for (int q = 0; q < tq; ++q) {
            float2 x2[D];
            {
                const float4* xp = reinterpret_cast<const float4*>(xs + (size_t)q * XS);
#pragma unroll
                for (int i = 0; i < D / 2; ++i) {
                    const float4 t = xp[i];
                    x2[2 * i] = make_float2(t.x, t.y);
                    x2[2 * i + 1] = make_float2(t.z, t.w);
                }
                if (D & 1) x2[D - 1] = reinterpret_cast<const float2*>(xs + (size_t)q * XS)[D - 1];
            }
            float2 dy2[OTS];
            if (OT > 0) {
                const float2* dp = reinterpret_cast<const float2*>(dys + (size_t)q * 2 * OT);
#pragma unroll
                for (int o = 0; o < OT; ++o) dy2[o] = dp[o];
            }
            const float* gq = gs + (size_t)q * GS;
            const float* dq = dhs + (size_t)q * Kspan;
#pragma unroll
            for (int p = 0; p < NKP; ++p) {
                const float2 g2 = make_float2(gq[ridx[p][0]], gq[ridx[p][1]]);
                float2 dh2;
                if (contig[p]) dh2 = *reinterpret_cast<const float2*>(dq + kidx[p][0]);
                else dh2 = make_float2(dq[kidx[p][0]], dq[kidx[p][1]]);
                float2 df[D];
                df[0] = add2(x2[0], nc2[p][0]);
                float2 d2 = mul2(df[0], df[0]);
#pragma unroll
                for (int j = 1; j < D; ++j) {
                    df[j] = add2(x2[j], nc2[p][j]);
                    d2 = fma2(df[j], df[j], d2);
                }
                float2 phi;
                if (BASIS == IRBFN_BASIS_GAUSSIAN) {
                    const float2 e = mul2(d2, sc2[p]);
                    phi = make_float2(ex2_approx(e.x), ex2_approx(e.y));
                } else {
                    const float2 e = fma2(d2, sc2[p], make_float2(1.f, 1.f));
                    phi = make_float2(rcp_approx(e.x), rcp_approx(e.y));
                }
                const float2 v = mul2(g2, phi);  // gamma * phi
#pragma unroll
                for (int o = 0; o < OT; ++o) aW[p][o] = fma2(v, dy2[o], aW[p][o]);
                float2 t = mul2(v, dh2);         // dh * gamma * phi   (= -dL/d(alpha^2), gaussian)
                if (BASIS == IRBFN_BASIS_INVERSE_QUADRATIC) t = mul2(t, phi);
#pragma unroll
                for (int j = 0; j < D; ++j) aC[p][j] = fma2(t, df[j], aC[p][j]);
                aL[p] = fma2(t, d2, aL[p]);
            }
        }
    }
    cp_async_wait_all();

    // ---- flush raw partial sums (the reduce kernel applies 2 exp(-2 log_sig)) ----
    float* part = a.part + (size_t)split * RK * PW;
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        const int pl = tid + p * NT;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int f = f_lo + 2 * pl + h;
            if (pl < a.pairs_per_chunk && f < f_hi) {
                float* o = part + (size_t)f * PW;
                if (a.write_cls) {
#pragma unroll
                    for (int j = 0; j < D; ++j) o[j] = h == 0 ? aC[p][j].x : aC[p][j].y;
                    o[D] = h == 0 ? aL[p].x : aL[p].y;
                }
#pragma unroll
                for (int oo = 0; oo < OT; ++oo) o[D + 1 + oo] = h == 0 ? aW[p][oo].x : aW[p][oo].y;
            }
        }
    }
}

template <int D, int OT, int NKP>
inline int launch_bwd_p_nkp(const BwdPArgs& a, int basis, int grid, int nt, size_t smem, cudaStream_t st) {
    cudaError_t e;
    if (basis == IRBFN_BASIS_GAUSSIAN) {
        auto kern = irbfn_bwd_p_kernel<D, OT, IRBFN_BASIS_GAUSSIAN, NKP>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, nt, smem, st>>>(a);
    } else {
        auto kern = irbfn_bwd_p_kernel<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, NKP>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, nt, smem, st>>>(a);
    }
    return (int)cudaGetLastError();
}

// kernel pairs per thread as a function of the output window: keeps the accumulators in registers
__host__ __device__ constexpr int bwd_p_nkp_for(int OT) { return OT <= 4 ? 3 : (OT <= 10 ? 2 : 1); }

template <int D>
inline int launch_bwd_p(const BwdPArgs& a, int OT, int basis, int grid, int nt, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return launch_bwd_p_nkp<D, 0, bwd_p_nkp_for(0)>(a, basis, grid, nt, smem, st);
        case 2: return launch_bwd_p_nkp<D, 2, bwd_p_nkp_for(2)>(a, basis, grid, nt, smem, st);
        case 4: return launch_bwd_p_nkp<D, 4, bwd_p_nkp_for(4)>(a, basis, grid, nt, smem, st);
        case 8: return launch_bwd_p_nkp<D, 8, bwd_p_nkp_for(8)>(a, basis, grid, nt, smem, st);
        case 10: return launch_bwd_p_nkp<D, 10, bwd_p_nkp_for(10)>(a, basis, grid, nt, smem, st);
        case 16: return launch_bwd_p_nkp<D, 16, bwd_p_nkp_for(16)>(a, basis, grid, nt, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
