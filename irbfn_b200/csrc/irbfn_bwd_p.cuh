// IRBFN backward, parameter gradients (sm_100a): dcenters, dlog_sigs, dkernel, dbias.
//
// Reference: jax.value_and_grad(loss_fn)(params) through WCRBFNet.apply
//   scripts/train_nmpc.py:297,416 ; scripts/train_nmpc_frenet.py:388,416,448
//   forward being differentiated: model.py:187-196, flax_rbf.py:275-283
//
// Design: the roles of the forward are swapped.
//   * parameter-stationary: a thread owns NKP packed PAIRS of kernels (f, f+1) of the flat
//     (region,kernel) index; their centres, width scales, readout rows and all of their gradient
//     accumulators (D + 1 + OT packed floats per pair) live in registers for the whole launch, so
//     the cross-query reduction of dC / dlog_sig / dW costs nothing per query: no shuffles, no atomics.
//   * queries stream through shared memory in tiles of TQ rows [x | dy | gamma slab], written
//     straight into their compute layout by cp.async (double buffered, one barrier per tile).  A
//     query's scalars are broadcast operands of the packed instructions (SASS `Rn.F32`), so nothing
//     is duplicated in shared memory.
//   * dh_k = dy . W[k,:] is formed in the thread from its own readout rows (DHT = 1) when the output
//     is narrow (O <= 4, the NMPC shapes) or when the CTA's kernel chunk lies inside one region, so
//     that nothing could share a table (K large: C2b, the C5 stress shape); otherwise (several
//     regions over the same kernels, or a dhidden cotangent) a per-tile table dh[q][k] is built once
//     per tile and shared by all regions of the chunk.
//   * phi is recomputed from x and the parameters (nothing saved by the forward).
//   * grid = splits x chunks: a chunk is a contiguous range of kernel pairs, a split a strided
//     set of query tiles.  Every CTA writes its raw partial sums to the workspace once; a second
//     small kernel reduces over splits (and over regions for dW) in a fixed order, in double, and
//     applies the 2 exp(-2 log_sig) factor.  Fixed partition + fixed order = deterministic.
#pragma once
#include "irbfn_common.cuh"
#include "irbfn_plan.h"

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
    float* part_db;         // [splits][O] column sums of dy (the bias gradient's partials), or nullptr
    long long B;
    int R, K, Dsrc, O;
    int o0, ocur;           // dW column window of this launch
    int nchunks, splits;
    int pairs_per_chunk;    // Pc
    int ppw;                // kernel pairs per warp: lanes >= ppw hold none (32: pair slot = thread index)
    int TQ;                 // queries per tile
    int Kspan;              // columns of the dh table (even)
    int Rspan;              // columns of the gate table, + 1 zero column at index Rspan
    int write_cls;          // 1: write dC / dLS partials (first output window only)
    // ---- read by the SPARSE instantiations only; behind the dense fields so their parameter offsets stay put ----
    float gate_thr;         // a (query, region) pair with gamma <= gate_thr is dropped (irbfn_backward_sparse)
};

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
    unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Bias gradient db[o] = sum_q dy[q][o], taken by the CTAs of chunk 0 (one per split) while a tile's dy rows
// sit in shared memory: thread (phase, column) adds every nph-th row of its column into a shared-memory
// accumulator; the phases are combined in phase order at the end.  Replaces a separate pass over dy
// (db_partial_kernel).  Everything is recomputed from the kernel arguments at each use and the running sums
// live in shared memory, so the hot loop keeps its registers (the 640-thread variant sits at its 96-register
// cap: holding this state in registers spilled and cost 12 % of the kernel).
constexpr int kDbScratch = 1024;   // floats reserved at the end of dynamic shared memory (>= threads per CTA)
__device__ __forceinline__ bool db_active(const float* part_db, const float* dy, int chunk, int O, int NT) {
    return part_db != nullptr && dy != nullptr && chunk == 0 && O > 0 && O <= NT;
}
__device__ __forceinline__ void db_tile(float* dbs, const float* dys, int DS, int tq, int O, int NT, int tid) {
    int nph = NT / O;
    if (nph > 32) nph = 32;
    if (tid < nph * O) {
        const int o = tid % O, ph = tid / O;
        float s = 0.f;
        for (int q = ph; q < tq; q += nph) s += dys[(size_t)q * DS + o];
        dbs[tid] += s;
    }
}
__device__ __forceinline__ void db_flush(const float* dbs, float* part_db, int split, int O, int NT, int tid) {
    __syncthreads();
    int nph = NT / O;
    if (nph > 32) nph = 32;
    if (tid < O) {
        float s = 0.f;
        for (int p = 0; p < nph; ++p) s += dbs[p * O + tid];
        part_db[(size_t)split * O + tid] = s;
    }
}

// shared-memory row strides (floats); must match bwdp_smem_floats in irbfn_plan.h
__host__ __device__ constexpr int bwdp_xs(int D) { return ((D + 3) / 4) * 4; }
__host__ __device__ inline int bwdp_ds(int O) { return O <= 0 ? 0 : (O <= 2 ? 2 : ((O + 3) / 4) * 4); }

// SPARSE = 1 (gate-sparse parameter gradients, irbfn_backward_sparse; NKP == 1 only): the threads of a warp hold
// kernel pairs of at most two neighbouring regions and a query is the broadcast operand of the whole CTA, so
// "this query's gate is <= gate_thr for every region this warp holds" is a WARP-UNIFORM predicate: the warp ballots
// it for 32 queries at a time and evaluates only the set bits.  No binning or sorting of the queries is needed (unlike the
// query-stationary forward, irbfn_sparse.cu), and the partial sums keep their fixed order: still deterministic.
template <int D, int OT, int BASIS, int NKP, int DHT, int SPARSE = 0>
__global__ void __launch_bounds__((NKP == 1 && OT <= 4) ? 640 : 384, 1) irbfn_bwd_p_kernel(const BwdPArgs a) {
    static_assert(SPARSE == 0 || NKP == 1, "the warp-uniform skip test reads the thread's single kernel pair");
    constexpr int OTS = OT > 0 ? OT : 1;
    constexpr int PW = D + 1 + OT;
    constexpr int XS = bwdp_xs(D);
    const int NT = blockDim.x;
    const int tid = threadIdx.x;
    const bool lane_ok = (tid & 31) < a.ppw;
    const int pslot = (tid >> 5) * a.ppw + (tid & 31);   // == tid when ppw == 32
    const int chunk = blockIdx.x % a.nchunks;
    const int split = blockIdx.x / a.nchunks;
    const int RK = a.R * a.K;
    const int TQ = a.TQ, O = a.O, Dsrc = a.Dsrc, Kspan = a.Kspan, Rspan = a.Rspan, K = a.K;
    const int GS = Rspan + 1;                   // gate row stride (last column stays zero)
    const int DS = bwdp_ds(O);                  // dy row stride (pad columns stay zero)

    const int f_lo = 2 * chunk * a.pairs_per_chunk;
    const int f_len = 2 * a.pairs_per_chunk;
    const int f_hi = min(f_lo + f_len, RK);     // exclusive
    const int r_lo = f_lo / K;
    const bool k_direct = f_len >= K;           // dh table indexed by k itself (else by f - f_lo)
    const bool has_dhid = a.dhidden != nullptr;

    extern __shared__ __align__(16) float smem[];
    // two stages of [xs | dys | gs | raw dhidden], then the dh table (DHT == 0 only)
    const int st_x = 0, st_dy = st_x + TQ * XS, st_g = st_dy + TQ * DS;
    const int st_dh = ((st_g + TQ * GS + 3) / 4) * 4;
    const int stage_floats = ((st_dh + (has_dhid ? TQ * Kspan : 0) + 3) / 4) * 4;
    float* dhs = smem + 2 * (size_t)stage_floats;  // [TQ][Kspan]

    // zero both stages once: cp.async never writes the pad columns, they must read as 0
    for (int i = tid; i < 2 * stage_floats; i += NT) smem[i] = 0.f;

    // ---- this thread's kernel pairs: parameters -> registers ----
    float2 nc2[NKP][D];   // (-c_j[f0], -c_j[f1])
    float2 sc2[NKP];      // width scale of (f0, f1)
    float2 w2[NKP][DHT ? OTS : 1];  // readout rows (W[k0,o], W[k1,o])
    int kidx[NKP][2], ridx[NKP][2];
    bool contig[NKP];
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        const int pl = pslot + p * NT;  // pair slot within the chunk
        sc2[p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int f = f_lo + 2 * pl + h;
            const bool valid = lane_ok && pl < a.pairs_per_chunk && f < f_hi;
            const int fc = valid ? f : f_lo;
            const int r = fc / K, k = fc - r * K;
            kidx[p][h] = k_direct ? k : (fc - f_lo);
            ridx[p][h] = valid ? (r - r_lo) : Rspan;  // zero column for padding kernels
            const float s = valid ? width_scale<BASIS>(__ldg(a.log_sigs + fc)) : 0.f;
            if (h == 0) sc2[p].x = s; else sc2[p].y = s;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const float c = (valid && j < Dsrc) ? -__ldg(a.centers + (size_t)fc * Dsrc + j) : 0.f;
                if (h == 0) nc2[p][j].x = c; else nc2[p][j].y = c;
            }
            if (DHT) {
#pragma unroll
                for (int o = 0; o < OT; ++o) {
                    const float w = (valid && o < O) ? __ldg(a.W + (size_t)k * O + o) : 0.f;
                    if (h == 0) w2[p][o].x = w; else w2[p][o].y = w;
                }
            }
        }
        contig[p] = (kidx[p][1] == kidx[p][0] + 1) && ((kidx[p][0] & 1) == 0);
    }

    // gate-sparse: the contiguous range of gate-table columns (regions) this warp's pairs belong to; a warp without
    // pairs gets an empty range and skips every query
    int wr_lo = Rspan, wr_hi = -1;
    if (SPARSE) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (ridx[0][h] < Rspan) { wr_lo = min(wr_lo, ridx[0][h]); wr_hi = max(wr_hi, ridx[0][h]); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            wr_lo = min(wr_lo, __shfl_xor_sync(0xffffffffu, wr_lo, o));
            wr_hi = max(wr_hi, __shfl_xor_sync(0xffffffffu, wr_hi, o));
        }
    }

    float2 aC[NKP][D], aL[NKP], aW[NKP][OTS];
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        aL[p] = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < D; ++j) aC[p][j] = make_float2(0.f, 0.f);
#pragma unroll
        for (int o = 0; o < OTS; ++o) aW[p][o] = make_float2(0.f, 0.f);
    }

    const long long ntiles = (a.B + TQ - 1) / TQ;
    float* dbs = dhs + (DHT ? 0 : (size_t)TQ * Kspan);   // bias-gradient accumulators: kDbScratch floats at the end
    if (tid < kDbScratch) dbs[tid] = 0.f;

    auto issue_tile = [&](long long tile, int stage) {
        float* sb = smem + (size_t)stage * stage_floats;
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        for (int i = tid; i < tq * Dsrc; i += NT) {
            const int q = i / Dsrc, j = i - q * Dsrc;
            cp_async4(sb + st_x + q * XS + j, a.x + q0 * Dsrc + i);
        }
        if (a.dy != nullptr)
            for (int i = tid; i < tq * O; i += NT) {
                const int q = i / O, o = i - q * O;
                cp_async4(sb + st_dy + q * DS + o, a.dy + q0 * O + i);
            }
        for (int i = tid; i < tq * Rspan; i += NT) {
            const int q = i / Rspan, rr = i - q * Rspan;
            const int r = min(r_lo + rr, a.R - 1);
            cp_async4(sb + st_g + q * GS + rr, a.gate + (q0 + q) * a.R + r);
        }
        if (has_dhid)
            for (int i = tid; i < tq * Kspan; i += NT) {
                const int q = i / Kspan, kk = i - q * Kspan;
                int k = k_direct ? kk : (f_lo + kk) % K;
                k = min(k, K - 1);
                cp_async4(sb + st_dh + i, a.dhidden + (q0 + q) * K + k);
            }
        cp_async_commit();
    };

    // one query of the current stage against all of this thread's kernel pairs
    auto eval = [&](const float* sb, int q) {
        float xv[XS];
        {
            const float4* xp = reinterpret_cast<const float4*>(sb + st_x + (size_t)q * XS);
#pragma unroll
            for (int i = 0; i < XS / 4; ++i) {
                const float4 t = xp[i];
                xv[4 * i] = t.x; xv[4 * i + 1] = t.y; xv[4 * i + 2] = t.z; xv[4 * i + 3] = t.w;
            }
        }
        float dyv[OTS];
        if (OT > 0) {
            const float* dp = sb + st_dy + (size_t)q * DS;
            if (DHT) {
                const float2* d2p = reinterpret_cast<const float2*>(dp);
#pragma unroll
                for (int o = 0; o < OT / 2; ++o) { const float2 t = d2p[o]; dyv[2 * o] = t.x; dyv[2 * o + 1] = t.y; }
            } else {
#pragma unroll
                for (int o = 0; o < OT; ++o) dyv[o] = o < a.ocur ? dp[a.o0 + o] : 0.f;
            }
        }
        const float* gq = sb + st_g + (size_t)q * GS;
        const float* dq = dhs + (size_t)q * Kspan;
#pragma unroll
        for (int p = 0; p < NKP; ++p) {
            const float2 g2 = make_float2(gq[ridx[p][0]], gq[ridx[p][1]]);
            float2 dh2;
            if (DHT) {
                dh2 = mul2(w2[p][0], dup2(dyv[0]));
#pragma unroll
                for (int o = 1; o < OT; ++o) dh2 = fma2(w2[p][o], dup2(dyv[o]), dh2);
            } else if (contig[p]) {
                dh2 = *reinterpret_cast<const float2*>(dq + kidx[p][0]);
            } else {
                dh2 = make_float2(dq[kidx[p][0]], dq[kidx[p][1]]);
            }
            float2 df[D];
            df[0] = add2(nc2[p][0], dup2(xv[0]));
            float2 d2 = mul2(df[0], df[0]);
#pragma unroll
            for (int j = 1; j < D; ++j) {
                df[j] = add2(nc2[p][j], dup2(xv[j]));
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
            for (int o = 0; o < OT; ++o) aW[p][o] = fma2(v, dup2(dyv[o]), aW[p][o]);
            float2 t = mul2(v, dh2);         // dh * gamma * phi   (= -dL/d(alpha^2), gaussian)
            if (BASIS == IRBFN_BASIS_INVERSE_QUADRATIC) t = mul2(t, phi);
#pragma unroll
            for (int j = 0; j < D; ++j) aC[p][j] = fma2(t, df[j], aC[p][j]);
            aL[p] = fma2(t, d2, aL[p]);
        }
    };

    __syncthreads();  // zero fill is visible before any cp.async lands on it
    long long tile = split;
    int stage = 0;
    if (tile < ntiles) issue_tile(tile, 0);

    for (; tile < ntiles; tile += a.splits, stage ^= 1) {
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        const float* sb = smem + (size_t)stage * stage_floats;
        cp_async_wait_all();
        __syncthreads();  // stage landed; everyone is done computing on the other stage
        if (tile + a.splits < ntiles) issue_tile(tile + a.splits, stage ^ 1);
        if (db_active(a.part_db, a.dy, chunk, O, NT)) db_tile(dbs, sb + st_dy, DS, tq, O, NT, tid);

        if (!DHT) {
            // ---- dh table of the tile: dh[q][k] = dy[q,:] . W[k,:] (+ dhidden[q][k]) ----
            for (int i = tid; i < tq * Kspan; i += NT) {
                const int q = i / Kspan, kk = i - q * Kspan;
                const int k = k_direct ? kk : (f_lo + kk) % K;
                float acc = has_dhid ? sb[st_dh + i] : 0.f;
                if (k < K && a.dy != nullptr) {
                    const float* dp = sb + st_dy + (size_t)q * DS;
                    for (int o = 0; o < O; ++o) acc = fmaf(dp[o], __ldg(a.W + (size_t)k * O + o), acc);
                }
                dhs[i] = acc;
            }
            __syncthreads();
        }

        if (SPARSE) {
            // ---- gate-sparse: 32 queries at a time, lane l looks at query qb + l and asks whether any region this
            // warp holds pairs of has gamma > gate_thr for it; the warp then walks the set bits of the ballot.  A
            // skipped query costs nothing beyond its share of the ballot (one LDS + compare per region of the warp).
            for (int qb = 0; qb < tq; qb += 32) {
                const int ql = qb + (tid & 31);
                bool live = false;
                if (ql < tq) {
                    const float* gq0 = sb + st_g + (size_t)ql * GS;    // row stride Rspan + 1: odd for the shipped shapes
                    for (int rr = wr_lo; rr <= wr_hi; ++rr) live |= gq0[rr] > a.gate_thr;
                }
                unsigned m = __ballot_sync(0xffffffffu, live);
                while (m) {
                    const int b0 = __ffs(m) - 1;
                    m &= m - 1;
                    if (m) {
                        const int b1 = __ffs(m) - 1;
                        m &= m - 1;
                        eval(sb, qb + b0);
                        eval(sb, qb + b1);
                    } else {
                        eval(sb, qb + b0);
                    }
                }
            }
        } else {
        // ---- hot loop over the queries of the tile, two at a time for ILP ----
        int q = 0;
        for (; q + 1 < tq; q += 2) {
            eval(sb, q);
            eval(sb, q + 1);
        }
        if (q < tq) eval(sb, q);
        }
        if (!DHT) __syncthreads();  // the dh table is rebuilt by the next tile
    }
    cp_async_wait_all();
    if (db_active(a.part_db, a.dy, chunk, O, NT)) db_flush(dbs, a.part_db, split, O, NT, tid);

    // ---- flush raw partial sums (the reduce kernel applies 2 exp(-2 log_sig)) ----
    float* part = a.part + (size_t)split * RK * PW;
#pragma unroll
    for (int p = 0; p < NKP; ++p) {
        const int pl = pslot + p * NT;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int f = f_lo + 2 * pl + h;
            if (lane_ok && pl < a.pairs_per_chunk && f < f_hi) {
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

// kernel pairs per thread as a function of the output window: keeps the accumulators in registers
__host__ __device__ constexpr int bwd_p_nkp_for(int OT) { return OT <= 4 ? IRBFN_BWDP_NKP_NARROW : 1; }

template <int D, int OT, int BASIS, int DHT, int SPARSE = 0>
inline int launch_bwd_p_inst(const BwdPArgs& a, int grid, int nt, size_t smem, cudaStream_t st) {
    auto kern = irbfn_bwd_p_kernel<D, OT, BASIS, bwd_p_nkp_for(OT), DHT, SPARSE>;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, nt, smem, st>>>(a);
    return (int)cudaGetLastError();
}

template <int D, int OT>
inline int launch_bwd_p_ot(const BwdPArgs& a, int basis, int dht, int grid, int nt, size_t smem, cudaStream_t st) {
    constexpr int CAN_DHT = OT > 0 ? 1 : 0;   // dh formed in-thread from the thread's own readout rows
    if (basis == IRBFN_BASIS_GAUSSIAN)
        return (dht && CAN_DHT) ? launch_bwd_p_inst<D, OT, IRBFN_BASIS_GAUSSIAN, CAN_DHT>(a, grid, nt, smem, st)
                                : launch_bwd_p_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 0>(a, grid, nt, smem, st);
    return (dht && CAN_DHT) ? launch_bwd_p_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, CAN_DHT>(a, grid, nt, smem, st)
                            : launch_bwd_p_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 0>(a, grid, nt, smem, st);
}

template <int D>
inline int launch_bwd_p(const BwdPArgs& a, int OT, int basis, int dht, int grid, int nt, size_t smem,
                        cudaStream_t st) {
    switch (OT) {
        case 0: return launch_bwd_p_ot<D, 0>(a, basis, dht, grid, nt, smem, st);
        case 2: return launch_bwd_p_ot<D, 2>(a, basis, dht, grid, nt, smem, st);
        case 4: return launch_bwd_p_ot<D, 4>(a, basis, dht, grid, nt, smem, st);
        case 8: return launch_bwd_p_ot<D, 8>(a, basis, dht, grid, nt, smem, st);
        case 10: return launch_bwd_p_ot<D, 10>(a, basis, dht, grid, nt, smem, st);
        case 16: return launch_bwd_p_ot<D, 16>(a, basis, dht, grid, nt, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

// Gate-sparse variants (SPARSE = 1): every output window and both dh forms, one kernel pair per thread.  Returns -1
// when no sparse instantiation exists (two pairs per thread): the caller launches the dense kernel, which computes
// the same sums without skipping.
template <int D, int OT>
inline int launch_bwd_p_sparse_ot(const BwdPArgs& a, int basis, int dht, int grid, int nt, size_t smem, cudaStream_t st) {
    constexpr int CAN_DHT = OT > 0 ? 1 : 0;
    if (basis == IRBFN_BASIS_GAUSSIAN)
        return (dht && CAN_DHT) ? launch_bwd_p_inst<D, OT, IRBFN_BASIS_GAUSSIAN, CAN_DHT, 1>(a, grid, nt, smem, st)
                                : launch_bwd_p_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 0, 1>(a, grid, nt, smem, st);
    return (dht && CAN_DHT) ? launch_bwd_p_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, CAN_DHT, 1>(a, grid, nt, smem, st)
                            : launch_bwd_p_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 0, 1>(a, grid, nt, smem, st);
}

template <int D>
inline int launch_bwd_p_sparse(const BwdPArgs& a, int OT, int basis, int dht, int grid, int nt, size_t smem,
                               cudaStream_t st) {
    if (bwd_p_nkp_for(OT) != 1) return -1;
    switch (OT) {
        case 2: return launch_bwd_p_sparse_ot<D, 2>(a, basis, dht, grid, nt, smem, st);
        case 4: return launch_bwd_p_sparse_ot<D, 4>(a, basis, dht, grid, nt, smem, st);
        case 8: return launch_bwd_p_sparse_ot<D, 8>(a, basis, dht, grid, nt, smem, st);
        case 10: return launch_bwd_p_sparse_ot<D, 10>(a, basis, dht, grid, nt, smem, st);
        case 16: return launch_bwd_p_sparse_ot<D, 16>(a, basis, dht, grid, nt, smem, st);
        default: return -1;
    }
}

}  // namespace irbfn
