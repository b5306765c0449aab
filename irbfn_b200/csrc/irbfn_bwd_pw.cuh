// IRBFN backward, parameter gradients for WIDE outputs and several regions (sm_100a).
//
// Reference: jax.value_and_grad(loss_fn)(params) through WCRBFNet.apply
//   scripts/train_nmpc_frenet.py:416 (train_step_fullint, O = 10) ; forward being differentiated:
//   model.py:187-196 (blend over regions, ONE shared readout), flax_rbf.py:275-283
//
// Why a second kernel.  irbfn_bwd_p_kernel keeps a (region, kernel) pair and ALL its
// gradient accumulators in one thread, including the pair's own copy of the readout-row gradient dW[k,:].
// But the readout is shared by the regions: dh_k = dy . W[k,:] does not depend on the region, and
// dW[k,:] = sum_q h_qk dy_q needs only the region-summed h_qk = sum_r gamma_qr phi_qrk.  Forming dh and a
// private dW per (query, region, kernel) costs 2*O lane-operations on top of the 3D+5 the distance,
// basis and centre/width gradients need: at O = 16, D = 6 that is 55 executed against 23 algorithmic.
//
// Layout here: a warp's lanes are the REGIONS of one kernel pair (groups of G = min(32, pow2ceil(R)) lanes;
// 32/G pairs per warp when R is small, several warps per pair when R > 32).
//   * dh[q][k] is built once per query tile for the CTA's few kernels (a shared-memory table, ~1 % of the
//     tile's work) and read as one broadcast LDS.64 per (query, pair);
//   * the region sum of v = gamma*phi crosses lanes through a per-warp shared-memory scratch, 16 queries at
//     a time: every lane stores its 16 packed v's (STS.64, conflict-free stride), then sums one query's
//     column over the lanes of its group (16 LDS.64 + 16 FADD2) and multiplies that h by dy[q,:] into ONE
//     private dW accumulator row per lane: O/16 packed FMAs per (query, pair) instead of 2*O;
//   * gamma is read straight from the (B,R) gate rows: a coalesced 128-byte line per (warp, query), shared
//     through L1 by the CTA's other kernel pairs, one query ahead of its use.
// Centre / width gradients stay thread-private exactly as in irbfn_bwd_p_kernel, and the partial sums go
// to the same workspace rows, so reduce_p_kernel is shared (a group's dW total is written to the row of its
// first region, zeros to the others).  Fixed partition, fixed order: deterministic.
#pragma once
#include <type_traits>

#include "irbfn_bwd_p.cuh"

namespace irbfn {

constexpr int kPwNQ = 16;                 // queries per cross-lane batch
constexpr int kPwLS = 2 * kPwNQ + 2;      // scratch floats per lane: 34 -> STS.64 / LDS.64 conflict-free

struct BwdPWArgs {
    const float* x;         // (B,Dsrc)
    const float* centers;   // (R,K,Dsrc)
    const float* log_sigs;  // (R,K)
    const float* W;         // (K,O)
    const float* gate;      // (B,R)
    const float* dy;        // (B,O)
    float* part;            // [splits][R*K][PW]   PW = D + 1 + OT
    float* part_db;         // [splits][O] column sums of dy (bias gradient partials), or nullptr
    long long B;
    int R, K, Dsrc, O;
    int G;                  // lanes per region group (power of two, <= 32)
    int NRG;                // region groups per kernel pair: ceil(R / G)
    int units;              // KP * NRG, unit u = kp * NRG + region-group
    int units_per_cta;
    int nchunks, splits;
    int TQ;                 // queries per tile (multiple of kPwNQ)
    int Kspan;              // dh table columns = 2 * (kernel pairs a CTA can touch), even
};

template <int D, int OT, int BASIS>
__global__ void __launch_bounds__((OT <= 10 && D <= 7) ? 512 : 384, 1) irbfn_bwd_pw_kernel(const BwdPWArgs a) {
    constexpr int PW = D + 1 + OT;
    constexpr int XS = bwdp_xs(D);
    constexpr int NQ = kPwNQ, LS = kPwLS;
    const int NT = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk = blockIdx.x % a.nchunks, split = blockIdx.x / a.nchunks;
    const int G = a.G, K = a.K, O = a.O, Dsrc = a.Dsrc, TQ = a.TQ, Kspan = a.Kspan;
    constexpr int WS = ((OT + 3) / 4) * 4;                    // readout-row stride of the staged W rows
    constexpr int DS = WS + (WS % 8 == 0 ? 4 : 0);            // dy row stride: pad columns stay zero; an odd
                                                              // multiple of 16 bytes keeps row-per-lane LDS.128 conflict-free
    const int upw = 32 / G;                                   // units per warp
    const int g = lane / G, rr = lane - g * G;                // group within the warp, lane within the group

    // ---- which (kernel pair, region) this lane owns ----
    const int u_lo = chunk * a.units_per_cta;
    const int u_hi = min(u_lo + a.units_per_cta, a.units);
    const int kp_lo = u_lo / a.NRG;
    const int u = u_lo + warp * upw + g;
    const bool unit_ok = u < u_hi;
    const int uc = unit_ok ? u : u_lo;
    const int kp = uc / a.NRG, rg = uc - kp * a.NRG;
    const int r = rg * G + rr;
    const bool lane_ok = unit_ok && r < a.R;
    const bool warp_full = __all_sync(0xffffffffu, lane_ok);  // warp-uniform: every lane owns a region
    const int kk = 2 * (kp - kp_lo);                          // dh table column of (k0, k1)
    const int k0 = 2 * kp;
    const bool k1_ok = k0 + 1 < K;

    extern __shared__ __align__(16) float smem[];
    const int st_x = 0, st_dy = TQ * XS;
    const int stage_floats = ((st_dy + TQ * DS + 3) / 4) * 4;
    float* dhs = smem + 2 * (size_t)stage_floats;             // [TQ][Kspan]
    float* Ws = dhs + (size_t)TQ * Kspan;                     // [Kspan][WS] readout rows of the CTA's kernels
    float* scr = Ws + (size_t)Kspan * WS + (size_t)warp * 32 * LS;   // this warp's [32][LS]

    // zero both stages (cp.async never writes the pad columns) and the dh table (rows beyond a ragged tile
    // are multiplied by gamma = 0 and must be finite)
    for (int i = tid; i < 2 * stage_floats + TQ * Kspan; i += NT) smem[i] = 0.f;
    for (int i = tid; i < Kspan * WS; i += NT) {
        const int c = i / WS, o = i - c * WS;
        const int k = 2 * kp_lo + c;
        Ws[i] = (k < K && o < O) ? __ldg(a.W + (size_t)k * O + o) : 0.f;
    }

    // ---- parameters of the lane's two kernels -> registers ----
    float2 nc2[D], sc2;
    {
        const size_t f0 = (size_t)(lane_ok ? r : 0) * K + k0;
        const bool v0 = lane_ok, v1 = lane_ok && k1_ok;
        sc2.x = v0 ? width_scale<BASIS>(__ldg(a.log_sigs + f0)) : 0.f;
        sc2.y = v1 ? width_scale<BASIS>(__ldg(a.log_sigs + f0 + 1)) : 0.f;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            nc2[j].x = (v0 && j < Dsrc) ? -__ldg(a.centers + f0 * Dsrc + j) : 0.f;
            nc2[j].y = (v1 && j < Dsrc) ? -__ldg(a.centers + (f0 + 1) * Dsrc + j) : 0.f;
        }
    }
    float2 aC[D], aL = make_float2(0.f, 0.f), aW[OT];
#pragma unroll
    for (int j = 0; j < D; ++j) aC[j] = make_float2(0.f, 0.f);
#pragma unroll
    for (int o = 0; o < OT; ++o) aW[o] = make_float2(0.f, 0.f);

    // ---- reader role in the cross-lane sum: which query column(s) and which source lanes ----
    // G >= NQ: one query (rr % NQ), NQ source lanes of the group starting at (rr / NQ) * NQ
    // G <  NQ: NQ / G queries (rr, rr + G, ...), all G lanes of the group
    const int rd_per = G >= NQ ? NQ : G;                      // source lanes per flush
    const int rd_n = NQ / rd_per;                             // flushes per batch
    const int rd_src = g * G + (G >= NQ ? (rr / NQ) * NQ : 0);
    const int rd_q0 = G >= NQ ? rr % NQ : rr;

    const long long gpitch = (long long)a.R * (long long)sizeof(float);   // bytes between gate rows
    const int dhpitch2 = 2 * Kspan;                                       // floats between dh rows two queries apart
    const long long ntiles = (a.B + TQ - 1) / TQ;
    float* dbs = Ws + (size_t)Kspan * WS + (size_t)(NT / 32) * 32 * LS;   // after the warps' scratch: kDbScratch floats
    if (tid < kDbScratch) dbs[tid] = 0.f;
    auto issue_tile = [&](long long tile, int stage) {
        float* sb = smem + (size_t)stage * stage_floats;
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        for (int i = tid; i < tq * Dsrc; i += NT) {
            const int q = i / Dsrc, j = i - q * Dsrc;
            cp_async4(sb + st_x + q * XS + j, a.x + q0 * Dsrc + i);
        }
        for (int i = tid; i < tq * O; i += NT) {
            const int q = i / O, o = i - q * O;
            cp_async4(sb + st_dy + q * DS + o, a.dy + q0 * O + i);
        }
        cp_async_commit();
    };

    __syncthreads();
    long long tile = split;
    int stage = 0;
    if (tile < ntiles) issue_tile(tile, 0);

    for (; tile < ntiles; tile += a.splits, stage ^= 1) {
        const long long q0 = tile * TQ;
        const int tq = (int)min((long long)TQ, a.B - q0);
        const float* sb = smem + (size_t)stage * stage_floats;
        cp_async_wait_all();
        __syncthreads();   // stage landed; everyone is done with the other stage and with the dh table
        if (tile + a.splits < ntiles) issue_tile(tile + a.splits, stage ^ 1);
        if (db_active(a.part_db, a.dy, chunk, O, NT)) db_tile(dbs, sb + st_dy, DS, tq, O, NT, tid);

        // ---- dh table of the tile: dh[q][c] = dy[q,:] . W[kp_lo*2 + c, :] ----
        // one thread per query: its dy row in registers, the readout rows as broadcast LDS.128
        for (int q = tid; q < tq; q += NT) {
            float dyv[DS];
#pragma unroll
            for (int t = 0; t < DS / 4; ++t) {
                const float4 v4 = reinterpret_cast<const float4*>(sb + st_dy + (size_t)q * DS)[t];
                dyv[4 * t] = v4.x; dyv[4 * t + 1] = v4.y; dyv[4 * t + 2] = v4.z; dyv[4 * t + 3] = v4.w;
            }
#pragma unroll 2
            for (int c = 0; c < Kspan; c += 2) {
                const float4* w0 = reinterpret_cast<const float4*>(Ws + (size_t)c * WS);
                const float4* w1 = reinterpret_cast<const float4*>(Ws + (size_t)(c + 1) * WS);
                float2 acc = make_float2(0.f, 0.f);
#pragma unroll
                for (int t = 0; t < WS / 4; ++t) {
                    const float4 u0 = w0[t], u1 = w1[t];
                    acc.x = fmaf(dyv[4 * t], u0.x, acc.x); acc.y = fmaf(dyv[4 * t], u1.x, acc.y);
                    acc.x = fmaf(dyv[4 * t + 1], u0.y, acc.x); acc.y = fmaf(dyv[4 * t + 1], u1.y, acc.y);
                    acc.x = fmaf(dyv[4 * t + 2], u0.z, acc.x); acc.y = fmaf(dyv[4 * t + 2], u1.z, acc.y);
                    acc.x = fmaf(dyv[4 * t + 3], u0.w, acc.x); acc.y = fmaf(dyv[4 * t + 3], u1.w, acc.y);
                }
                *reinterpret_cast<float2*>(dhs + (size_t)q * Kspan + c) = acc;
            }
        }
        __syncthreads();

        // gamma of (query, this lane's region): a ring of four loads in flight, four queries ahead of use
        // (a per-lane pointer bumped by the row pitch: two integer instructions per load)
        const char* gptr = reinterpret_cast<const char*>(a.gate + q0 * a.R + (lane_ok ? r : 0));
        float gr[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            gr[t] = (lane_ok && t < tq) ? __ldg(reinterpret_cast<const float*>(gptr)) : 0.f;
            gptr += gpitch;
        }
        // x rows and dh pairs of the next TWO queries, loaded while the current two are computed (row TQ and
        // TQ + 1 of either table are finite padding)
        float4 xn[2][XS / 4];
        float2 dhn[2];
        const float* dhp = dhs + kk;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
#pragma unroll
            for (int t = 0; t < XS / 4; ++t) xn[u][t] = reinterpret_cast<const float4*>(sb + st_x + u * XS)[t];
            dhn[u] = *reinterpret_cast<const float2*>(dhp + u * Kspan);
        }

        // one batch of 16 queries, two at a time with their instructions interleaved (two independent
        // dependency chains per warp: with three warps per scheduler a single chain leaves the FP32 pipe
        // waiting on its own 4-cycle latencies): thread-private centre / width gradients, v = gamma * phi to
        // the scratch.  FAST: every lane of the warp owns a region and the ring's loads stay inside the tile.
        auto batch = [&](int qb, auto fast_tag) {
            constexpr bool FAST = decltype(fast_tag)::value;
            const float* xb = sb + st_x + (size_t)qb * XS;
            float* sw = scr + lane * LS;
#pragma unroll
            for (int i = 0; i < NQ; i += 2) {
                float gq[2], xv[2][XS];
                float2 dh2[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    gq[u] = gr[(i + u) & 3];
                    if (FAST) gr[(i + u) & 3] = __ldg(reinterpret_cast<const float*>(gptr));
                    else gr[(i + u) & 3] = (lane_ok && qb + i + u + 4 < tq) ? __ldg(reinterpret_cast<const float*>(gptr)) : 0.f;
                    gptr += gpitch;
#pragma unroll
                    for (int t = 0; t < XS / 4; ++t) {
                        xv[u][4 * t] = xn[u][t].x; xv[u][4 * t + 1] = xn[u][t].y;
                        xv[u][4 * t + 2] = xn[u][t].z; xv[u][4 * t + 3] = xn[u][t].w;
                    }
                    dh2[u] = dhn[u];
                }
                dhp += dhpitch2;
#pragma unroll
                for (int u = 0; u < 2; ++u) {
#pragma unroll
                    for (int t = 0; t < XS / 4; ++t) xn[u][t] = reinterpret_cast<const float4*>(xb + (i + 2 + u) * XS)[t];
                }
                dhn[0] = *reinterpret_cast<const float2*>(dhp);
                dhn[1] = *reinterpret_cast<const float2*>(dhp + Kspan);
                float2 df[2][D], d2[2], d2b[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    df[u][0] = add2(nc2[0], dup2(xv[u][0]));
                    d2[u] = mul2(df[u][0], df[u][0]);
                    if (D > 1) {
                        df[u][1] = add2(nc2[1], dup2(xv[u][1]));
                        d2b[u] = mul2(df[u][1], df[u][1]);
                    }
                }
#pragma unroll
                for (int j = 2; j < D; ++j) {
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        df[u][j] = add2(nc2[j], dup2(xv[u][j]));
                        if (j & 1) d2b[u] = fma2(df[u][j], df[u][j], d2b[u]);
                        else d2[u] = fma2(df[u][j], df[u][j], d2[u]);
                    }
                }
                float2 phi[2], v[2], t[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (D > 1) d2[u] = add2(d2[u], d2b[u]);
                    if (BASIS == IRBFN_BASIS_GAUSSIAN) {
                        const float2 e = mul2(d2[u], sc2);
                        phi[u] = make_float2(ex2_approx(e.x), ex2_approx(e.y));
                    } else {
                        const float2 e = fma2(d2[u], sc2, make_float2(1.f, 1.f));
                        phi[u] = make_float2(rcp_approx(e.x), rcp_approx(e.y));
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    v[u] = mul2(dup2(gq[u]), phi[u]);            // gamma * phi (0 beyond the tile: gamma = 0)
                    t[u] = mul2(v[u], dh2[u]);
                    if (BASIS == IRBFN_BASIS_INVERSE_QUADRATIC) t[u] = mul2(t[u], phi[u]);
                }
                *reinterpret_cast<float2*>(sw + 2 * i) = v[0];
                *reinterpret_cast<float2*>(sw + 2 * i + 2) = v[1];
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    aC[j] = fma2(t[0], df[0][j], aC[j]);
                    aC[j] = fma2(t[1], df[1][j], aC[j]);
                }
                aL = fma2(t[0], d2[0], aL);
                aL = fma2(t[1], d2[1], aL);
            }
        };

        for (int qb = 0; qb < tq; qb += NQ) {
            if (warp_full && qb + NQ + 4 <= tq) batch(qb, std::true_type{});
            else batch(qb, std::false_type{});
            __syncwarp();
            // ---- region sum h (one query column over the group's lanes) and its dW contribution ----
#pragma unroll 1
            for (int f = 0; f < rd_n; ++f) {
                const int qi = rd_q0 + f * G;                 // query within the batch (f == 0 when G >= NQ)
                const float* sp = scr + (size_t)rd_src * LS + 2 * qi;
                float2 h = make_float2(0.f, 0.f);
                if (rd_per == NQ) {                            // G >= 16: always 16 source lanes, fully unrolled
                    float2 h1 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int s = 0; s < NQ; s += 2) {
                        h = add2(h, *reinterpret_cast<const float2*>(sp + s * LS));
                        h1 = add2(h1, *reinterpret_cast<const float2*>(sp + (s + 1) * LS));
                    }
                    h = add2(h, h1);
                } else {
#pragma unroll 4
                    for (int s = 0; s < rd_per; ++s) h = add2(h, *reinterpret_cast<const float2*>(sp + (size_t)s * LS));
                }
                const float* dp = sb + st_dy + (size_t)(qb + qi) * DS;
                float dyv[DS];
#pragma unroll
                for (int t = 0; t < DS / 4; ++t) {
                    const float4 v4 = reinterpret_cast<const float4*>(dp)[t];
                    dyv[4 * t] = v4.x; dyv[4 * t + 1] = v4.y; dyv[4 * t + 2] = v4.z; dyv[4 * t + 3] = v4.w;
                }
#pragma unroll
                for (int o = 0; o < OT; ++o) aW[o] = fma2(h, dup2(dyv[o]), aW[o]);
            }
            __syncwarp();
        }
    }
    cp_async_wait_all();
    if (db_active(a.part_db, a.dy, chunk, O, NT)) db_flush(dbs, a.part_db, split, O, NT, tid);

    // ---- dW: total over the group's lanes (fixed butterfly order), kept by the group's first lane ----
    for (int m = G >> 1; m >= 1; m >>= 1) {
#pragma unroll
        for (int o = 0; o < OT; ++o) {
            aW[o].x += __shfl_xor_sync(0xffffffffu, aW[o].x, m);
            aW[o].y += __shfl_xor_sync(0xffffffffu, aW[o].y, m);
        }
    }
    if (!lane_ok) return;
    float* part = a.part + (size_t)split * a.R * K * PW;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (h == 1 && !k1_ok) break;
        float* o = part + ((size_t)r * K + k0 + h) * PW;
#pragma unroll
        for (int j = 0; j < D; ++j) o[j] = h == 0 ? aC[j].x : aC[j].y;
        o[D] = h == 0 ? aL.x : aL.y;
#pragma unroll
        for (int oo = 0; oo < OT; ++oo) o[D + 1 + oo] = rr == 0 ? (h == 0 ? aW[oo].x : aW[oo].y) : 0.f;
    }
}

template <int D, int OT>
inline int launch_bwd_pw_ot(const BwdPWArgs& a, int basis, int grid, int nt, size_t smem, cudaStream_t st) {
    auto run = [&](auto kern) {
        const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, nt, smem, st>>>(a);
        return (int)cudaGetLastError();
    };
    if (basis == IRBFN_BASIS_GAUSSIAN) return run(irbfn_bwd_pw_kernel<D, OT, IRBFN_BASIS_GAUSSIAN>);
    return run(irbfn_bwd_pw_kernel<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC>);
}

template <int D>
inline int launch_bwd_pw(const BwdPWArgs& a, int OT, int basis, int grid, int nt, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 8: return launch_bwd_pw_ot<D, 8>(a, basis, grid, nt, smem, st);
        case 10: return launch_bwd_pw_ot<D, 10>(a, basis, grid, nt, smem, st);
        case 16: return launch_bwd_pw_ot<D, 16>(a, basis, grid, nt, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
