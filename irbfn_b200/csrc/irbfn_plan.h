// Host-side launch planning: how the (R,K) parameter set is cut into shared-memory stages.
// Plain C++ (no CUDA types) so it can be unit-tested on a machine without a GPU.
#pragma once
#include <stddef.h>

namespace irbfn {

// Forward tile geometry; must match FwdShape in irbfn_fwd.cuh.
constexpr int kFwdQ = 4;      // queries per thread
constexpr int kFwdKT = 4;     // kernels per register tile and per transposed shared-memory tile
constexpr int kFwdTQ = 1024;  // queries per CTA tile of the 256-thread kernels (backward X)

// floats of one transposed 4-kernel tile: rows -c_0..-c_{D-1} and the width scales
inline int fwd_tile_floats(int D) { return (D + 1) * 4; }

// Threads per CTA of the forward kernel: 256 for every output window (the one compiled CTA size; 512-thread CTAs at
// <= 128 registers per thread were measured slower in round 1).
inline int fwd_threads_for(int OT) { (void)OT; return 256; }

// Compiled output-window widths of the forward kernel.
inline int fwd_ot_for(int ocur) {
    if (ocur <= 0) return 0;
    if (ocur <= 2) return 2;
    if (ocur <= 4) return 4;
    if (ocur <= 8) return 8;
    if (ocur <= 10) return 10;
    return 16;
}

// Small-batch forward (irbfn_fwd_small.cu): a CTA owns a few queries and its threads share the (region, kernel)
// pairs.  A CTA re-reads every parameter, so the parameter set must be small (it stays in L2) and the batch a few
// hundred queries; beyond that the query-stationary kernel's register tiling wins (measured: scripts/dev_latency.py).
constexpr long long kFwdSmallMaxPairs = 16384;        // R*K
constexpr long long kFwdSmallMaxWork = 5ll << 20;     // B*R*K (C2 shape: up to 4,369 queries)
constexpr long long kFwdSmallQ1MaxB = 148;            // one query per CTA up to one CTA per SM, then four per CTA
inline bool fwd_small_applies(long long B, int R, int K, int D, int O) {
    const long long pairs = (long long)R * K;
    if (B <= 0 || O <= 0 || O > 16 || D > 8 || R > 2048 || K < 1) return false;
    if (pairs > kFwdSmallMaxPairs) return false;
    return B * pairs <= kFwdSmallMaxWork;
}

// Split of the kernels of every region over gridDim.y for batches that do not fill the device with one CTA
// per 1,024-query tile.  Returns the number of splits (1 = none) and the kernels per split (multiple of
// kFwdKT).  The choice minimises a cost model fitted to B200 measurements (scripts/dev_fwd_split.py), in
// units of one (kernel tile, region) step of a CTA:
//     cost(ns) = ceil(tiles * ns / sms) * (c0 + ceil(Kt / ns) * R) + (ns > 1 ? c_reduce : 0)
// c0 = per-CTA fixed work (query load, gate, parameter staging, launch ramp), c_reduce = the partial-output
// round trip.  CTAs beyond one per SM share an SM, which costs the same as a second round.
inline int fwd_split(long long B, int R, int K, int sms, int* k_per_split) {
    const int Kt = (K + kFwdKT - 1) / kFwdKT;
    *k_per_split = Kt * kFwdKT;
    const long long tiles = (B + kFwdTQ - 1) / kFwdTQ;
    if (tiles <= 0 || tiles >= 4ll * sms || Kt < 2 || sms <= 0) return 1;
    const double c0 = 24.0 + R, c_reduce = 10.0;
    double best = 0.0;
    int best_kt = Kt;
    for (int ns = 1; ns <= Kt && ns <= 64; ++ns) {
        const int kt = (Kt + ns - 1) / ns;                 // kernel tiles per split
        const int nsp = (Kt + kt - 1) / kt;                // splits that size gives
        if (nsp != ns) continue;                           // same as a smaller ns already seen
        const long long rounds = (tiles * nsp + sms - 1) / sms;
        const double cost = (double)rounds * (c0 + (double)kt * R) + (nsp > 1 ? c_reduce : 0.0);
        if (ns == 1 || cost < best) { best = cost; best_kt = kt; }
    }
    const int nsp = (Kt + best_kt - 1) / best_kt;
    if (nsp <= 1) return 1;
    *k_per_split = best_kt * kFwdKT;
    return nsp;
}

struct FwdPlan {
    int RC;          // regions per stage
    int KC;          // kernels per stage (multiple of kFwdKT)
    int stage_once;  // whole parameter set resident: staged once per CTA
    size_t smem_bytes;
    int NT;          // threads per CTA (forward only)
};

// stage [RC][KC/4] tiles + readout rows [KC][OT] + gate table [RC][NT] float4
inline size_t fwd_smem_bytes(int RC, int KC, int D, int OT, int NT) {
    return ((size_t)RC * (KC / kFwdKT) * fwd_tile_floats(D) + (size_t)KC * OT + (size_t)RC * NT * kFwdQ) *
           sizeof(float);
}

inline bool plan_forward(int R, int K, int D, int OT, int smem_limit, FwdPlan* p) {
    const int NT = fwd_threads_for(OT);
    const int Kp = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
    if (fwd_smem_bytes(R, Kp, D, OT, NT) <= (size_t)smem_limit) {
        *p = FwdPlan{R, Kp, 1, fwd_smem_bytes(R, Kp, D, OT, NT), NT};
        return true;
    }
    const int want = Kp < 64 ? Kp : 64;
    const int cands[6] = {R < 32 ? R : 32, 16, 8, 4, 2, 1};
    for (int i = 0; i < 6; ++i) {
        const int RC = cands[i] < R ? cands[i] : R;
        if (RC <= 0) continue;
        const size_t fixed = (size_t)RC * NT * kFwdQ * sizeof(float);
        if (fixed >= (size_t)smem_limit) continue;
        const size_t per_k = ((size_t)RC * fwd_tile_floats(D) / kFwdKT + OT) * sizeof(float);
        long long KC = (long long)(((size_t)smem_limit - fixed) / per_k);
        KC = (KC / kFwdKT) * kFwdKT;
        if (KC > Kp) KC = Kp;
        if (KC >= want || (RC == 1 && KC >= kFwdKT)) {
            *p = FwdPlan{RC, (int)KC, 0, fwd_smem_bytes(RC, (int)KC, D, OT, NT), NT};
            return true;
        }
    }
    return false;
}

// ---------------------------------------------------------------------------------------
// backward X (query-side): same staging scheme as the forward, wider records, two gate tables
// ---------------------------------------------------------------------------------------
inline int bwdx_tile_floats(int D) { return (D + 2) * 4; }   // rows -c_0..-c_{D-1}, s, m of 4 kernels

// stage [RC][KC/4] tiles + readout rows [KC][OT] + gate and d-gate tables [RC][1024] each
inline size_t bwdx_smem_bytes(int RC, int KC, int D, int OT) {
    return ((size_t)RC * (KC / kFwdKT) * bwdx_tile_floats(D) + (size_t)KC * OT + (size_t)2 * RC * kFwdTQ) *
           sizeof(float);
}

inline bool plan_bwd_x(int R, int K, int D, int OT, int smem_limit, FwdPlan* p) {
    const int Kp = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
    if (bwdx_smem_bytes(R, Kp, D, OT) <= (size_t)smem_limit) {
        *p = FwdPlan{R, Kp, 1, bwdx_smem_bytes(R, Kp, D, OT), 256};
        return true;
    }
    const int want = Kp < 64 ? Kp : 64;
    const int cands[5] = {R < 16 ? R : 16, 8, 4, 2, 1};
    for (int i = 0; i < 5; ++i) {
        const int RC = cands[i] < R ? cands[i] : R;
        if (RC <= 0) continue;
        const size_t fixed = (size_t)2 * RC * kFwdTQ * sizeof(float);
        if (fixed >= (size_t)smem_limit) continue;
        const size_t per_k = ((size_t)RC * bwdx_tile_floats(D) / kFwdKT + OT) * sizeof(float);
        long long KC = (long long)(((size_t)smem_limit - fixed) / per_k);
        KC = (KC / kFwdKT) * kFwdKT;
        if (KC > Kp) KC = Kp;
        if (KC >= want || (RC == 1 && KC >= kFwdKT)) {
            *p = FwdPlan{RC, (int)KC, 0, bwdx_smem_bytes(RC, (int)KC, D, OT), 256};
            return true;
        }
    }
    return false;
}

// ---------------------------------------------------------------------------------------
// backward P (parameter gradients): kernel pairs per thread, chunks of pairs per CTA,
// query splits, tile size.  Must match irbfn_bwd_p.cuh.
// ---------------------------------------------------------------------------------------
#ifndef IRBFN_BWDP_NKP_NARROW
#define IRBFN_BWDP_NKP_NARROW 1
#endif
inline int bwdp_nkp_for(int OT) { return OT <= 4 ? IRBFN_BWDP_NKP_NARROW : 1; }
// Threads per CTA: a thread that holds ONE kernel pair and few outputs needs ~95 registers, so 640 threads
// (20 warps, five on every SM sub-partition) fit one SM; the wide-output variants (~165 registers) get 384
// (12 warps, three per sub-partition).
inline int bwdp_max_threads(int OT) { return (bwdp_nkp_for(OT) == 1 && OT <= 4) ? 640 : 384; }

struct BwdPPlan {
    int OT, NKP, NT;
    int ppw;             // kernel pairs per warp (32 = every lane holds pairs)
    int nchunks, pairs_per_chunk, splits;
    int TQ, Kspan, Rspan;
    int PW;              // partial-sum row width: D + 1 + OT
    int dht;             // dh formed in-thread from the thread's own readout rows (narrow outputs)
    size_t smem_bytes;
    size_t part_floats;  // splits * R*K * PW
};

// row strides of the shared-memory stage; must match irbfn_bwd_p.cuh
inline int bwdp_xs_h(int D) { return ((D + 3) / 4) * 4; }
inline int bwdp_ds_h(int O) { return O <= 0 ? 0 : (O <= 2 ? 2 : ((O + 3) / 4) * 4); }

constexpr int kDbScratch_h = 1024;   // bias-gradient accumulators at the end of shared memory (kDbScratch in irbfn_bwd_p.cuh)

inline size_t bwdp_smem_floats(int TQ, int D, int O, int Kspan, int Rspan, int has_dhid, int dht) {
    size_t st = (size_t)TQ * bwdp_xs_h(D) + (size_t)TQ * bwdp_ds_h(O) + (size_t)TQ * (Rspan + 1);
    st = ((st + 3) / 4) * 4;
    st += has_dhid ? (size_t)TQ * Kspan : 0;
    st = ((st + 3) / 4) * 4;
    return 2 * st + (dht ? 0 : (size_t)TQ * Kspan) + kDbScratch_h;
}

// sparse = 1 (irbfn_backward_sparse): dh is formed in the thread whenever one output window covers O, also when
// several regions share the chunk's kernels -- the shared per-tile dh table serves every (query, region) pair, and a
// gate-sparse call evaluates only a fraction of them.
inline bool plan_bwd_p(long long B, int R, int K, int D, int Dsrc, int O, int OT, int has_dhid, int sms,
                       int smem_limit, BwdPPlan* p, int sparse = 0) {
    const long long RK = (long long)R * K;
    if (RK <= 0 || RK > (1ll << 30)) return false;
    const int P = (int)((RK + 1) / 2);
    const int NKP = bwdp_nkp_for(OT);
    const int maxt = bwdp_max_threads(OT);
    const int per_cta = maxt * NKP;
    const int nchunks = (P + per_cta - 1) / per_cta;
    const int Pc = (P + nchunks - 1) / nchunks;
    int NT = (((Pc + NKP - 1) / NKP + 31) / 32) * 32;
    if (NT > maxt) NT = maxt;
    int ppw = 32;
    if (NKP == 1) {
        // The SM places warp w on sub-partition w % 4: a warp count that is a multiple of four with the pairs
        // spread evenly over the warps keeps the four FP32 pipes equally loaded.  (600 pairs on 10 warps of 2
        // pairs per lane put 3/3/2/1.4 warps of work on the four pipes; the barrier per query tile then held
        // the kernel to the slowest: 65 % FMA-pipe utilisation in ncu with math_pipe_throttle AND barrier on top.)
        int nw = (((Pc + 31) / 32 + 3) / 4) * 4;
        if (nw * 32 > maxt) nw = maxt / 32;
        ppw = (Pc + nw - 1) / nw;
        NT = nw * 32;
    }
    const long long flen = 2ll * Pc;
    int Kspan = flen >= K ? K : (int)flen;
    Kspan = (Kspan + 1) & ~1;
    int Rspan = (int)((flen + K - 1) / K) + 1;
    if (Rspan > R) Rspan = R;
    (void)Dsrc;
    // in-thread dh: narrow outputs always; wide single-window outputs when the chunk's kernels belong to
    // one region (a per-tile dh table would be used once, and costs 227 KB-limited tiles of ~80 queries)
    const bool narrow = O >= 1 && O <= 4 && (OT == 2 || OT == 4);
    const bool lone = O >= 1 && O <= 16 && OT >= O && flen < K;
    const bool any1 = sparse && O >= 1 && O <= 16 && OT >= O;
    const int dht = (!has_dhid && (narrow || lone || any1)) ? 1 : 0;
    int TQ = 256;
    while (TQ >= 4 && bwdp_smem_floats(TQ, D, O, Kspan, Rspan, has_dhid, dht) * sizeof(float) > (size_t)smem_limit)
        TQ -= 4;
    if (TQ < 4) return false;
    const int slots = sms;
    long long splits = nchunks <= slots ? slots / nchunks : (4ll * slots + nchunks - 1) / nchunks;
    if (splits < 1) splits = 1;
    // Every split walks m = ceil(tiles / splits) tiles; shrink the tile so that the last of those rounds is
    // full too (32,768 queries on 148 SMs: 147 tiles of 224 in one round instead of 128 tiles of 256).
    if (B > 0) {
        const long long m = (B + splits * TQ - 1) / (splits * TQ);
        long long tqb = (B + splits * m - 1) / (splits * m);
        tqb = (tqb + 3) & ~3ll;
        if (tqb < 4) tqb = 4;
        if (tqb < TQ) TQ = (int)tqb;
    }
    const long long ntiles = (B + TQ - 1) / TQ;
    if (splits > ntiles) splits = ntiles;
    if (splits < 1) splits = 1;
    p->OT = OT; p->NKP = NKP; p->NT = NT; p->ppw = ppw;
    p->nchunks = nchunks; p->pairs_per_chunk = Pc; p->splits = (int)splits;
    p->TQ = TQ; p->Kspan = Kspan; p->Rspan = Rspan;
    p->PW = D + 1 + OT;
    p->dht = dht;
    p->smem_bytes = bwdp_smem_floats(TQ, D, O, Kspan, Rspan, has_dhid, dht) * sizeof(float);
    p->part_floats = (size_t)splits * (size_t)RK * (size_t)p->PW;
    return true;
}

// ---------------------------------------------------------------------------------------
// backward P, wide outputs over several regions (irbfn_bwd_pw.cuh): lanes = regions of a kernel pair.
// Chosen when the readout is wide enough (O > 4) for the per-(region, kernel) dh / dW work of the
// thread-private kernel to dominate, several regions share the readout (R >= 2), one output window
// suffices (O <= 16) and the cotangent is dy (no dhidden).
// ---------------------------------------------------------------------------------------
constexpr int kPwNQ_h = 16;              // must match kPwNQ / kPwLS in irbfn_bwd_pw.cuh
constexpr int kPwLS_h = 2 * kPwNQ_h + 2;

struct BwdPWPlan {
    int OT, NT;
    int G, NRG, units, units_per_cta, nchunks, splits;
    int TQ, Kspan, PW;
    size_t smem_bytes;
    size_t part_floats;
};

inline bool bwdpw_applies(int R, int O, int has_dhid) { return R >= 2 && O > 4 && O <= 16 && !has_dhid; }

inline size_t bwdpw_smem_floats(int TQ, int D, int O, int OT, int Kspan, int NW) {
    (void)O;
    const int WS = ((OT + 3) / 4) * 4, DS = WS + (WS % 8 == 0 ? 4 : 0);   // must match irbfn_bwd_pw.cuh
    size_t st = (size_t)TQ * bwdp_xs_h(D) + (size_t)TQ * DS;
    st = ((st + 3) / 4) * 4;
    return 2 * st + (size_t)TQ * Kspan + (size_t)Kspan * WS + (size_t)NW * 32 * kPwLS_h + kDbScratch_h;
}

inline bool plan_bwd_pw(long long B, int R, int K, int D, int O, int OT, int sms, int smem_limit, BwdPWPlan* p) {
    if (!bwdpw_applies(R, O, 0) || OT < O || (long long)R * K > (1ll << 30)) return false;
    int G = 32;
    if (R < 32) { G = 2; while (G < R) G <<= 1; }
    const int NRG = (R + G - 1) / G;
    const int KP = (K + 1) / 2;
    const long long units_ll = (long long)KP * NRG;
    if (units_ll > (1ll << 30)) return false;
    const int units = (int)units_ll;
    const int upw = 32 / G;                                  // units per warp
    const int max_warps = (OT <= 10 && D <= 7) ? 16 : 12;    // register budget: 128 / 168 per thread (irbfn_bwd_pw.cuh)
    const int per_cta = max_warps * upw;
    const int nchunks = (units + per_cta - 1) / per_cta;
    const int upc = (units + nchunks - 1) / nchunks;         // units per CTA
    int NW = (((upc + upw - 1) / upw + 3) / 4) * 4;          // a multiple of four warps: one per SM sub-partition
    if (NW > max_warps) NW = max_warps;
    // kernel pairs one CTA can touch: its units span at most ceil(upc / NRG) + 1 pairs
    int Kspan = 2 * ((upc + NRG - 1) / NRG + (NRG > 1 ? 1 : 0));
    if (Kspan > 2 * KP) Kspan = 2 * KP;
    int TQ = 256;
    while (TQ >= kPwNQ_h && bwdpw_smem_floats(TQ, D, O, OT, Kspan, NW) * sizeof(float) > (size_t)smem_limit) TQ -= kPwNQ_h;
    if (TQ < kPwNQ_h) return false;
    // splits: fill the device, and prefer a grid close to a whole number of waves
    long long splits = nchunks <= sms ? sms / nchunks : (4ll * sms + nchunks - 1) / nchunks;
    if (splits < 1) splits = 1;
    if (nchunks > sms) {
        double best = 0.0;
        long long best_s = splits;
        for (long long s = splits; s <= splits + 6; ++s) {
            const long long grid = s * nchunks, waves = (grid + sms - 1) / sms;
            const double eff = (double)grid / (double)(waves * sms);
            if (eff > best + 1e-9) { best = eff; best_s = s; }
        }
        splits = best_s;
    }
    if (B > 0) {
        const long long m = (B + splits * TQ - 1) / (splits * TQ);
        long long tqb = (B + splits * m - 1) / (splits * m);
        tqb = ((tqb + kPwNQ_h - 1) / kPwNQ_h) * kPwNQ_h;
        if (tqb < TQ) TQ = (int)tqb;
    }
    const long long ntiles = (B + TQ - 1) / TQ;
    if (splits > ntiles) splits = ntiles;
    if (splits < 1) splits = 1;
    p->OT = OT; p->NT = NW * 32;
    p->G = G; p->NRG = NRG; p->units = units; p->units_per_cta = upc; p->nchunks = nchunks; p->splits = (int)splits;
    p->TQ = TQ; p->Kspan = Kspan; p->PW = D + 1 + OT;
    p->smem_bytes = bwdpw_smem_floats(TQ, D, O, OT, Kspan, NW) * sizeof(float);
    p->part_floats = (size_t)splits * (size_t)R * (size_t)K * (size_t)p->PW;
    return true;
}

}  // namespace irbfn
