// IRBFN forward kernel (sm_100a), the two variants that ADD something to the plain dense forward of irbfn_fwd.cuh:
//   MODE kFwdSparse   gate-sparse, region-binned evaluation (irbfn_sparse.cu)
//   MODE kFwdGateOut  training forward that also writes the gate for the backward (irbfn_forward_gate)
// Same pipeline as irbfn_fwd.cuh (that file stays byte-for-byte what round 1 tuned: even code that is skipped at
// run time changed ptxas' schedule of the wide-output dense variants, C5 forward 2.06 -> 2.33 s).
//
// Reference arithmetic (paths relative to the reference checkout):
//   gate    src/irbfn_mpc/model.py:70-95      blend   src/irbfn_mpc/model.py:187-193
//   layer   flax_rbf/flax_rbf.py:275-283      readout src/irbfn_mpc/model.py:196
//
// Design:
//   * query-stationary: each thread owns Q = 4 queries, held in registers as two packed pairs
//     (x_{q0,j}, x_{q1,j}); every fp32 operation of the distance loop is a packed FADD2/FFMA2
//     whose centre operand is a scalar register broadcast to both lanes (SASS `Rn.F32`), so a
//     (query, region, kernel) evaluation costs D+2 issue slots for 2D+2 FP32-pipe cycles and the
//     kernel is bound by the FP32 pipe, not by instruction issue.
//   * parameters are staged in shared memory once per block (or once per (region,kernel) chunk
//     when they do not fit) as tiles of KT = 4 kernels, TRANSPOSED: row j of a tile holds -c_j of
//     its four kernels, row D their width scales s = exp(-2 log_sig) (times -log2 e for the
//     gaussian).  One broadcast LDS.128 per feature feeds 8 independent packed chains
//     (4 kernels x 2 query pairs): feature-outer, chain-inner, so latency hides inside one warp.
//   * the gate lives in a thread-private column of a shared-memory table gtab[region][thread],
//     computed once per query tile from the packed bound tables (no (B,R) array in HBM).
//   * loop order: kernel tile outer, region inner, so the blended activation h_k is complete in
//     registers before the readout touches it; (B,R,K) and (B,K) never exist.  The basis function
//     (MUFU) of one (tile, region) step is evaluated DURING the next step, two MUFUs after each
//     feature of its distance loop, and blended at that step's end (software pipeline): the XU
//     never sees a burst and its latency is covered by packed arithmetic of the same warp.
//   * persistent grid: one NT-thread CTA per SM walks query tiles of NT*4 queries.
#pragma once
#include "irbfn_common.cuh"
#include "irbfn_fwd.cuh"   // FwdShape; the plain dense kernel lives there, untouched

namespace irbfn {


// MODE 0: the plain dense forward.  MODE 1: gate-sparse (perm, masks, compact staging).  MODE 2: dense + the gate
// written out for a training step's backward.  Separate instantiations: the extra code of modes 1 / 2, even when
// skipped at run time, changed the register allocation of the wide-output dense variants (C5 forward 2.06 -> 2.33 s).
// MODE 3: dense, input rows streamed in while the kernel runs (irbfn_forward_streamed): a tile waits for its flag.
constexpr int kFwdDense = 0, kFwdSparse = 1, kFwdGateOut = 2, kFwdStream = 3;

template <int D, int OT, int BASIS, int NT, int MODE = kFwdDense>
__global__ void __launch_bounds__(NT, 1) irbfn_fwd_mode_kernel(const FwdArgs a) {
    constexpr bool SPARSE = MODE == kFwdSparse;
    constexpr bool GOUT = MODE == kFwdGateOut;
    constexpr bool STREAM = MODE == kFwdStream;
    constexpr int Q = FwdShape::Q, TQ = NT * Q, KT = FwdShape::KT;
    constexpr int OP = OT / 2;            // packed output pairs
    constexpr int OTS = OT > 0 ? OT : 1;
    constexpr int OPS = OP > 0 ? OP : 1;

    extern __shared__ __align__(16) float smem[];
    const int KCT = a.KC / KT;                                  // kernel tiles per region in a stage
    float4* stage4 = reinterpret_cast<float4*>(smem);           // [RC][KCT][D+1] float4
    float* Ws = smem + (size_t)a.RC * KCT * (D + 1) * 4;        // [KC][OT]
    float* gtab = Ws + (size_t)a.KC * (OT > 0 ? OT : 0);        // [RC][NT] float4
    float* stage = smem;

    // gate-sparse evaluation: region masks of the current region chunk (<= 64 regions per chunk)
    __shared__ unsigned long long s_cmask[3];   // union over the CTA's warps; three buffers in rotation so that
                                                // re-zeroing one never races with a later chunk's atomicOr
    __shared__ int s_slot2r[64];                // compact stage slot -> region of the chunk
    __shared__ unsigned s_tile;                 // STREAM: the tile this CTA fetched from the counter
    float gthr = 0.f;
    unsigned long long n_skip = 0, n_done = 0;
    int parity = 0;
    if (SPARSE) {
        gthr = __ldg(a.gate_thr);
        if (threadIdx.x < 3) s_cmask[threadIdx.x] = 0ull;
        __syncthreads();
    }

    const int tid = threadIdx.x;
    const int R = a.R, K = a.K, RC = a.RC, KC = a.KC;
    const int n_rch = (R + RC - 1) / RC;
    const int kb = blockIdx.y * a.k_per_split;              // this CTA's kernel range (split-K for small batches)
    const int ke = min(K, kb + a.k_per_split);
    const int n_kch = (ke - kb + KC - 1) / KC;
    float* const ybase = a.y != nullptr ? a.y + (size_t)blockIdx.y * a.split_stride : nullptr;
    const bool add_bias = blockIdx.y == 0;
    bool staged = false;

    // SPARSE with resident parameters: no barrier after the first staging, so every WARP schedules itself: its
    // first 128-query unit is static (all warps meet in the staging barrier), the rest come from a counter.
    const bool warp_sched = SPARSE && a.stage_once && a.work_counter != nullptr;
    const long long num_units = (a.B + 32 * Q - 1) / (32 * Q);
    long long iter = 0;
    for (long long tile = blockIdx.x;; tile += gridDim.x, ++iter) {
        long long q0;
        if (warp_sched) {
            long long u;
            if (iter == 0) {
                u = (long long)blockIdx.x * (NT / 32) + (tid >> 5);
            } else {
                unsigned v = 0;
                if ((tid & 31) == 0) v = atomicAdd(a.work_counter, 1u);
                v = __shfl_sync(0xffffffffu, v, 0);
                u = (long long)gridDim.x * (NT / 32) + v;
                if (u >= num_units) break;
            }
            q0 = u * (32 * Q) + (long long)(tid & 31) * Q;      // beyond B on the first round: an empty unit
        } else if (SPARSE && a.work_counter != nullptr) {
            // SPARSE with chunk-staged parameters: the staging barriers keep the CTA together, so the CTA fetches its
            // tiles from the counter (the first one is static).  Binned tiles differ several-fold in active regions; a
            // static round-robin left the SMs 83 % busy on average at C5 (sm__cycles_active.avg / elapsed.max).
            if (iter > 0) {
                __syncthreads();                               // everyone has read s_tile of the previous round
                if (tid == 0) s_tile = gridDim.x + atomicAdd(a.work_counter, 1u);
                __syncthreads();
                tile = s_tile;
            }
            if (tile >= a.num_tiles) break;
            q0 = tile * TQ + (long long)tid * Q;
        } else if (STREAM) {
            // Streamed input: the rows of a tile are uploaded WHILE earlier tiles compute.  Tiles are handed out by a
            // counter (CTAs whose first rows arrived early take more tiles than those that had to wait), and a CTA
            // waits for the flag of the tile's upload piece -- written behind the piece's data, in stream order --
            // before it touches the rows.  The rows are then read with COHERENT loads (ld.global.cg): ld.global.nc marks data as constant for
            // the kernel's lifetime and ptxas hoists such loads above the barrier that follows the wait.
            __syncthreads();                                   // everyone has consumed s_tile of the previous round
            if (tid == 0) {
                unsigned t = atomicAdd(a.work_counter, 1u);
                if (t < (unsigned)a.num_tiles) {
                    const long long t0 = clock64();
                    unsigned v;
                    const unsigned* f = a.ready + t / (unsigned)a.tiles_per_flag;   // the flag of the tile's upload piece
                    for (unsigned spin = 0;; ++spin) {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                        if (v != 0u) break;
                        // ~2 s: the upload never came.  The first CTA to find out raises the status word and every
                        // CTA gives the launch up when it sees it (the result is undefined from here on), so a
                        // stalled upload costs one budget, not one per tile.
                        const bool gave_up = a.stream_status != nullptr && (spin & 63u) == 63u &&
                                             *(volatile int*)a.stream_status != 0;
                        if (gave_up || clock64() - t0 > 4000000000ll) {
                            if (a.stream_status != nullptr) atomicExch(a.stream_status, 1);
                            t = 0xffffffffu;
                            break;
                        }
                        __nanosleep(128);
                    }
                }
                s_tile = t;
            }
            __syncthreads();
            tile = s_tile;
            if (tile >= a.num_tiles) break;
            q0 = tile * TQ + (long long)tid * Q;
        } else {
            if (tile >= a.num_tiles) break;
            q0 = tile * TQ + (long long)tid * Q;
        }

        // ---- queries -> registers, packed in pairs over queries ----
        float2 x2[2][D];
        long long qi[Q];                     // SPARSE: the rows this thread evaluates (permuted order), -1 = none
        if (SPARSE) {
#pragma unroll
            for (int i = 0; i < Q; ++i) qi[i] = q0 + i < a.B ? (long long)__ldg(a.perm + q0 + i) : -1ll;
#pragma unroll
            for (int p = 0; p < 2; ++p) {
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    x2[p][j].x = (qi[2 * p] >= 0 && j < a.Dsrc) ? __ldg(a.x + qi[2 * p] * a.Dsrc + j) : 0.f;
                    x2[p][j].y = (qi[2 * p + 1] >= 0 && j < a.Dsrc) ? __ldg(a.x + qi[2 * p + 1] * a.Dsrc + j) : 0.f;
                }
            }
        } else if (STREAM) {
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    x2[p][j].x = (qa < a.B && j < a.Dsrc) ? __ldcg(a.x + qa * a.Dsrc + j) : 0.f;
                    x2[p][j].y = (qb < a.B && j < a.Dsrc) ? __ldcg(a.x + qb * a.Dsrc + j) : 0.f;
                }
            }
        } else {
            load_queries4<D>(a.x, q0, a.B, a.Dsrc, x2);
        }

        // ---- output accumulators start at the bias ----
        float2 y2[Q][OPS];
        if (OT > 0) {
#pragma unroll
            for (int op = 0; op < OP; ++op) {
                float2 bv;
                bv.x = (add_bias && 2 * op < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op) : 0.f;
                bv.y = (add_bias && 2 * op + 1 < a.ocur) ? __ldg(a.bias + a.o0 + 2 * op + 1) : 0.f;
#pragma unroll
                for (int q = 0; q < Q; ++q) y2[q][op] = bv;
            }
        }

        for (int rch = 0; rch < n_rch; ++rch) {
            const int r0 = rch * RC;
            const int rc = min(RC, R - r0);
            float4* gcol = reinterpret_cast<float4*>(gtab) + tid;  // + rr * NT

            // ---- gate table for this region chunk: thread-private columns, no barrier ----
            if (a.gate_ext != nullptr) {
                for (int rr = 0; rr < rc; ++rr) {
                    float4 g;
                    g.x = (q0 + 0 < a.B) ? __ldg(a.gate_ext + (q0 + 0) * R + r0 + rr) : 0.f;
                    g.y = (q0 + 1 < a.B) ? __ldg(a.gate_ext + (q0 + 1) * R + r0 + rr) : 0.f;
                    g.z = (q0 + 2 < a.B) ? __ldg(a.gate_ext + (q0 + 2) * R + r0 + rr) : 0.f;
                    g.w = (q0 + 3 < a.B) ? __ldg(a.gate_ext + (q0 + 3) * R + r0 + rr) : 0.f;
                    gcol[rr * NT] = g;
                }
            } else if (SPARSE) {
                gate_columns_rows(a.gd, a.x, qi, a.Dsrc, r0, rc, gcol, NT);
            } else {
                if (STREAM) {
                    gate_columns_impl(a.gd, [&](int j) {
                        float4 v;
                        v.x = q0 + 0 < a.B ? __ldcg(a.x + (q0 + 0) * a.Dsrc + j) : 0.f;
                        v.y = q0 + 1 < a.B ? __ldcg(a.x + (q0 + 1) * a.Dsrc + j) : 0.f;
                        v.z = q0 + 2 < a.B ? __ldcg(a.x + (q0 + 2) * a.Dsrc + j) : 0.f;
                        v.w = q0 + 3 < a.B ? __ldcg(a.x + (q0 + 3) * a.Dsrc + j) : 0.f;
                        return v;
                    }, r0, rc, gcol, NT);
                } else
                gate_columns(a.gd, a.x, q0, a.B, a.Dsrc, r0, rc, gcol, NT);
            }
            // ---- SPARSE: which regions of the chunk matter to this warp / to the CTA ----
            unsigned long long wmask = 0ull, cmask = 0ull;
            if (SPARSE) {
                for (int rr = 0; rr < rc; ++rr) {
                    // the drop rule is applied PER QUERY (a gate at or below the threshold becomes 0), so a query's
                    // result does not depend on which other queries share its warp; the warp then skips the
                    // regions that are dropped for all of its queries
                    float4 g = gcol[rr * NT];
                    g.x = g.x > gthr ? g.x : 0.f;
                    g.y = g.y > gthr ? g.y : 0.f;
                    g.z = g.z > gthr ? g.z : 0.f;
                    g.w = g.w > gthr ? g.w : 0.f;
                    gcol[rr * NT] = g;
                    const bool act = fmaxf(fmaxf(g.x, g.y), fmaxf(g.z, g.w)) > 0.f;
                    if (__any_sync(0xffffffffu, act)) wmask |= 1ull << rr;
                }
                n_skip += (unsigned long long)(rc - __popcll(wmask));
                n_done += (unsigned long long)__popcll(wmask);
                if (!a.stage_once) {
                    // parameters staged per (tile, chunk): the CTA stages only the regions some warp needs
                    if ((tid & 31) == 0 && wmask) atomicOr(&s_cmask[parity], wmask);
                    __syncthreads();
                    cmask = s_cmask[parity];
                    // buffer of the chunk after next: last read before this barrier, next written after the next one
                    if (tid == 0) s_cmask[(parity + 2) % 3] = 0ull;
                    parity = (parity + 1) % 3;
                    if (cmask == 0ull) continue;            // CTA-uniform: nobody needs this region chunk
                    if (tid < 64 && ((cmask >> tid) & 1ull)) s_slot2r[__popcll(cmask & ((1ull << tid) - 1ull))] = tid;
                } else {
                    cmask = ~0ull;      // everything is resident: warps skip on their own, no barrier per tile
                }
            }
            const bool compact = SPARSE && !a.stage_once;     // stage only the CTA's active regions
            const int n_stage_r = compact ? __popcll(cmask) : rc;
            if (GOUT && a.gate_out != nullptr && blockIdx.y == 0) {
                // training: the gate goes out to (B,R) for the parameter-gradient kernel.  A warp owns 128
                // consecutive queries, i.e. 128 * rc consecutive floats of the (B,R) rows when rc == R: its lanes
                // copy them out of the warp's own table columns, coalesced (thread-private scattered 4-byte
                // stores cost 40 % of the kernel; a CTA-wide copy needs two barriers per tile).
                __syncwarp();
                const int w0 = (tid >> 5) * 32;                         // first thread of this warp
                const long long qw0 = (long long)tile * TQ + (long long)w0 * Q;
                const int nq = (int)max(0ll, min((long long)(32 * Q), a.B - qw0));
                for (int e = tid & 31; e < nq * rc; e += 32) {
                    const int q = e / rc, rr = e - q * rc;
                    a.gate_out[(qw0 + q) * R + r0 + rr] = gtab[((size_t)rr * NT + w0 + (q >> 2)) * 4 + (q & 3)];
                }
                __syncwarp();
            }

            for (int kch = 0; kch < n_kch; ++kch) {
                const int k0 = kb + kch * KC;
                const int kc = min(KC, ke - k0);
                const int kcp = ((kc + KT - 1) / KT) * KT;

                // ---- stage parameters of (region chunk, kernel chunk), transposed per 4-kernel tile ----
                if (!(a.stage_once && staged)) {
                    __syncthreads();  // everyone is done with the previous stage (and s_slot2r is visible)
                    for (int idx = tid; idx < n_stage_r * kcp; idx += NT) {
                        const int slot = idx / kcp, kk = idx - slot * kcp;
                        const int rr = compact ? s_slot2r[slot] : slot;
                        float* t = stage + ((size_t)slot * KCT + (kk >> 2)) * ((D + 1) * 4) + (kk & 3);
                        if (kk < kc) {
                            const size_t rk = (size_t)(r0 + rr) * K + (k0 + kk);
#pragma unroll
                            for (int j = 0; j < D; ++j)
                                t[4 * j] = j < a.Dsrc ? -__ldg(a.centers + rk * a.Dsrc + j) : 0.f;
                            t[4 * D] = width_scale<BASIS>(__ldg(a.log_sigs + rk));
                        } else {
#pragma unroll
                            for (int j = 0; j <= D; ++j) t[4 * j] = 0.f;
                        }
                    }
                    if (OT > 0) {
                        for (int idx = tid; idx < kcp * OTS; idx += NT) {
                            const int kk = idx / OTS, o = idx - kk * OTS;
                            Ws[idx] = (kk < kc && o < a.ocur)
                                          ? __ldg(a.W + (size_t)(k0 + kk) * a.ldw + a.o0 + o) : 0.f;
                        }
                    }
                    __syncthreads();
                    staged = true;
                }

                // ---- hot loop: (kernel tile, region) steps, blend pipelined by one step ----
                float2 h2[KT][2], ph[KT][2], ea[KT][2];   // blend accumulators, basis values, basis arguments
                float2 gp[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
                for (int kk = 0; kk < KT; ++kk) {
                    h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);
                    ea[kk][0] = ea[kk][1] = make_float2(1.f, 1.f);   // finite basis value, blended with gate 0
                }

                // y += h . W[kt .. kt+3, :] for the finished kernel tile at chunk offset kt
                auto readout = [&](int kt) {
                    if (OT > 0) {
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
                            const float2* w2 = reinterpret_cast<const float2*>(Ws + (size_t)(kt + kk) * OTS);
                            float2 wv[OPS];
#pragma unroll
                            for (int op = 0; op < OP; ++op) wv[op] = w2[op];
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const float2 hx = dup2(h2[kk][p].x), hy = dup2(h2[kk][p].y);
#pragma unroll
                                for (int op = 0; op < OP; ++op) {
                                    y2[2 * p][op] = fma2(hx, wv[op], y2[2 * p][op]);
                                    y2[2 * p + 1][op] = fma2(hy, wv[op], y2[2 * p + 1][op]);
                                }
                            }
                        }
                    }
                };
                auto store_hidden = [&](int kt) {
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        const int k = k0 + kt + kk;
                        if (k < K) {
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const long long qa = q0 + 2 * p, qb = qa + 1;
                                if (qa < a.B) {
                                    float* hp = a.hidden + qa * K + k;
                                    *hp = (rch > 0 ? *hp : 0.f) + h2[kk][p].x;
                                }
                                if (qb < a.B) {
                                    float* hp = a.hidden + qb * K + k;
                                    *hp = (rch > 0 ? *hp : 0.f) + h2[kk][p].y;
                                }
                            }
                        }
                    }
                };
                // basis function of the PREVIOUS step's arguments: MUFU number i of 16 (4 kernels x 4 queries)
                auto mufu = [&](int i) {
                    const int kk = i >> 2, p = (i >> 1) & 1;
                    float& src = (i & 1) ? ea[kk][p].y : ea[kk][p].x;
                    float& dst = (i & 1) ? ph[kk][p].y : ph[kk][p].x;
                    dst = BASIS == IRBFN_BASIS_GAUSSIAN ? ex2_approx(src) : rcp_approx(src);
                };
                // distances of one step: feature-outer, 8 independent packed chains; the 16 MUFUs of the
                // previous step are spread between the features so that the XU never sees a burst and the
                // issue gaps a MUFU leaves are filled with this step's packed arithmetic.
                auto distances = [&](const float4* rec, float2 (&acc)[KT][2]) {
#pragma unroll
                    for (int j = 0; j < D; ++j) {
                        const float4 c4 = rec[j];
                        const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                const float2 d = add2(x2[p][j], dup2(c[kk]));
                                acc[kk][p] = j == 0 ? mul2(d, d) : fma2(d, d, acc[kk][p]);
                            }
                        }
#pragma unroll
                        for (int i = (16 * j) / D; i < (16 * (j + 1)) / D; ++i) mufu(i);
                    }
                };
                // blend of the PREVIOUS step
                auto blend = [&]() {
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
                        h2[kk][0] = fma2(gp[0], ph[kk][0], h2[kk][0]);
                        h2[kk][1] = fma2(gp[1], ph[kk][1], h2[kk][1]);
                    }
                };
                // arguments of this step's basis function; evaluated during the next step's distances
                auto scale = [&](const float4* rec, const float2 (&acc)[KT][2], const float4 g4) {
                    const float4 s4 = rec[D];
                    const float sc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                    for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
                        for (int p = 0; p < 2; ++p) {
                            if (BASIS == IRBFN_BASIS_GAUSSIAN) ea[kk][p] = mul2(acc[kk][p], dup2(sc[kk]));
                            else ea[kk][p] = fma2(acc[kk][p], dup2(sc[kk]), make_float2(1.f, 1.f));
                        }
                    }
                    gp[0] = make_float2(g4.x, g4.y);
                    gp[1] = make_float2(g4.z, g4.w);
                };

                const size_t rstride = (size_t)KCT * (D + 1);
                if (SPARSE) {
                    // the same pipeline over the warp's active regions only (warp-uniform trip counts; the CTA's
                    // barriers all sit in the staging above, which every warp still joins)
                    if (wmask != 0ull) {
                        const int rr0 = __ffsll((long long)wmask) - 1;
                        const unsigned long long rest = wmask & (wmask - 1ull);
                        const size_t slot0 = compact ? (size_t)__popcll(cmask & ((1ull << rr0) - 1ull)) : (size_t)rr0;
                        for (int kt = 0; kt < kcp; kt += KT) {
                            const float4* rec0 = stage4 + (size_t)(kt >> 2) * (D + 1);
                            {   // first active region of the tile (peeled): also closes the previous kernel tile
                                const float4* rec = rec0 + slot0 * rstride;
                                const float4 g4 = gcol[rr0 * NT];
                                float2 acc[KT][2];
                                distances(rec, acc);
                                blend();
                                readout(kt > 0 ? kt - KT : 0);
                                scale(rec, acc, g4);
#pragma unroll
                                for (int kk = 0; kk < KT; ++kk) h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);
                            }
                            unsigned long long m = rest;
                            while (m) {   // branch-free steady state over the other active regions
                                const int rr = __ffsll((long long)m) - 1;
                                m &= m - 1ull;
                                const size_t slot = compact ? (size_t)__popcll(cmask & ((1ull << rr) - 1ull)) : (size_t)rr;
                                const float4* rec = rec0 + slot * rstride;
                                const float4 g4 = gcol[rr * NT];
                                float2 acc[KT][2];
                                distances(rec, acc);
                                blend();
                                scale(rec, acc, g4);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 16; ++i) mufu(i);
                        blend();
                        readout(kcp - KT);
                    }
                    continue;
                }
                for (int kt = 0; kt < kcp; kt += KT) {
                    const float4* rec = stage4 + (size_t)(kt >> 2) * (D + 1);
                    {   // first region of the tile (peeled): also closes the previous kernel tile.
                        // At kt == 0 the pipeline is empty (h and gp are zero): blend and readout add 0.
                        const float4 g4 = gcol[0];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        blend();
                        readout(kt > 0 ? kt - KT : 0);
                        scale(rec, acc, g4);
                        if (a.hidden != nullptr && kt > 0) store_hidden(kt - KT);
#pragma unroll
                        for (int kk = 0; kk < KT; ++kk) h2[kk][0] = h2[kk][1] = make_float2(0.f, 0.f);
                    }
                    for (int rr = 1; rr < rc; ++rr) {   // branch-free steady state
                        rec += rstride;
                        const float4 g4 = gcol[rr * NT];
                        float2 acc[KT][2];
                        distances(rec, acc);
                        blend();
                        scale(rec, acc, g4);
                    }
                }
                // drain the pipeline: basis + blend of the last step, close the last kernel tile
#pragma unroll
                for (int i = 0; i < 16; ++i) mufu(i);
                blend();
                readout(kcp - KT);
                if (a.hidden != nullptr) store_hidden(kcp - KT);
            }
        }

        // ---- store y ----
        // STREAM: y may be pinned host memory.  A thread's four rows are 4 * O consecutive floats, 16-byte aligned when
        // the launch covers whole rows: written as float4s, so that every PCIe write carries full 32-byte sectors
        // (scalar 4-byte stores to host memory cost the kernel +600 us per 2^20 queries at O = 2).
        if (STREAM && OT > 0 && a.ocur == OT && a.ldw == OT && a.o0 == 0 && q0 + Q - 1 < a.B) {
            float4* yp4 = reinterpret_cast<float4*>(ybase + q0 * OT);
            float v[Q * OTS];
#pragma unroll
            for (int q = 0; q < Q; ++q)
#pragma unroll
                for (int op = 0; op < OP; ++op) { v[q * OT + 2 * op] = y2[q][op].x; v[q * OT + 2 * op + 1] = y2[q][op].y; }
#pragma unroll
            for (int i = 0; i < OT; ++i) yp4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        } else
        if (OT > 0) {
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const long long qq = SPARSE ? qi[q] : q0 + q;
                if (SPARSE ? qq >= 0 : qq < a.B) {
                    float* yp = ybase + qq * a.ldw + a.o0;
#pragma unroll
                    for (int op = 0; op < OP; ++op) {
                        if (2 * op < a.ocur) yp[2 * op] = y2[q][op].x;
                        if (2 * op + 1 < a.ocur) yp[2 * op + 1] = y2[q][op].y;
                    }
                }
            }
        }
    }
    if (SPARSE && a.skip_count != nullptr && (tid & 31) == 0) {
        atomicAdd(a.skip_count, n_skip);
        atomicAdd(a.skip_count + 1, n_done);
    }
}

// Host-side: pick the instantiation for (OT, basis, threads) and launch.  Returns a cudaError_t as int.
template <int D, int OT, int BASIS, int NT, int MODE = kFwdDense>
inline int launch_fwdm_inst(const FwdArgs& a, int grid, size_t smem, cudaStream_t st) {
    auto kern = irbfn_fwd_mode_kernel<D, OT, BASIS, NT, MODE>;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<dim3(grid, a.nsplit > 0 ? a.nsplit : 1), NT, smem, st>>>(a);
    return (int)cudaGetLastError();
}

template <int D, int OT>
inline int launch_fwdm_ot(const FwdArgs& a, int basis, int nt, int grid, size_t smem, cudaStream_t st) {
    if (nt != 256) return (int)cudaErrorInvalidValue;   // the one compiled CTA size (fwd_threads_for)
    if (a.perm != nullptr) {                            // gate-sparse instantiation
        if (OT == 0) return (int)cudaErrorInvalidValue;
        if (basis == IRBFN_BASIS_GAUSSIAN) return launch_fwdm_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 256, kFwdSparse>(a, grid, smem, st);
        return launch_fwdm_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 256, kFwdSparse>(a, grid, smem, st);
    }
    if (a.ready != nullptr) {                           // streamed input
        if (OT == 0) return (int)cudaErrorInvalidValue;
        if (basis == IRBFN_BASIS_GAUSSIAN) return launch_fwdm_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 256, kFwdStream>(a, grid, smem, st);
        return launch_fwdm_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 256, kFwdStream>(a, grid, smem, st);
    }
    if (a.gate_out != nullptr) {                        // training forward: also writes the gate
        if (OT == 0) return (int)cudaErrorInvalidValue;
        if (basis == IRBFN_BASIS_GAUSSIAN) return launch_fwdm_inst<D, OT, IRBFN_BASIS_GAUSSIAN, 256, kFwdGateOut>(a, grid, smem, st);
        return launch_fwdm_inst<D, OT, IRBFN_BASIS_INVERSE_QUADRATIC, 256, kFwdGateOut>(a, grid, smem, st);
    }
    return (int)cudaErrorInvalidValue;                  // the plain dense forward is irbfn_fwd_kernel (irbfn_fwd.cuh)
}

template <int D>
inline int launch_fwdm(const FwdArgs& a, int OT, int basis, int nt, int grid, size_t smem, cudaStream_t st) {
    switch (OT) {
        case 0: return (int)cudaErrorInvalidValue;      // both modes produce y
        case 2: return launch_fwdm_ot<D, 2>(a, basis, nt, grid, smem, st);
        case 4: return launch_fwdm_ot<D, 4>(a, basis, nt, grid, smem, st);
        case 8: return launch_fwdm_ot<D, 8>(a, basis, nt, grid, smem, st);
        case 10: return launch_fwdm_ot<D, 10>(a, basis, nt, grid, smem, st);
        case 16: return launch_fwdm_ot<D, 16>(a, basis, nt, grid, smem, st);
        default: return (int)cudaErrorInvalidValue;
    }
}

}  // namespace irbfn
