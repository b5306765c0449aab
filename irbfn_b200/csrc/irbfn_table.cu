// Table-side kernels of the IRBFN workflow (sm_100a) -- SURVEY.md 8f ranks 3 and 4.
//
//   irbfn_gather_rows   the training input pipeline on the device: batch_x = X[perm], batch_y = Y[perm]
//                       (scripts/train_nmpc_frenet.py:499-501 does this gather in NumPy on the host and
//                       copies the batch to the device every step)
//   irbfn_nearest_row   brute-force nearest table row per query, jnp.argmin(jnp.linalg.norm(all - q, axis=1))
//                       (src/irbfn_mpc/explicit_planner.py:21-25; irbfn_vs_explicit_frenet.ipynb cell 11)
//   irbfn_grid_index    per-dimension np.searchsorted(keys_d, v, side="right") clamped to shape-1
//                       (src/irbfn_mpc/explicit_planner.py:165-175)
//
// gather and grid_index are HBM-bound index work (bit-exact).  nearest_row has the RBF distance's
// shape -- (B,N,D) subtract / FMA -- and is FP32-pipe bound: it reuses the forward's layout (four
// queries per thread as two packed pairs, table rows broadcast from transposed 4-row shared-memory
// tiles) with a running (min, argmin) instead of the basis function.
#include "irbfn_common.cuh"

namespace irbfn {
int device_info(int* sms, int* smem_optin);   // irbfn_api.cu

namespace {

// One thread per output float; a row of W floats is read by W consecutive threads (one 32-byte sector
// for the 8-feature state tables), written fully coalesced.
__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ X, const float* __restrict__ Y,
                                                          const long long* __restrict__ idx, float* __restrict__ xb,
                                                          float* __restrict__ yb, long long n, long long N, int D,
                                                          int O, int* __restrict__ bad) {
    const int W = D + O;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * W) return;
    const long long i = t / W;
    const int c = (int)(t - i * W);
    const long long r = __ldg(idx + i);
    if (r < 0 || r >= N) {                      // never read out of bounds; flag it for the host
        if (c == 0) atomicExch(bad, 1);
        if (c < D) xb[i * D + c] = 0.f;
        else yb[i * O + (c - D)] = 0.f;
        return;
    }
    if (c < D) xb[i * D + c] = __ldg(X + r * D + c);
    else yb[i * O + (c - D)] = __ldg(Y + r * O + (c - D));
}

// searchsorted(keys_d, v, side="right") = number of keys <= v, then min(len_d - 1, .)
// T = float (tables narrowed to fp32) or double (the reference's float64 tables and values: index-exact).
template <typename T>
__global__ void __launch_bounds__(256) grid_index_kernel(const T* __restrict__ q, const T* __restrict__ keys,
                                                         const int* __restrict__ offset, int* __restrict__ out,
                                                         long long B, int nd) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= B * nd) return;
    const int d = (int)(t % nd);
    const T v = __ldg(q + t);
    const int lo0 = __ldg(offset + d), hi0 = __ldg(offset + d + 1);
    int lo = lo0, hi = hi0;                     // first position with key > v
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(keys + mid) <= v) lo = mid + 1;
        else hi = mid;
    }
    const int len = hi0 - lo0;
    const int pos = v != v ? len : lo - lo0;     // numpy orders NaN after every key
    out[t] = pos < len - 1 ? pos : len - 1;
}


// ---------------------------------------------------------------------------------------------
// nearest table row.  grid = (query tiles, table splits); every CTA scans its slice of the table in
// chunks of TR rows staged (negated, transposed in tiles of four rows) through shared memory.
struct NearestArgs {
    const float* q;        // (B,Dsrc)
    const float* table;    // (N,Dsrc)
    long long* idx;        // (B) or partial [nsplit][B]
    float* dist2;          // (B) or partial [nsplit][B]; squared distance of the winner
    long long B, N;
    long long rows_per_split;   // multiple of 4
    int Dsrc;
    int num_tiles;
    int TR;                // rows per shared-memory chunk (multiple of 4)
};

constexpr int kNearNT = 256, kNearQ = 4;

template <int D>
__global__ void __launch_bounds__(kNearNT, 1) nearest_row_kernel(const NearestArgs a) {
    extern __shared__ __align__(16) float smem[];          // [TR/4][D] float4: row j of a tile = -t_j of 4 rows
    float4* tile4 = reinterpret_cast<float4*>(smem);
    const int tid = threadIdx.x;
    const long long n_lo = (long long)blockIdx.y * a.rows_per_split;
    const long long n_hi = min(a.N, n_lo + a.rows_per_split);
    const float kInf = __int_as_float(0x7f800000);

    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        const long long q0 = (long long)tile * (kNearNT * kNearQ) + (long long)tid * kNearQ;
        float2 x2[2][D];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const long long qa = q0 + 2 * p, qb = qa + 1;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                x2[p][j].x = (qa < a.B && j < a.Dsrc) ? __ldg(a.q + qa * a.Dsrc + j) : 0.f;
                x2[p][j].y = (qb < a.B && j < a.Dsrc) ? __ldg(a.q + qb * a.Dsrc + j) : 0.f;
            }
        }
        float best[4] = {kInf, kInf, kInf, kInf};
        long long bidx[4] = {n_lo, n_lo, n_lo, n_lo};

        for (long long c0 = n_lo; c0 < n_hi; c0 += a.TR) {
            const int rows = (int)min((long long)a.TR, n_hi - c0);
            const int rows4 = (rows + 3) & ~3;
            __syncthreads();
            for (int i = tid; i < rows4 * D; i += kNearNT) {       // coalesced read, transposed write
                const int r = i / D, j = i - r * D;
                // padding rows sit at +inf distance: -t = inf makes (x - t)^2 = inf, never the minimum
                const float v = r < rows ? (j < a.Dsrc ? -__ldg(a.table + (c0 + r) * a.Dsrc + j) : 0.f) : kInf;
                smem[((r >> 2) * D + j) * 4 + (r & 3)] = v;
            }
            __syncthreads();
            const float4* rec = tile4;
            for (int r = 0; r < rows4; r += 4, rec += D) {
                float2 acc[4][2];
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    const float4 c4 = rec[j];
                    const float c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
                        for (int p = 0; p < 2; ++p) {
                            const float2 d = add2(x2[p][j], dup2(c[kk]));
                            acc[kk][p] = j == 0 ? mul2(d, d) : fma2(d, d, acc[kk][p]);
                        }
                    }
                }
                // running (min, argmin): a min-tree over the tile's four rows and ONE rarely-taken branch per
                // tile keep the compare/select chain off the ALU pipe (a query's minimum improves ~ln N times
                // in N rows).  Rows are visited in increasing order and the update is a strict <, so fp32 ties
                // keep the lowest index, like argmin.
                float m[4];
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    m[2 * p] = fminf(fminf(acc[0][p].x, acc[1][p].x), fminf(acc[2][p].x, acc[3][p].x));
                    m[2 * p + 1] = fminf(fminf(acc[0][p].y, acc[1][p].y), fminf(acc[2][p].y, acc[3][p].y));
                }
                if ((m[0] < best[0]) | (m[1] < best[1]) | (m[2] < best[2]) | (m[3] < best[3])) {
                    const long long base = c0 + r;
#pragma unroll
                    for (int p = 0; p < 2; ++p) {
#pragma unroll
                        for (int kk = 3; kk >= 0; --kk) {          // descending: the first row that attains m wins
                            if (acc[kk][p].x == m[2 * p] && m[2 * p] < best[2 * p]) bidx[2 * p] = base + kk;
                            if (acc[kk][p].y == m[2 * p + 1] && m[2 * p + 1] < best[2 * p + 1]) bidx[2 * p + 1] = base + kk;
                        }
                        best[2 * p] = fminf(best[2 * p], m[2 * p]);
                        best[2 * p + 1] = fminf(best[2 * p + 1], m[2 * p + 1]);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (q0 + u < a.B) {
                a.idx[(size_t)blockIdx.y * a.B + q0 + u] = bidx[u];
                a.dist2[(size_t)blockIdx.y * a.B + q0 + u] = best[u];
            }
        }
    }
}

// winner over the table splits, in split order with strict <: the lowest row index wins ties
__global__ void __launch_bounds__(256) nearest_reduce_kernel(const long long* __restrict__ pidx,
                                                             const float* __restrict__ pd, long long* __restrict__ idx,
                                                             float* __restrict__ dist, int nsplit, long long B) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float best = pd[b];
    long long bi = pidx[b];
    for (int s = 1; s < nsplit; ++s) {
        const float v = pd[(size_t)s * B + b];
        if (v < best) { best = v; bi = pidx[(size_t)s * B + b]; }
    }
    idx[b] = bi;
    if (dist) dist[b] = sqrtf(best);
}

__global__ void __launch_bounds__(256) sqrt_inplace_kernel(float* d, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = sqrtf(d[i]);
}

template <int D>
int launch_nearest(const NearestArgs& a, dim3 grid, size_t smem, cudaStream_t st) {
    auto kern = nearest_row_kernel<D>;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, kNearNT, smem, st>>>(a);
    return (int)cudaGetLastError();
}

// table splits for batches too small to give every SM a query tile
int nearest_split(long long B, long long N, int sms, long long* rows_per_split) {
    const long long tiles = (B + kNearNT * kNearQ - 1) / (kNearNT * kNearQ);
    long long ns = 1;
    if (tiles > 0 && 2 * tiles <= sms) ns = (sms + tiles - 1) / tiles;
    long long rps = (N + ns - 1) / ns;
    rps = ((rps + 3) / 4) * 4;
    if (rps < 256) rps = 256;
    *rows_per_split = rps;
    return (int)((N + rps - 1) / rps);
}

}  // namespace
}  // namespace irbfn

using namespace irbfn;

extern "C" int irbfn_gather_rows(const float* X, const float* Y, const int64_t* index, float* x_out, float* y_out,
                                 int64_t n, int64_t N, int D, int O, int* bad_index_flag, irbfn_stream_t stream) {
    if (n < 0 || N <= 0 || D <= 0 || O < 0) return IRBFN_ERR_BAD_DIMS;
    if (n == 0) return IRBFN_OK;
    if (!X || !index || !x_out || !bad_index_flag || (O > 0 && (!Y || !y_out))) return IRBFN_ERR_NULL_POINTER;
    const long long total = (long long)n * (D + O);
    gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        X, Y, (const long long*)index, x_out, y_out, n, N, D, O, bad_index_flag);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? IRBFN_OK : note_cuda_error("gather_rows_kernel", (int)e);
}

template <typename T>
static int grid_index_impl(const T* q, const T* keys, const int32_t* offset, int32_t* index, int64_t B, int num_dims,
                           irbfn_stream_t stream) {
    if (B < 0 || num_dims <= 0) return IRBFN_ERR_BAD_DIMS;
    if (B == 0) return IRBFN_OK;
    if (!q || !keys || !offset || !index) return IRBFN_ERR_NULL_POINTER;
    const long long total = (long long)B * num_dims;
    grid_index_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q, keys, offset, index, B,
                                                                                           num_dims);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? IRBFN_OK : note_cuda_error("grid_index_kernel", (int)e);
}

extern "C" int irbfn_grid_index(const float* q, const float* keys, const int32_t* offset, int32_t* index,
                                int64_t B, int num_dims, irbfn_stream_t stream) {
    return grid_index_impl<float>(q, keys, offset, index, B, num_dims, stream);
}

extern "C" int irbfn_grid_index_f64(const double* q, const double* keys, const int32_t* offset, int32_t* index,
                                    int64_t B, int num_dims, irbfn_stream_t stream) {
    return grid_index_impl<double>(q, keys, offset, index, B, num_dims, stream);
}

extern "C" size_t irbfn_nearest_row_workspace_bytes(int64_t B, int64_t N) {
    if (B <= 0 || N <= 0) return 0;
    int sms = 148, smem = 0;
    device_info(&sms, &smem);
    long long rps;
    const int ns = nearest_split(B, N, sms, &rps);
    // partial (index, squared distance) per split; one split still needs the squared-distance row
    return (size_t)ns * (size_t)B * (sizeof(long long) + sizeof(float));
}

extern "C" int irbfn_nearest_row(const float* q, const float* table, int64_t* index, float* distance, int64_t B,
                                 int64_t N, int D, void* workspace, size_t workspace_bytes, irbfn_stream_t stream) {
    if (B < 0 || N <= 0 || D <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (B == 0) return IRBFN_OK;
    if (!q || !table || !index) return IRBFN_ERR_NULL_POINTER;
    int sms = 0, smem_limit = 0;
    int rc = device_info(&sms, &smem_limit);
    if (rc != IRBFN_OK) return rc;
    if (!workspace || workspace_bytes < irbfn_nearest_row_workspace_bytes(B, N)) return IRBFN_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    NearestArgs a{};
    const int ns = nearest_split(B, N, sms, &a.rows_per_split);
    long long* pidx = (long long*)workspace;
    float* pd = (float*)(pidx + (size_t)ns * B);
    a.q = q; a.table = table; a.B = B; a.N = N; a.Dsrc = D;
    a.idx = ns > 1 ? pidx : (long long*)index;
    a.dist2 = pd;
    const int DT = D < 4 ? 4 : D;
    a.TR = 2048;                                             // 64 KB of shared memory at D = 8
    if ((long long)a.TR > a.rows_per_split) a.TR = (int)a.rows_per_split;
    const size_t smem = (size_t)a.TR * DT * sizeof(float);
    const long long tiles = (B + kNearNT * kNearQ - 1) / (kNearNT * kNearQ);
    a.num_tiles = (int)tiles;
    const dim3 grid((unsigned)(tiles < sms ? tiles : sms), (unsigned)ns);
    int e;
    switch (DT) {
        case 4: e = launch_nearest<4>(a, grid, smem, st); break;
        case 5: e = launch_nearest<5>(a, grid, smem, st); break;
        case 6: e = launch_nearest<6>(a, grid, smem, st); break;
        case 7: e = launch_nearest<7>(a, grid, smem, st); break;
        default: e = launch_nearest<8>(a, grid, smem, st); break;
    }
    if (e != 0) return note_cuda_error("nearest_row_kernel", e);
    if (ns > 1) {
        nearest_reduce_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(pidx, pd, (long long*)index, distance, ns, B);
    } else if (distance) {
        cudaMemcpyAsync(distance, pd, (size_t)B * sizeof(float), cudaMemcpyDeviceToDevice, st);
        sqrt_inplace_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(distance, B);
    }
    e = (int)cudaGetLastError();
    return e != 0 ? note_cuda_error("nearest_row reduce", e) : IRBFN_OK;
}


// ---------------------------------------------------------------------------------------------------------
// Lattice queries on the device: the meshgrid of irbfn_vs_explicit_frenet.ipynb cell 8 (np.meshgrid(...,
// indexing="ij") over some columns, the others held) written straight into HBM.  A 2^20-point sweep then
// uploads a few hundred bytes of axis values instead of 33.5 MB of queries -- on an 8-GPU node the host link
// (measured ~190 GB/s aggregate) is what bounds the host-buffer path.  HBM-bound: 4 D bytes per query.
namespace irbfn {
namespace {
struct LatticeArgs {
    const float* values;     // all axes' values concatenated
    const int* offset;       // (naxes + 1) prefix sums into values
    const int* column;       // (naxes) query column of each axis; the FIRST axis varies slowest
    const float* fixed;      // (D) values of the columns no axis covers
    float* x;                // (n, D)
    long long n;
    int naxes, D;
};
__global__ void __launch_bounds__(256) lattice_kernel(const LatticeArgs a) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < a.n; q += stride) {
        float row[IRBFN_MAX_D];
#pragma unroll
        for (int j = 0; j < IRBFN_MAX_D; ++j) row[j] = j < a.D ? __ldg(a.fixed + j) : 0.f;
        long long rem = q;
        for (int i = a.naxes - 1; i >= 0; --i) {          // last axis varies fastest ("ij" order, C-contiguous)
            const int lo = __ldg(a.offset + i), len = __ldg(a.offset + i + 1) - lo;
            const int idx = (int)(rem % len);
            rem /= len;
            const float v = __ldg(a.values + lo + idx);
            const int c = __ldg(a.column + i);
#pragma unroll
            for (int j = 0; j < IRBFN_MAX_D; ++j)
                if (j == c) row[j] = v;
        }
#pragma unroll
        for (int j = 0; j < IRBFN_MAX_D; ++j)
            if (j < a.D) a.x[q * a.D + j] = row[j];
    }
}
}  // namespace
}  // namespace irbfn

extern "C" int irbfn_lattice_queries(const float* values, const int32_t* offset, const int32_t* column, int num_axes,
                                     const float* fixed, float* x, int64_t n, int D, irbfn_stream_t stream) {
    if (n < 0 || D <= 0 || num_axes < 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (n == 0) return IRBFN_OK;
    if (!fixed || !x || (num_axes > 0 && (!values || !offset || !column))) return IRBFN_ERR_NULL_POINTER;
    irbfn::LatticeArgs a{values, offset, column, fixed, x, n, num_axes, D};
    const long long nb = (n + 255) / 256;
    irbfn::lattice_kernel<<<(unsigned)(nb < 148 * 16 ? nb : 148 * 16), 256, 0, (cudaStream_t)stream>>>(a);
    const int e = (int)cudaGetLastError();
    return e != 0 ? irbfn::note_cuda_error("lattice_kernel", e) : IRBFN_OK;
}
