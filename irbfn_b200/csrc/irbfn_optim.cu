// Optimiser update on the flat parameter buffer (sm_100a): optax.chain(clip_by_global_norm(c), adam(lr))
// as the reference builds it (scripts/train_nmpc.py:231-236, applied at :299 by state.apply_gradients), in
// two launches instead of ~12 element-wise library launches -- at 8 GPUs the C3 step holds ~0.15 ms of
// kernels per rank, so a dozen 3 us launches were a tenth of it.
//
//   g_norm = sqrt(sum g^2)                       optax.global_norm
//   g      = g_norm < c ? g : (g / g_norm) * c   clip_by_global_norm (lax.select on the trigger)
//   mu     = b1 mu + (1 - b1) g ;  nu = b2 nu + (1 - b2) g^2           scale_by_adam (eps_root = 0)
//   p     -= lr * (mu / (1 - b1^t)) / (sqrt(nu / (1 - b2^t)) + eps)    t = step count after increment
//
// Kernel 1 writes per-block partial sums of g^2 (fixed slabs, fixed tree: deterministic) and advances the
// device-side step count; kernel 2 re-adds the partials in order in every block and applies the update.
// The step count lives on the device so that the pair can sit inside a CUDA graph.
#include "irbfn_common.cuh"

namespace irbfn {
namespace {

constexpr int kOptThreads = 256;
constexpr int kOptMaxBlocks = 148;

int opt_blocks(long long n) {
    long long nb = (n + 4 * kOptThreads - 1) / (4 * kOptThreads);
    return (int)(nb < 1 ? 1 : (nb > kOptMaxBlocks ? kOptMaxBlocks : nb));
}

__global__ void __launch_bounds__(kOptThreads) sumsq_kernel(const float* __restrict__ g, long long n,
                                                            double* __restrict__ part, double* step_count) {
    __shared__ double red[kOptThreads];
    const long long per = (n + gridDim.x - 1) / gridDim.x;
    const long long lo = (long long)blockIdx.x * per;
    const long long hi = lo + per < n ? lo + per : n;
    double s = 0.0;
    for (long long i = lo + threadIdx.x; i < hi; i += kOptThreads) {
        const double v = (double)g[i];
        s += v * v;
    }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int h = kOptThreads / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        part[blockIdx.x] = red[0];
        if (blockIdx.x == 0) *step_count += 1.0;
    }
}

struct AdamArgs {
    float* p;
    float* g;
    float* mu;
    float* nu;
    const double* part;
    const double* step_count;
    float* gnorm_out;
    long long n;
    int nparts;
    double b1d, b2d;                    // for the bias corrections 1 - b^t
    float lr, b1, b2, omb1, omb2, eps, max_norm;   // each rounded from the double hyper-parameter, as JAX
};                                      // rounds a weak-typed Python scalar: (1 - b2) is float32(0.001), not 1 - float32(0.999)

__global__ void __launch_bounds__(kOptThreads) clip_adam_kernel(const AdamArgs a) {
    __shared__ float sh[4];   // clipping active, g_norm, 1 - b1^t, 1 - b2^t
    if (threadIdx.x == 0) {
        double ss = 0.0;
        for (int i = 0; i < a.nparts; ++i) ss += a.part[i];
        const float gnf = (float)sqrt(ss);
        sh[0] = (a.max_norm > 0.f && !(gnf < a.max_norm)) ? 1.f : 0.f;   // optax: trigger = g_norm < max_norm
        sh[1] = gnf;
        const double t = *a.step_count;
        sh[2] = (float)(1.0 - pow(a.b1d, t));
        sh[3] = (float)(1.0 - pow(a.b2d, t));
        if (blockIdx.x == 0 && a.gnorm_out != nullptr) *a.gnorm_out = gnf;
    }
    __syncthreads();
    const bool clip = sh[0] != 0.f;
    const float gn = sh[1], c1 = sh[2], c2 = sh[3];
    const float b1 = a.b1, b2 = a.b2;
    for (long long i = (long long)blockIdx.x * kOptThreads + threadIdx.x; i < a.n;
         i += (long long)gridDim.x * kOptThreads) {
        float g = a.g[i];
        if (clip) g = (g / gn) * a.max_norm;                       // select(trigger, t, (t / g_norm) * max_norm)
        const float m = a.omb1 * g + b1 * a.mu[i];                 // update_moment
        const float v = a.omb2 * (g * g) + b2 * a.nu[i];
        a.g[i] = g;
        a.mu[i] = m;
        a.nu[i] = v;
        a.p[i] += -a.lr * ((m / c1) / (sqrtf(v / c2) + a.eps));    // bias_correction, scale_by_adam, scale(-lr)
    }
}

// ---------------------------------------------------------------------------------------------------------
// One launch for the tail of a data-parallel step: [all-reduce over NVLink peer memory] -> global norm ->
// clip -> Adam.  Replaces ncclAllReduce + sumsq_kernel + clip_adam_kernel when the gradient is small
// (n <= kFusedMaxN: the shipped NMPC shapes are 10-20 k floats, where all three are pure latency).
//
// One CTA of 1,024 threads per rank.  Every rank's backward has written its local gradient (plus the loss
// scalar riding behind it) into ITS symmetric buffer; `peers[r]` is rank r's buffer mapped into this
// process (torch symmetric memory; peers[rank] is the local one), laid out
//     [ n + n_extra floats | ready[world] | done[world] | seq ]   (the flags are uint32 words).
//   1. seq = ++local seq word (device-side, so the launch can sit in a CUDA graph);
//      publish ready[rank] = seq in every peer's buffer (release, system scope);
//   2. wait until every peer published ready >= seq in MY buffer (acquire);
//   3. reduced[i] = sum_r peers[r][i] in rank order -- the same order on every rank, so the replicated
//      parameters stay bit-identical -- written to the LOCAL `reduced` buffer (peers still read mine);
//   4. publish done[rank] = seq to every peer, run the optimiser update on the local replica meanwhile, and
//      leave only when every peer published done >= seq: stream order then guarantees that the next step's
//      backward does not overwrite a buffer a peer is still reading.  No second buffer, no host sync.
// A rank that waits longer than ~20 s sets *status = 1 and gives up (a peer died); the caller checks it.
constexpr int kFusedThreads = 1024;
constexpr long long kFusedMaxN = 1 << 17;

struct FusedArgs {
    float* p;
    float* mu;
    float* nu;
    float* reduced;              // [n + n_extra] local: the all-reduced gradient (clipped in place by the update)
    float* const* peers;         // device array of world pointers, or nullptr (single process)
    double* step_count;
    float* gnorm_out;
    int* status;
    long long n;
    int n_extra, rank, world, apply_update;
    double b1d, b2d;
    float lr, b1, b2, omb1, omb2, eps, max_norm;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {   // peer memory: never from a stale L1 line
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_peer1(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(kFusedThreads) fused_update_kernel(const FusedArgs a) {
    __shared__ double red[kFusedThreads / 32];
    __shared__ float sh[4];
    __shared__ unsigned sh_seq;
    __shared__ int sh_fail;
    const int tid = threadIdx.x;
    const long long nt = a.n + a.n_extra;
    const long long flag_off = ((nt + 3) / 4) * 4;                       // uint32 words after the padded data
    unsigned seq = 0;
    if (a.peers != nullptr) {
        unsigned* mine = reinterpret_cast<unsigned*>(a.peers[a.rank]) + flag_off;
        if (tid == 0) {
            sh_seq = mine[2 * a.world] + 1u;
            mine[2 * a.world] = sh_seq;
            sh_fail = 0;
        }
        __syncthreads();
        seq = sh_seq;
        // 1. my gradient (written by the previous kernel of this stream) is ready for every peer
        if (tid < a.world) {
            __threadfence_system();
            st_release_sys(reinterpret_cast<unsigned*>(a.peers[tid]) + flag_off + a.rank, seq);
        }
        // 2. wait for every peer's gradient
        if (tid < a.world) {
            const long long t0 = clock64();
            while ((int)(ld_acquire_sys(mine + tid) - seq) < 0) {
                if (clock64() - t0 > 40000000000ll) { sh_fail = 1; break; }
            }
        }
        __syncthreads();
        if (sh_fail) { if (tid == 0 && a.status) *a.status = 1; return; }
        // 3. sum in rank order (float4 where the buffers allow it)
        const long long n4 = nt / 4;
        for (long long i = tid; i < n4; i += kFusedThreads) {
            float4 s = ld_peer4(a.peers[0] + 4 * i);
            for (int r = 1; r < a.world; ++r) {
                const float4 v = ld_peer4(a.peers[r] + 4 * i);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            reinterpret_cast<float4*>(a.reduced)[i] = s;
        }
        for (long long i = 4 * n4 + tid; i < nt; i += kFusedThreads) {
            float s = ld_peer1(a.peers[0] + i);
            for (int r = 1; r < a.world; ++r) s += ld_peer1(a.peers[r] + i);
            a.reduced[i] = s;
        }
        __syncthreads();
        // 4. I am done reading: tell every peer (the update below hides the round trip)
        if (tid < a.world) {
            __threadfence_system();
            st_release_sys(reinterpret_cast<unsigned*>(a.peers[tid]) + flag_off + a.world + a.rank, seq);
        }
    }

    if (a.apply_update) {
        // global norm over the first n floats (fixed per-thread strides, fixed tree: deterministic); four
        // independent loads in flight per thread -- one CTA is latency-bound, not bandwidth-bound
        constexpr long long T = kFusedThreads;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        long long i = tid;
        for (; i + 3 * T < a.n; i += 4 * T) {
            const float v0 = a.reduced[i], v1 = a.reduced[i + T], v2 = a.reduced[i + 2 * T], v3 = a.reduced[i + 3 * T];
            s0 += (double)v0 * v0; s1 += (double)v1 * v1; s2 += (double)v2 * v2; s3 += (double)v3 * v3;
        }
        for (; i < a.n; i += T) { const float v = a.reduced[i]; s0 += (double)v * v; }
        double s = (s0 + s1) + (s2 + s3);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if ((tid & 31) == 0) red[tid >> 5] = s;
        __syncthreads();
        if (tid == 0) {
            double ss = 0.0;
            for (int w = 0; w < kFusedThreads / 32; ++w) ss += red[w];
            const float gnf = (float)sqrt(ss);
            *a.step_count += 1.0;
            const double t = *a.step_count;
            sh[0] = (a.max_norm > 0.f && !(gnf < a.max_norm)) ? 1.f : 0.f;
            sh[1] = gnf;
            sh[2] = (float)(1.0 - pow(a.b1d, t));
            sh[3] = (float)(1.0 - pow(a.b2d, t));
            if (a.gnorm_out != nullptr) *a.gnorm_out = gnf;
        }
        __syncthreads();
        const bool clip = sh[0] != 0.f;
        const float gn = sh[1], c1 = sh[2], c2 = sh[3];
        auto upd = [&](float g, float mu, float nu, float p, long long at) {
            if (clip) g = (g / gn) * a.max_norm;
            const float m = a.omb1 * g + a.b1 * mu;
            const float v = a.omb2 * (g * g) + a.b2 * nu;
            a.reduced[at] = g;
            a.mu[at] = m;
            a.nu[at] = v;
            a.p[at] = p + -a.lr * ((m / c1) / (sqrtf(v / c2) + a.eps));
        };
        i = tid;
        for (; i + 3 * T < a.n; i += 4 * T) {
            float g[4], mu[4], nu[4], p[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                g[u] = a.reduced[i + u * T]; mu[u] = a.mu[i + u * T]; nu[u] = a.nu[i + u * T]; p[u] = a.p[i + u * T];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) upd(g[u], mu[u], nu[u], p[u], i + u * T);
        }
        for (; i < a.n; i += T) upd(a.reduced[i], a.mu[i], a.nu[i], a.p[i], i);
    }

    if (a.peers != nullptr) {
        unsigned* mine = reinterpret_cast<unsigned*>(a.peers[a.rank]) + flag_off;
        if (tid < a.world) {
            const long long t0 = clock64();
            while ((int)(ld_acquire_sys(mine + a.world + tid) - seq) < 0) {
                if (clock64() - t0 > 40000000000ll) { if (a.status) *a.status = 1; break; }
            }
        }
    }
}

}  // namespace
}  // namespace irbfn

using namespace irbfn;

extern "C" size_t irbfn_clip_adam_workspace_bytes(int64_t n) {
    return n <= 0 ? 0 : (size_t)opt_blocks(n) * sizeof(double);
}

extern "C" int irbfn_clip_adam(float* params, float* grad, float* mu, float* nu, int64_t n, double* step_count,
                               double lr, double b1, double b2, double eps, double max_norm, float* grad_norm,
                               void* workspace, size_t workspace_bytes, irbfn_stream_t stream) {
    if (n <= 0) return IRBFN_ERR_BAD_DIMS;
    if (!params || !grad || !mu || !nu || !step_count) return IRBFN_ERR_NULL_POINTER;
    if (!(b1 >= 0.0 && b1 < 1.0) || !(b2 >= 0.0 && b2 < 1.0)) return IRBFN_ERR_BAD_DIMS;
    if (!workspace || workspace_bytes < irbfn_clip_adam_workspace_bytes(n)) return IRBFN_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = opt_blocks(n);
    sumsq_kernel<<<nb, kOptThreads, 0, st>>>(grad, n, (double*)workspace, step_count);
    int e = (int)cudaGetLastError();
    if (e != 0) return note_cuda_error("sumsq_kernel", e);
    AdamArgs a;
    a.p = params; a.g = grad; a.mu = mu; a.nu = nu; a.part = (const double*)workspace; a.step_count = step_count;
    a.gnorm_out = grad_norm; a.n = n; a.nparts = nb; a.b1d = b1; a.b2d = b2;
    a.lr = (float)lr; a.b1 = (float)b1; a.b2 = (float)b2; a.omb1 = (float)(1.0 - b1); a.omb2 = (float)(1.0 - b2);
    a.eps = (float)eps; a.max_norm = (float)max_norm;
    clip_adam_kernel<<<nb, kOptThreads, 0, st>>>(a);
    e = (int)cudaGetLastError();
    return e != 0 ? note_cuda_error("clip_adam_kernel", e) : IRBFN_OK;
}

// Fused tail of a training step: optional one-shot all-reduce over peer memory, then clip + Adam, ONE launch.
//   reduced      [n + n_extra] local output: the summed gradient (first n floats: the flat parameter gradient,
//                clipped in place like irbfn_clip_adam does; the n_extra trailing floats are summed only)
//   peers        device array of `world` float pointers (peers[rank] = this rank's own symmetric buffer holding
//                its local gradient) or NULL with world == 1: then `reduced` already holds the gradient
//   apply_update 0: all-reduce only
// Returns IRBFN_ERR_BAD_DIMS for n + n_extra > irbfn_fused_update_max_n() (use NCCL + irbfn_clip_adam).
extern "C" int64_t irbfn_fused_update_max_n(void) { return kFusedMaxN; }

extern "C" size_t irbfn_fused_update_buffer_bytes(int64_t n_total, int world) {
    if (n_total <= 0 || world <= 0) return 0;
    return (size_t)(((n_total + 3) / 4) * 4 + 2 * world + 1 + 3) * sizeof(float);
}

extern "C" int irbfn_fused_update(float* params, float* mu, float* nu, float* reduced, int64_t n, int n_extra,
                                  float* const* peers, int rank, int world, int apply_update, double* step_count,
                                  double lr, double b1, double b2, double eps, double max_norm, float* grad_norm,
                                  int* status, irbfn_stream_t stream) {
    if (n <= 0 || n_extra < 0 || n + n_extra > kFusedMaxN) return IRBFN_ERR_BAD_DIMS;
    if (world < 1 || world > 32 || rank < 0 || rank >= world) return IRBFN_ERR_BAD_DIMS;
    if (!reduced) return IRBFN_ERR_NULL_POINTER;
    if (apply_update && (!params || !mu || !nu || !step_count)) return IRBFN_ERR_NULL_POINTER;
    if (world > 1 && !peers) return IRBFN_ERR_NULL_POINTER;
    if (!(b1 >= 0.0 && b1 < 1.0) || !(b2 >= 0.0 && b2 < 1.0)) return IRBFN_ERR_BAD_DIMS;
    FusedArgs a;
    a.p = params; a.mu = mu; a.nu = nu; a.reduced = reduced; a.peers = world > 1 ? peers : nullptr;
    a.step_count = step_count; a.gnorm_out = grad_norm; a.status = status;
    a.n = n; a.n_extra = n_extra; a.rank = rank; a.world = world; a.apply_update = apply_update;
    a.b1d = b1; a.b2d = b2;
    a.lr = (float)lr; a.b1 = (float)b1; a.b2 = (float)b2; a.omb1 = (float)(1.0 - b1); a.omb2 = (float)(1.0 - b2);
    a.eps = (float)eps; a.max_norm = (float)max_norm;
    fused_update_kernel<<<1, kFusedThreads, 0, (cudaStream_t)stream>>>(a);
    const int e = (int)cudaGetLastError();
    return e != 0 ? note_cuda_error("fused_update_kernel", e) : IRBFN_OK;
}
