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
