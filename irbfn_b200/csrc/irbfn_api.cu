// C ABI of libirbfn_b200.so (include/irbfn_b200.h): argument validation, launch planning
// (how parameters are chunked through shared memory) and the two small gate kernels.
// The heavy kernels live in irbfn_fwd.cuh / irbfn_bwd_x.cuh / irbfn_bwd_p.cuh and are
// instantiated per feature count in their own translation units.
#include <stdlib.h>
#include <cuda_runtime.h>
#include <stdio.h>

#include "irbfn_common.cuh"
#include "irbfn_fwd.cuh"
#include "irbfn_plan.h"

namespace irbfn {
static thread_local char g_detail[256] = "";
int note_cuda_error(const char* where, int cuda_error) {
    snprintf(g_detail, sizeof(g_detail), "%s: %s (cudaError %d)", where,
             cudaGetErrorString((cudaError_t)cuda_error), cuda_error);
    cudaGetLastError();  // clear the non-sticky error state
    return IRBFN_ERR_LAUNCH;
}
int launch_fwd_d4(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwd_d5(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwd_d6(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwd_d7(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwd_d8(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwd_small(const FwdSmallArgs&, int basis, cudaStream_t);               // irbfn_fwd_small.cu
int launch_fwdm_d4(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);   // gate-writing / gate-sparse variants
int launch_fwdm_d5(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d6(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d7(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);
int launch_fwdm_d8(const FwdArgs&, int, int, int, int, size_t, cudaStream_t);

// ---------------------------------------------------------------------------------------
// device properties, cached per device (immutable after first query; benign race)
// ---------------------------------------------------------------------------------------
struct DevInfo {
    int ok;
    int sms;
    int smem_optin;
    int cc_major;
};

static int query_device(DevInfo* out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) {
        cudaGetLastError();
        return IRBFN_ERR_NO_DEVICE;
    }
    static DevInfo cache[64];
    if (dev < 64 && cache[dev].ok) {
        *out = cache[dev];
        return IRBFN_OK;
    }
    DevInfo d{};
    if (cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&d.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
        cudaGetLastError();
        return IRBFN_ERR_NO_DEVICE;
    }
    if (d.cc_major != 10) return IRBFN_ERR_NO_DEVICE;  // sm_100a SASS only
    d.ok = 1;
    if (dev < 64) cache[dev] = d;
    *out = d;
    return IRBFN_OK;
}

int device_info(int* sms, int* smem_optin) {
    DevInfo d;
    int rc = query_device(&d);
    if (rc != IRBFN_OK) return rc;
    *sms = d.sms;
    *smem_optin = d.smem_optin;
    return IRBFN_OK;
}

int check_gate(const float* gate, const IrbfnGateDesc* gd, int D, GateDev* out) {
    if ((gate == nullptr) == (gd == nullptr)) return IRBFN_ERR_BAD_GATE;
    *out = GateDev{};
    if (gd != nullptr) {
        if (gd->num_split <= 0 || gd->num_split > D || gd->num_split > IRBFN_MAX_S ||
            gd->num_ranges < 0 || gd->num_bounds <= 0)
            return IRBFN_ERR_BAD_GATE;
        if (!gd->lower || !gd->upper || !gd->offset || !gd->delta || !gd->ranges)
            return IRBFN_ERR_NULL_POINTER;
        out->S = gd->num_split;
        out->num_ranges = gd->num_ranges;
        out->num_bounds = gd->num_bounds;
        out->lower = gd->lower;
        out->upper = gd->upper;
        out->offset = gd->offset;
        out->delta = gd->delta;
        out->ranges = gd->ranges;
    }
    return IRBFN_OK;
}

// ---------------------------------------------------------------------------------------
// gamma (B,R) of the tanh region gate alone (model.py:42-95): one thread per query.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) region_gate_kernel(const float* __restrict__ x, GateDev gd,
                                                          float* __restrict__ gamma, long long B,
                                                          int R, int D) {
    extern __shared__ float fac[];  // [num_bounds][256] thread-private columns + the region index table
    gate_rows_block(x, gd, gamma, B, R, D, fac);
}

// ---------------------------------------------------------------------------------------
// ClusterWCRBFNet gate (model.py:402-406): logits = x Wc + bc ; gate = softmax(logits).
// One warp per query; lanes stride over the R regions.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) softmax_gate_kernel(const float* __restrict__ x,
                                                           const float* __restrict__ Wc,
                                                           const float* __restrict__ bc,
                                                           float* __restrict__ gate,
                                                           float* __restrict__ logits, long long B,
                                                           int D, int R) {
    const int lane = threadIdx.x & 31;
    const long long q = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (q >= B) return;
    float xv[IRBFN_MAX_D];
#pragma unroll
    for (int j = 0; j < IRBFN_MAX_D; ++j) xv[j] = j < D ? __ldg(x + q * D + j) : 0.f;
    float m = -INFINITY;
    for (int r = lane; r < R; r += 32) {
        float l = __ldg(bc + r);
#pragma unroll
        for (int j = 0; j < IRBFN_MAX_D; ++j)
            if (j < D) l = fmaf(xv[j], __ldg(Wc + j * R + r), l);
        gate[q * R + r] = l;  // parked; rewritten below
        if (logits) logits[q * R + r] = l;
        m = fmaxf(m, l);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int r = lane; r < R; r += 32) {
        const float e = expf(gate[q * R + r] - m);
        gate[q * R + r] = e;
        s += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float inv = 1.0f / s;
    for (int r = lane; r < R; r += 32) gate[q * R + r] *= inv;
}

// y[i] = sum_s part[s][i] in split order (deterministic) -- closes the small-batch split-K forward
__global__ void __launch_bounds__(256) split_reduce_kernel(const float* __restrict__ part, float* __restrict__ y,
                                                           int nsplit, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = part[i];
    for (int k = 1; k < nsplit; ++k) s += part[(size_t)k * n + i];
    y[i] = s;
}

}  // namespace irbfn

using namespace irbfn;

// ---------------------------------------------------------------------------------------
// irbfn_forward_noalloc: scratch for the split-over-kernels partial outputs lives in the library image.
// In HOST code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the address is
// looked up once per device, with that device current, by cudaGetSymbolAddress (no allocation: the array is
// part of the module's global segment).
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[IRBFN_NOALLOC_SCRATCH_BYTES];

static int noalloc_scratch_address(void** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) {
        cudaGetLastError();
        return IRBFN_ERR_NO_DEVICE;
    }
    static void* cache[64];   // immutable per device after the first lookup; benign race (same value)
    if (dev < 64 && cache[dev] != nullptr) {
        *out = cache[dev];
        return IRBFN_OK;
    }
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
        cudaGetLastError();
        return note_cuda_error("cudaGetSymbolAddress(g_noalloc_scratch)", (int)e);
    }
    if (dev < 64) cache[dev] = ptr;
    *out = ptr;
    return IRBFN_OK;
}

extern "C" {

int irbfn_abi_version(void) { return IRBFN_ABI_VERSION; }

const char* irbfn_last_error_detail(void) { return g_detail; }

const char* irbfn_error_string(int code) {
    switch (code) {
        case IRBFN_OK: return "ok";
        case IRBFN_ERR_BAD_DIMS: return "a dimension is <= 0 or inconsistent";
        case IRBFN_ERR_UNSUPPORTED_D: return "in_features exceeds IRBFN_MAX_D";
        case IRBFN_ERR_BAD_BASIS: return "unsupported basis function id";
        case IRBFN_ERR_NULL_POINTER: return "a required pointer is NULL";
        case IRBFN_ERR_BAD_GATE: return "give exactly one of gate / gate descriptor; 0 < S <= D";
        case IRBFN_ERR_WORKSPACE: return "workspace missing or too small";
        case IRBFN_ERR_LAUNCH: return "CUDA launch / runtime error";
        case IRBFN_ERR_NO_DEVICE: return "no usable sm_100 device is current";
        case IRBFN_ERR_BAD_MASK: return "gradient mask empty or names an output whose pointer is NULL";
        default: return "unknown irbfn error code";
    }
}

size_t irbfn_forward_workspace_bytes(int64_t B, int R, int K, int D, int O, int fused_gate) {
    (void)D; (void)fused_gate;
    if (B <= 0 || K <= 0 || O <= 0) return 0;
    int sms = 148, smem = 0;
    device_info(&sms, &smem);   // keeps the nominal B200 value when no device is present
    int kps;
    int ns = fwd_split(B, R, K, sms, &kps);
    return ns > 1 ? (size_t)ns * (size_t)B * (size_t)O * sizeof(float) : 0;
}

int irbfn_forward_path(int64_t B, int R, int K, int D, int O, int y_only) {
    if (B <= 0 || R <= 0 || K <= 0 || D <= 0 || D > IRBFN_MAX_D || O < 0) return IRBFN_ERR_BAD_DIMS;
    if (y_only && fwd_small_applies(B, R, K, D, O)) return IRBFN_PATH_SMALL;
    int sms = 148, smem = 0;
    device_info(&sms, &smem);   // keeps the nominal B200 value when no device is present
    int kps;
    return fwd_split(B, R, K, sms, &kps) > 1 && O > 0 ? IRBFN_PATH_SPLIT : IRBFN_PATH_TILED;
}

static int forward_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                        const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                        float* hidden, float* gate_out, int64_t B, int R, int K, int D, int O, int basis,
                        void* workspace, size_t workspace_bytes, irbfn_stream_t stream,
                        unsigned* ready = nullptr, int tiles_per_flag = 0, int* stream_status = nullptr);

int irbfn_forward(const float* x, const float* centers, const float* log_sigs, const float* W,
                  const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                  float* hidden, int64_t B, int R, int K, int D, int O, int basis, void* workspace,
                  size_t workspace_bytes, irbfn_stream_t stream) {
    return forward_impl(x, centers, log_sigs, W, bias, gate, gd, y, hidden, nullptr, B, R, K, D, O, basis,
                        workspace, workspace_bytes, stream);
}

int irbfn_forward_gate(const float* x, const float* centers, const float* log_sigs, const float* W,
                       const float* bias, const IrbfnGateDesc* gd, float* y, float* gate_out, int64_t B,
                       int R, int K, int D, int O, int basis, void* workspace, size_t workspace_bytes,
                       irbfn_stream_t stream) {
    if (gd == nullptr || gate_out == nullptr) return IRBFN_ERR_NULL_POINTER;
    return forward_impl(x, centers, log_sigs, W, bias, nullptr, gd, y, nullptr, gate_out, B, R, K, D, O, basis,
                        workspace, workspace_bytes, stream);
}

static int forward_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                        const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                        float* hidden, float* gate_out, int64_t B, int R, int K, int D, int O, int basis,
                        void* workspace, size_t workspace_bytes, irbfn_stream_t stream,
                        unsigned* ready, int tiles_per_flag, int* stream_status) {
    if (B < 0 || R <= 0 || K <= 0 || D <= 0) return IRBFN_ERR_BAD_DIMS;
    if (y != nullptr && O <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (basis != IRBFN_BASIS_GAUSSIAN && basis != IRBFN_BASIS_INVERSE_QUADRATIC)
        return IRBFN_ERR_BAD_BASIS;
    if (y == nullptr && hidden == nullptr) return IRBFN_ERR_NULL_POINTER;
    if (!centers || !log_sigs) return IRBFN_ERR_NULL_POINTER;
    if (y != nullptr && (!W || !bias)) return IRBFN_ERR_NULL_POINTER;
    GateDev gdev;
    int rc = check_gate(gate, gd, D, &gdev);
    if (rc != IRBFN_OK) return rc;
    if (B == 0) return IRBFN_OK;
    if (!x) return IRBFN_ERR_NULL_POINTER;

    DevInfo dev;
    rc = query_device(&dev);
    if (rc != IRBFN_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    // A few queries against a small parameter set (pred_step): the pair-parallel kernel, one launch, no workspace.
    if (y != nullptr && hidden == nullptr && gate_out == nullptr && ready == nullptr &&
        fwd_small_applies(B, R, K, D, O)) {
        FwdSmallArgs sa{};
        sa.x = x;
        sa.centers = centers;
        sa.log_sigs = log_sigs;
        sa.W = W;
        sa.bias = bias;
        sa.gate_ext = gate;
        sa.gd = gdev;
        sa.y = y;
        sa.B = B;
        sa.R = R;
        sa.K = K;
        sa.D = D;
        sa.O = O;
        const int e = launch_fwd_small(sa, basis, st);
        return e == 0 ? IRBFN_OK : note_cuda_error("irbfn_fwd_small_kernel", e);
    }

    const int DT = D < 4 ? 4 : D;  // compiled feature count; extra columns are zero padding
    const int Ototal = y ? O : 0;
    int o0 = 0;
    bool first = true;
    // Small batches leave most SMs idle with one CTA per 1,024 queries: split the kernels of every
    // region over gridDim.y, partial readouts into the workspace, fixed-order reduction afterwards.
    // Without (enough) workspace the unsplit path runs -- same result, more latency.
    int k_per_split = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
    int nsplit = ready != nullptr ? 1 : fwd_split(B, R, K, dev.sms, &k_per_split);   // streamed input: whole tiles per CTA
    float* part = nullptr;
    if (nsplit > 1 && y != nullptr) {
        if (workspace != nullptr && workspace_bytes >= (size_t)nsplit * (size_t)B * (size_t)O * sizeof(float))
            part = (float*)workspace;
        else {
            nsplit = 1;
            k_per_split = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
        }
    }
    // Output windows of at most 16 columns; the hidden activations ride along the first one.
    do {
        const int ocur = Ototal - o0 > 16 ? 16 : Ototal - o0;
        const int OT = fwd_ot_for(ocur);
        FwdPlan plan;
        if (!plan_forward(R, nsplit > 1 ? k_per_split : K, DT, OT, dev.smem_optin, &plan)) return IRBFN_ERR_BAD_DIMS;
        FwdArgs a{};
        a.nsplit = nsplit;
        a.k_per_split = k_per_split;
        a.split_stride = part ? (long long)B * O : 0;
        a.x = x;
        a.centers = centers;
        a.log_sigs = log_sigs;
        a.W = W;
        a.bias = bias;
        a.gate_ext = gate;
        a.gd = gdev;
        a.y = ocur > 0 ? (part ? part : y) : nullptr;
        a.hidden = first ? hidden : nullptr;
        a.gate_out = first ? gate_out : nullptr;
        a.ready = ready;
        a.tiles_per_flag = tiles_per_flag;
        a.stream_status = stream_status;
        if (ready != nullptr)   // the tile counter sits behind the flags
            a.work_counter = ready + ((B + kFwdTQ - 1) / kFwdTQ + tiles_per_flag - 1) / tiles_per_flag;
        a.B = B;
        a.R = R;
        a.K = K;
        a.Dsrc = D;
        a.ldw = O;
        a.o0 = o0;
        a.ocur = ocur;
        a.RC = plan.RC;
        a.KC = plan.KC;
        a.stage_once = plan.stage_once;
        const long long tq = (long long)plan.NT * FwdShape::Q;
        const long long tiles = (B + tq - 1) / tq;
        a.num_tiles = (int)tiles;
        const int grid = (int)(tiles < dev.sms ? tiles : dev.sms);
        int e;
        if (a.gate_out != nullptr || a.ready != nullptr) {   // the variants that also write the gate / wait for streamed rows
            switch (DT) {
                case 4: e = launch_fwdm_d4(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 5: e = launch_fwdm_d5(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 6: e = launch_fwdm_d6(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 7: e = launch_fwdm_d7(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                default: e = launch_fwdm_d8(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            }
        } else {
            switch (DT) {
                case 4: e = launch_fwd_d4(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 5: e = launch_fwd_d5(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 6: e = launch_fwd_d6(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                case 7: e = launch_fwd_d7(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
                default: e = launch_fwd_d8(a, OT, basis, plan.NT, grid, plan.smem_bytes, st); break;
            }
        }
        if (e != 0) return note_cuda_error("irbfn_fwd_kernel", e);
        o0 += ocur;
        first = false;
    } while (o0 < Ototal);
    if (part) {
        const long long n = (long long)B * O;
        split_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(part, y, nsplit, n);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return note_cuda_error("split_reduce_kernel", (int)e);
    }
    return IRBFN_OK;
}

int irbfn_forward_streamed(const float* x, const float* centers, const float* log_sigs, const float* W,
                           const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y, int64_t B,
                           int R, int K, int D, int O, int basis, uint32_t* ready, int64_t queries_per_flag,
                           int* status, irbfn_stream_t stream) {
    if (ready == nullptr || y == nullptr) return IRBFN_ERR_NULL_POINTER;
    if (queries_per_flag <= 0 || queries_per_flag % kFwdTQ != 0 || queries_per_flag / kFwdTQ > (1 << 30))
        return IRBFN_ERR_BAD_DIMS;
    if (O <= 0 || O > 16) return IRBFN_ERR_BAD_DIMS;   // one output window: a second pass would not wait again
    if ((B + kFwdTQ - 1) / kFwdTQ >= (1ll << 31)) return IRBFN_ERR_BAD_DIMS;
    return forward_impl(x, centers, log_sigs, W, bias, gate, gd, y, nullptr, nullptr, B, R, K, D, O, basis, nullptr, 0,
                        stream, ready, (int)(queries_per_flag / kFwdTQ), status);
}

// cuStreamWriteValue32 through the runtime's driver entry point (no link against libcuda): a 32-bit store executed
// by the stream's front end in stream order -- no copy engine, no SM.
typedef int (*StreamWrite32Fn)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
static StreamWrite32Fn stream_write32_fn() {
    static StreamWrite32Fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            p = nullptr;
        }
        return (StreamWrite32Fn)p;
    }();
    return fn;
}

int irbfn_upload_streamed_available(void) { return stream_write32_fn() != nullptr ? 1 : 0; }

int irbfn_upload_streamed(const float* x_host, float* x_dev, int64_t B, int D, uint32_t* ready,
                          int64_t queries_per_flag, const irbfn_stream_t* copy_streams, int num_streams) {
    if (!x_host || !x_dev || !ready || !copy_streams) return IRBFN_ERR_NULL_POINTER;
    if (B < 0 || D <= 0 || queries_per_flag <= 0 || queries_per_flag % kFwdTQ != 0 || num_streams <= 0)
        return IRBFN_ERR_BAD_DIMS;
    StreamWrite32Fn write32 = stream_write32_fn();
    if (write32 == nullptr) {
        snprintf(g_detail, sizeof(g_detail), "irbfn_upload_streamed: the driver has no cuStreamWriteValue32");
        return IRBFN_ERR_LAUNCH;
    }
    int64_t p = 0;
    for (int64_t lo = 0; lo < B; lo += queries_per_flag, ++p) {
        const int64_t hi = lo + queries_per_flag < B ? lo + queries_per_flag : B;
        cudaStream_t st = (cudaStream_t)copy_streams[p % num_streams];   // pieces alternate over the copy streams: one
                                                                         // stream's per-copy set-up hides behind another's transfer
        const cudaError_t e = cudaMemcpyAsync(x_dev + lo * D, x_host + lo * D, (size_t)(hi - lo) * D * sizeof(float),
                                              cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return note_cuda_error("cudaMemcpyAsync (streamed upload)", (int)e);
        const int r = write32(st, (unsigned long long)(uintptr_t)(ready + p), 1u, 0u);
        if (r != 0) {
            snprintf(g_detail, sizeof(g_detail), "cuStreamWriteValue32: CUresult %d", r);
            return IRBFN_ERR_LAUNCH;
        }
    }
    return IRBFN_OK;
}

// No counterpart in the reference (include/irbfn_b200.h).
int irbfn_forward_noalloc(const float* x, const float* centers, const float* log_sigs, const float* W,
                          const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                          float* hidden, int64_t B, int R, int K, int D, int O, int basis,
                          irbfn_stream_t stream) {
    const size_t need = (y != nullptr && O > 0) ? irbfn_forward_workspace_bytes(B, R, K, D, O, gd != nullptr) : 0;
    if (need > (size_t)IRBFN_NOALLOC_SCRATCH_BYTES) {
        snprintf(g_detail, sizeof(g_detail),
                 "irbfn_forward_noalloc: this shape needs %zu bytes of scratch, the built-in array holds %u; "
                 "use irbfn_forward with a caller-owned workspace", need, (unsigned)IRBFN_NOALLOC_SCRATCH_BYTES);
        return IRBFN_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    if (need > 0) {
        const int rc = noalloc_scratch_address(&scratch);
        if (rc != IRBFN_OK) return rc;
    }
    return forward_impl(x, centers, log_sigs, W, bias, gate, gd, y, hidden, nullptr, B, R, K, D, O, basis, scratch,
                        need > 0 ? (size_t)IRBFN_NOALLOC_SCRATCH_BYTES : 0, stream);
}

int irbfn_forward_plan(int R, int K, int D, int O, int smem_limit_bytes, int* regions_per_stage,
                       int* kernels_per_stage, int* threads_per_block, size_t* smem_bytes) {
    if (R <= 0 || K <= 0 || D <= 0 || O < 0 || smem_limit_bytes <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    FwdPlan plan;
    if (!plan_forward(R, K, D < 4 ? 4 : D, fwd_ot_for(O > 16 ? 16 : O), smem_limit_bytes, &plan))
        return IRBFN_ERR_BAD_DIMS;
    if (regions_per_stage) *regions_per_stage = plan.RC;
    if (kernels_per_stage) *kernels_per_stage = plan.KC;
    if (threads_per_block) *threads_per_block = plan.NT;
    if (smem_bytes) *smem_bytes = plan.smem_bytes;
    return IRBFN_OK;
}

int irbfn_region_gate(const float* x, const IrbfnGateDesc* gd, float* gamma, int64_t B, int R,
                      int D, irbfn_stream_t stream) {
    if (B < 0 || R <= 0 || D <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    GateDev gdev;
    int rc = check_gate(nullptr, gd, D, &gdev);
    if (rc != IRBFN_OK) return rc;
    if (B == 0) return IRBFN_OK;
    if (!x || !gamma) return IRBFN_ERR_NULL_POINTER;
    DevInfo dev;
    rc = query_device(&dev);
    if (rc != IRBFN_OK) return rc;
    const size_t smem = gate_rows_smem_bytes(gdev.num_bounds, 1, gdev.num_ranges, gdev.S);
    if (smem > (size_t)dev.smem_optin) return IRBFN_ERR_BAD_GATE;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(region_gate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem) != cudaSuccess)
        return IRBFN_ERR_LAUNCH;
    const long long blocks = (B + 255) / 256;
    region_gate_kernel<<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>(x, gdev, gamma, B, R, D);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? IRBFN_OK : note_cuda_error("region_gate_kernel", (int)e);
}

int irbfn_softmax_gate(const float* x, const float* Wc, const float* bc, float* gate,
                       float* logits, int64_t B, int D, int R, irbfn_stream_t stream) {
    if (B < 0 || R <= 0 || D <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (B == 0) return IRBFN_OK;
    if (!x || !Wc || !bc || !gate) return IRBFN_ERR_NULL_POINTER;
    DevInfo dev;
    int rc = query_device(&dev);
    if (rc != IRBFN_OK) return rc;
    const long long blocks = (B + 7) / 8;
    softmax_gate_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, Wc, bc, gate, logits,
                                                                           B, D, R);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? IRBFN_OK : note_cuda_error("softmax_gate_kernel", (int)e);
}

}  // extern "C"
