// C ABI of the one-pass Jacobian (include/irbfn_b200.h: irbfn_jacobian).
#include "irbfn_jac.cuh"
#include "irbfn_plan.h"

namespace irbfn {
int device_info(int* sms, int* smem_optin);                                        // irbfn_api.cu
int check_gate(const float* gate, const IrbfnGateDesc* gd, int D, GateDev* out);   // irbfn_api.cu
int launch_jac_d4(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jac_d5(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jac_d6(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jac_d7(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jac_d8(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jacs_d4(const JacArgs&, int, int, int, size_t, cudaStream_t);   // gate-sparse instantiations
int launch_jacs_d5(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jacs_d6(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jacs_d7(const JacArgs&, int, int, int, size_t, cudaStream_t);
int launch_jacs_d8(const JacArgs&, int, int, int, size_t, cudaStream_t);
size_t sparse_bin_workspace_bytes(long long B, int num_bounds);             // irbfn_sparse.cu
int sparse_bin_queries(const float*, const GateDev&, float, long long, int, int, void*, const int**, cudaStream_t,
                       const float** edge_out);
int note_cuda_error(const char* where, int cuda_error);

// stage [RC][KC/4] transposed tiles + readout rows [KC][OT] + d log gamma columns [D][NT] float2
static size_t jac_smem_bytes(int RC, int KC, int D, int OT) {
    return ((size_t)RC * (KC / kFwdKT) * fwd_tile_floats(D) + (size_t)KC * OT + (size_t)D * JacShape::NT * 2) *
           sizeof(float);
}
}  // namespace irbfn

using namespace irbfn;

static int jacobian_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                         const float* bias, const float* gate, const IrbfnGateDesc* gd, float* jac,
                         float* y, float* region_y, int64_t B, int R, int K, int D, int O, int basis,
                         float gate_tol, void* workspace, size_t workspace_bytes, irbfn_stream_t stream);

extern "C" int irbfn_jacobian(const float* x, const float* centers, const float* log_sigs, const float* W,
                              const float* bias, const float* gate, const IrbfnGateDesc* gd, float* jac,
                              float* y, float* region_y, int64_t B, int R, int K, int D, int O, int basis,
                              irbfn_stream_t stream) {
    return jacobian_impl(x, centers, log_sigs, W, bias, gate, gd, jac, y, region_y, B, R, K, D, O, basis, -1.0f, nullptr,
                         0, stream);
}

extern "C" size_t irbfn_jacobian_sparse_workspace_bytes(int64_t B, int num_bounds) {
    if (B <= 0 || num_bounds < 0) return 0;
    return sparse_bin_workspace_bytes(B, num_bounds);
}

extern "C" int irbfn_jacobian_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                                     const float* bias, const IrbfnGateDesc* gd, float* jac, float* y, int64_t B,
                                     int R, int K, int D, int O, int basis, float gate_tol, void* workspace,
                                     size_t workspace_bytes, irbfn_stream_t stream) {
    if (!(gate_tol >= 0.0f)) return IRBFN_ERR_BAD_DIMS;
    if (gd == nullptr) return IRBFN_ERR_BAD_GATE;                     // the tanh gate only
    return jacobian_impl(x, centers, log_sigs, W, bias, nullptr, gd, jac, y, nullptr, B, R, K, D, O, basis, gate_tol,
                         workspace, workspace_bytes, stream);
}

static int jacobian_impl(const float* x, const float* centers, const float* log_sigs, const float* W,
                         const float* bias, const float* gate, const IrbfnGateDesc* gd, float* jac,
                         float* y, float* region_y, int64_t B, int R, int K, int D, int O, int basis,
                         float gate_tol, void* workspace, size_t workspace_bytes, irbfn_stream_t stream) {
    if (B < 0 || R <= 0 || K <= 0 || D <= 0 || O <= 0) return IRBFN_ERR_BAD_DIMS;
    if (D > IRBFN_MAX_D) return IRBFN_ERR_UNSUPPORTED_D;
    if (basis != IRBFN_BASIS_GAUSSIAN && basis != IRBFN_BASIS_INVERSE_QUADRATIC) return IRBFN_ERR_BAD_BASIS;
    GateDev gdev;
    int rc = check_gate(gate, gd, D, &gdev);
    if (rc != IRBFN_OK) return rc;
    if (!jac || !centers || !log_sigs || !W || (y && !bias)) return IRBFN_ERR_NULL_POINTER;
    if (B == 0) return IRBFN_OK;
    if (!x) return IRBFN_ERR_NULL_POINTER;
    int sms = 0, smem_limit = 0;
    rc = device_info(&sms, &smem_limit);
    if (rc != IRBFN_OK) return rc;
    const int DT = D < 4 ? 4 : D;
    const int Kp = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
    // gate-sparse: bin the queries once (perm, interval edges); shapes the sparse kernel does not cover run dense
    const int* perm = nullptr;
    const float* edge = nullptr;
    if (gate_tol >= 0.0f && R > 1 && gdev.num_bounds <= 64 && B <= 0x7fffffffll) {
        if (workspace == nullptr || workspace_bytes < sparse_bin_workspace_bytes(B, gdev.num_bounds)) return IRBFN_ERR_WORKSPACE;
        const int be = sparse_bin_queries(x, gdev, gate_tol, B, D, sms, workspace, &perm, (cudaStream_t)stream, &edge);
        if (be != 0) return note_cuda_error("sparse binning (jacobian)", be);
    }
    for (int o0 = 0; o0 < O;) {
        const int ocur = O - o0 > 4 ? 4 : O - o0;
        const int OT = ocur <= 2 ? 2 : 4;
        JacArgs a{};
        if (jac_smem_bytes(R, Kp, DT, OT) <= (size_t)smem_limit) {
            a.RC = R; a.KC = Kp; a.stage_once = 1;
        } else {
            const size_t fixed = (size_t)DT * JacShape::NT * 2 * sizeof(float);
            const size_t per_k = ((size_t)fwd_tile_floats(DT) / kFwdKT + OT) * sizeof(float);
            long long KC = ((long long)(((size_t)smem_limit - fixed) / per_k) / kFwdKT) * kFwdKT;
            if (KC > Kp) KC = Kp;
            if (KC < kFwdKT) return IRBFN_ERR_BAD_DIMS;
            a.RC = 1; a.KC = (int)KC; a.stage_once = 0;
        }
        const size_t smem = jac_smem_bytes(a.RC, a.KC, DT, OT);
        a.x = x; a.centers = centers; a.log_sigs = log_sigs; a.W = W; a.bias = bias; a.gate_ext = gate; a.gd = gdev;
        a.jac = jac; a.y = y; a.region_y = region_y;
        a.B = B; a.R = R; a.K = K; a.Dsrc = D; a.O = O; a.o0 = o0; a.ocur = ocur;
        a.num_tiles = (int)((B + JacShape::TQ - 1) / JacShape::TQ);
        int e;
        if (perm != nullptr && a.stage_once) {
            a.perm = perm;
            a.edge = edge;
            switch (DT) {
                case 4: e = launch_jacs_d4(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
                case 5: e = launch_jacs_d5(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
                case 6: e = launch_jacs_d6(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
                case 7: e = launch_jacs_d7(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
                default: e = launch_jacs_d8(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
            }
        } else
        switch (DT) {
            case 4: e = launch_jac_d4(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
            case 5: e = launch_jac_d5(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
            case 6: e = launch_jac_d6(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
            case 7: e = launch_jac_d7(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
            default: e = launch_jac_d8(a, OT, basis, sms, smem, (cudaStream_t)stream); break;
        }
        if (e != 0) return note_cuda_error("irbfn_jac_kernel", e);
        o0 += ocur;
    }
    return IRBFN_OK;
}
