// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_jac_inst.cu.
#include "irbfn_jac.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
namespace {
template <int OT, int BASIS, int SPARSE = 0>
int launch_inst(const JacArgs& a, int sms, size_t smem, int* grid_out, cudaStream_t st) {
    auto kern = irbfn_jac_kernel<IRBFN_D, OT, BASIS, SPARSE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, JacShape::NT, smem);
    if (e != cudaSuccess) return (int)e;
    if (per_sm < 1) per_sm = 1;
    const long long cap = (long long)sms * per_sm;      // persistent: every resident slot gets one CTA
    const int grid = (int)(a.num_tiles < cap ? a.num_tiles : cap);
    if (grid_out) *grid_out = grid;
    kern<<<grid, JacShape::NT, smem, st>>>(a);
    return (int)cudaGetLastError();
}
}  // namespace

int IRBFN_CAT(launch_jac_d, IRBFN_D)(const JacArgs& a, int OT, int basis, int sms, size_t smem, cudaStream_t st) {
    if (basis == IRBFN_BASIS_GAUSSIAN)
        return OT == 2 ? launch_inst<2, IRBFN_BASIS_GAUSSIAN>(a, sms, smem, nullptr, st)
                       : launch_inst<4, IRBFN_BASIS_GAUSSIAN>(a, sms, smem, nullptr, st);
    return OT == 2 ? launch_inst<2, IRBFN_BASIS_INVERSE_QUADRATIC>(a, sms, smem, nullptr, st)
                   : launch_inst<4, IRBFN_BASIS_INVERSE_QUADRATIC>(a, sms, smem, nullptr, st);
}

// gate-sparse instantiations (irbfn_jacobian_sparse)
int IRBFN_CAT(launch_jacs_d, IRBFN_D)(const JacArgs& a, int OT, int basis, int sms, size_t smem, cudaStream_t st) {
    if (basis == IRBFN_BASIS_GAUSSIAN)
        return OT == 2 ? launch_inst<2, IRBFN_BASIS_GAUSSIAN, 1>(a, sms, smem, nullptr, st)
                       : launch_inst<4, IRBFN_BASIS_GAUSSIAN, 1>(a, sms, smem, nullptr, st);
    return OT == 2 ? launch_inst<2, IRBFN_BASIS_INVERSE_QUADRATIC, 1>(a, sms, smem, nullptr, st)
                   : launch_inst<4, IRBFN_BASIS_INVERSE_QUADRATIC, 1>(a, sms, smem, nullptr, st);
}
}  // namespace irbfn
