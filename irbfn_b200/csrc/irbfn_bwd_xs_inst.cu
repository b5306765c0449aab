// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_bwd_xs_inst.cu.
// The gate-sparse instantiations of the query-side backward (irbfn_bwd_x.cuh, SPARSE = 1).
#include "irbfn_bwd_x.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
int IRBFN_CAT(launch_bwd_xs_d, IRBFN_D)(const BwdXArgs& a, int OT, int basis, int grid, size_t smem, cudaStream_t st) {
    return launch_bwd_x_sparse<IRBFN_D>(a, OT, basis, grid, smem, st);
}
}  // namespace irbfn
