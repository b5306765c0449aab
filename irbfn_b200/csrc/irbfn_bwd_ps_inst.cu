// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_bwd_ps_inst.cu.
// The gate-sparse instantiations of the parameter-gradient kernel (irbfn_bwd_p.cuh, SPARSE = 1).
#include "irbfn_bwd_p.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
int IRBFN_CAT(launch_bwd_ps_d, IRBFN_D)(const BwdPArgs& a, int OT, int basis, int dht, int grid, int nt,
                                        size_t smem, cudaStream_t st) {
    return launch_bwd_p_sparse<IRBFN_D>(a, OT, basis, dht, grid, nt, smem, st);
}
}  // namespace irbfn
