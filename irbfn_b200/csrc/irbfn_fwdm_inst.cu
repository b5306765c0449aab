// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_fwdm_inst.cu.
// The gate-sparse and gate-writing variants of the forward (irbfn_fwd_modes.cuh).
#include "irbfn_fwd_modes.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
int IRBFN_CAT(launch_fwdm_d, IRBFN_D)(const FwdArgs& a, int OT, int basis, int nt, int grid,
                                      size_t smem, cudaStream_t st) {
    return launch_fwdm<IRBFN_D>(a, OT, basis, nt, grid, smem, st);
}
}  // namespace irbfn
