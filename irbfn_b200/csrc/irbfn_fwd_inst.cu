// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_fwd_inst.cu.
// Keeps each TU small so the in-tree build parallelises (irbfn_b200/build.py).
#include "irbfn_fwd.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
int IRBFN_CAT(launch_fwd_d, IRBFN_D)(const FwdArgs& a, int OT, int basis, int nt, int grid,
                                     size_t smem, cudaStream_t st) {
    return launch_fwd<IRBFN_D>(a, OT, basis, nt, grid, smem, st);
}
}  // namespace irbfn
