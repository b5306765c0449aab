// One translation unit per compiled feature count: nvcc -DIRBFN_D=<4..8> irbfn_bwd_pw_inst.cu.
#include "irbfn_bwd_pw.cuh"

#ifndef IRBFN_D
#error "compile with -DIRBFN_D=<4..8>"
#endif

#define IRBFN_CAT_(a, b) a##b
#define IRBFN_CAT(a, b) IRBFN_CAT_(a, b)

namespace irbfn {
int IRBFN_CAT(launch_bwd_pw_d, IRBFN_D)(const BwdPWArgs& a, int OT, int basis, int grid, int nt, size_t smem,
                                        cudaStream_t st) {
    return launch_bwd_pw<IRBFN_D>(a, OT, basis, grid, nt, smem, st);
}
}  // namespace irbfn
