// Host-side launch planning: how the (R,K) parameter set is cut into shared-memory stages.
// Plain C++ (no CUDA types) so it can be unit-tested on a machine without a GPU.
#pragma once
#include <stddef.h>

namespace irbfn {

// Forward tile geometry; must match FwdShape in irbfn_fwd.cuh.
constexpr int kFwdTQ = 1024;  // queries per CTA tile
constexpr int kFwdKT = 4;     // kernels per register tile

inline int fwd_rec_floats(int D) { return ((2 * D + 2 + 3) / 4) * 4; }

// Compiled output-window widths of the forward kernel.
inline int fwd_ot_for(int ocur) {
    if (ocur <= 0) return 0;
    if (ocur <= 2) return 2;
    if (ocur <= 4) return 4;
    if (ocur <= 8) return 8;
    if (ocur <= 10) return 10;
    return 16;
}

struct FwdPlan {
    int RC;          // regions per stage
    int KC;          // kernels per stage (multiple of kFwdKT)
    int stage_once;  // whole parameter set resident: staged once per CTA
    size_t smem_bytes;
};

inline size_t fwd_smem_bytes(int RC, int KC, int D, int OT) {
    return ((size_t)RC * KC * fwd_rec_floats(D) + (size_t)KC * OT + (size_t)RC * kFwdTQ) * sizeof(float);
}

inline bool plan_forward(int R, int K, int D, int OT, int smem_limit, FwdPlan* p) {
    const int Kp = ((K + kFwdKT - 1) / kFwdKT) * kFwdKT;
    if (fwd_smem_bytes(R, Kp, D, OT) <= (size_t)smem_limit) {
        *p = FwdPlan{R, Kp, 1, fwd_smem_bytes(R, Kp, D, OT)};
        return true;
    }
    const int want = Kp < 64 ? Kp : 64;
    const int cands[5] = {R < 16 ? R : 16, 8, 4, 2, 1};
    for (int i = 0; i < 5; ++i) {
        const int RC = cands[i] < R ? cands[i] : R;
        if (RC <= 0) continue;
        const size_t fixed = (size_t)RC * kFwdTQ * sizeof(float);
        if (fixed >= (size_t)smem_limit) continue;
        const size_t per_k = ((size_t)RC * fwd_rec_floats(D) + OT) * sizeof(float);
        long long KC = (long long)(((size_t)smem_limit - fixed) / per_k);
        KC = (KC / kFwdKT) * kFwdKT;
        if (KC > Kp) KC = Kp;
        if (KC >= want || (RC == 1 && KC >= kFwdKT)) {
            *p = FwdPlan{RC, (int)KC, 0, fwd_smem_bytes(RC, (int)KC, D, OT)};
            return true;
        }
    }
    return false;
}

}  // namespace irbfn
