// Loss-side vehicle dynamics of the IRBFN training steps (sm_100a), device helpers.
//
// Reference (paths relative to the reference checkout):
//   dynamic_frenet_onestep      src/irbfn_mpc/dynamics.py:190-281   (x + deriv_x_ls * dt; the high-speed
//                               branch is commented out in the reference, :279-280)
//   dynamic_frenet_onestep_aux  src/irbfn_mpc/dynamics.py:293-383
//   dynamic_st_onestep          src/irbfn_mpc/dynamics.py:9-91      (select(V > 3, f, f_ks))
//   dynamic_st_onestep_aux      src/irbfn_mpc/dynamics.py:103-187   (f_ks only)
//
// Everything here is element-wise per sample: one thread owns one sample, state in registers, the
// five Euler steps unrolled.  These kernels are HBM-bound: a sample is 4(D + 2 O)
// bytes read and 4 O written for ~300 flops.
#pragma once
#include "irbfn_common.cuh"

namespace irbfn {

struct DynParams {  // dyn_params of scripts/train_nmpc_frenet.py:322-338, same order
    float mu, m, I, lf, lr, csf, csr, h, dt, sv_max, a_max, s_max, v_max;
};

__device__ __forceinline__ float clipf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
// d clip / d v with jnp.clip = minimum(maximum(v, lo), hi): 1 inside, 0 outside, 1/2 on a bound
__device__ __forceinline__ float clip_grad(float v, float lo, float hi) {
    return (v > lo && v < hi) ? 1.f : ((v == lo || v == hi) ? 0.5f : 0.f);
}
__device__ __forceinline__ float signf(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

// Frenet state that the low-speed model actually moves; vy, wz and cur are constants of the rollout.
struct FrState {
    float s, ey, dl, vx, ep;
};

// One Euler step of dynamic_frenet_onestep (dynamics.py:226-281), low-speed derivatives.
__device__ __forceinline__ FrState frenet_step(const FrState z, float cu, float a_raw, float dv_raw,
                                               const DynParams& p) {
    const float dc = clipf(z.dl, -p.s_max, p.s_max);
    const float a = clipf(a_raw, -p.a_max, p.a_max);
    const float dv = clipf(dv_raw, -p.sv_max, p.sv_max);
    float sn, cs;
    sincosf(z.ep, &sn, &cs);
    const float sd = (z.vx * cs) / (1.f - z.ey * cu);
    FrState o;
    o.s = z.s + sd * p.dt;
    o.ey = z.ey + z.vx * sn * p.dt;
    o.dl = z.dl + dv * p.dt;
    o.vx = z.vx + a * p.dt;
    o.ep = z.ep + ((z.vx * tanf(dc)) / (p.lr + p.lf) - cu * sd) * p.dt;
    return o;
}

// Adjoint of frenet_step: g holds d loss / d (output state) on entry and d loss / d (input state) on
// exit; ga / gdv receive d loss / d (raw controls).
__device__ __forceinline__ void frenet_step_adjoint(const FrState z, float cu, float a_raw, float dv_raw,
                                                    const DynParams& p, FrState& g, float& ga, float& gdv) {
    const float dc = clipf(z.dl, -p.s_max, p.s_max);
    float sn, cs;
    sincosf(z.ep, &sn, &cs);
    const float q = 1.f - z.ey * cu;
    const float sd = (z.vx * cs) / q;
    const float L = p.lr + p.lf;
    const float td = tanf(dc);
    const float gsd = p.dt * g.s - p.dt * cu * g.ep;            // s' and epsi' both read sd
    FrState gi;
    gi.s = g.s;
    gi.ey = g.ey + (sd / q) * cu * gsd;                          // d sd / d ey = sd * cu / q
    gi.dl = g.dl + p.dt * z.vx * (1.f + td * td) / L * g.ep * clip_grad(z.dl, -p.s_max, p.s_max);
    gi.vx = g.vx + p.dt * sn * g.ey + p.dt * td / L * g.ep + (cs / q) * gsd;
    gi.ep = g.ep + p.dt * z.vx * cs * g.ey - (z.vx * sn / q) * gsd;
    ga = p.dt * g.vx * clip_grad(a_raw, -p.a_max, p.a_max);
    gdv = p.dt * g.dl * clip_grad(dv_raw, -p.sv_max, p.sv_max);
    g = gi;
}

// Single-track derivatives (dynamics.py:47-88): returns x + select(V > 3, f, f_ks) * dt.
__device__ __forceinline__ void st_step(float (&x)[7], float a_raw, float sv_raw, const DynParams& p,
                                        bool kinematic_only) {
    const float g = 9.81f;
    const float DELTA = clipf(x[2], -p.s_max, p.s_max);
    const float V = clipf(x[3], -p.v_max, p.v_max);
    const float PSI = x[4], PSI_DOT = x[5], BETA = x[6];
    const float ACCL = clipf(a_raw, -p.a_max, p.a_max);
    const float SV = clipf(sv_raw, -p.sv_max, p.sv_max);
    float f[7];
    if (!kinematic_only && V > 3.0f) {
        const float rf = g * p.lr - ACCL * p.h, rr = g * p.lf + ACCL * p.h;
        float sn, cs;
        sincosf(PSI + BETA, &sn, &cs);
        f[0] = V * cs;
        f[1] = V * sn;
        f[2] = SV;
        f[3] = ACCL;
        f[4] = PSI_DOT;
        f[5] = ((p.mu * p.m) / (p.I * (p.lf + p.lr))) *
               (p.lf * p.csf * rf * DELTA + (p.lr * p.csr * rr - p.lf * p.csf * rf) * BETA -
                (p.lf * p.lf * p.csf * rf + p.lr * p.lr * p.csr * rr) * (PSI_DOT / V));
        f[6] = (p.mu / (V * (p.lr + p.lf))) *
                   (p.csf * rf * DELTA - (p.csr * rr + p.csf * rf) * BETA +
                    (p.csr * rr * p.lr - p.csf * rf * p.lf) * (PSI_DOT / V)) - PSI_DOT;
    } else {
        float sn, cs;
        sincosf(PSI, &sn, &cs);
        f[0] = V * cs;
        f[1] = V * sn;
        f[2] = SV;
        f[3] = ACCL;
        f[4] = (V / (p.lr + p.lf)) * tanf(DELTA);
        f[5] = 0.f;
        f[6] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) x[i] += f[i] * p.dt;
}

}  // namespace irbfn
