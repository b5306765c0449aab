// Fused training losses with the vehicle dynamics inside (sm_100a) and the plain rollouts.
//
// One launch replaces, per training step of the reference, the whole tail of loss_fn after
// wcrbf.apply: hstack, two vmapped lax.scan rollouts, abs/l2, mean, and -- in the backward -- the
// autodiff transposes of all of them (scripts/train_nmpc_frenet.py:343-384, 399-413;
// scripts/train_nmpc.py:260-293, 306-413).  It reads x, y_pred, y once and writes the cotangent
// dloss/dy_pred (B,O) that irbfn_backward consumes, plus the two loss terms.
#include "irbfn_dynamics.cuh"

namespace irbfn {
namespace {

constexpr int kLossThreads = 256;

struct LossArgs {
    const float* x;       // (B,D)
    const float* yp;      // (B,O)
    const float* y;       // (B,O)
    float* dy;            // (B,O)
    double* part;         // [gridDim.x][2] partial sums (pred, int), unscaled
    long long B;
    int D, O;
    float inv_pred;       // 1 / (global_batch * prediction columns)
    float inv_int;        // weight / (global_batch * integrated-state elements)
    DynParams p;
};

template <int O_>
__device__ __forceinline__ void load_row(const float* base, long long b, int O, float (&v)[O_]) {
#pragma unroll
    for (int i = 0; i < O_; ++i) v[i] = i < O ? __ldg(base + b * O + i) : 0.f;
}

// MODE 0: Frenet, one aux step          scripts/train_nmpc_frenet.py:343-384
// MODE 1: Frenet, five-step rollout     scripts/train_nmpc_frenet.py:399-413
// MODE 2: single track, one aux step    scripts/train_nmpc.py:260-293
// MODE 3: kinematic five-step rollout   scripts/train_nmpc.py:306-413
template <int MODE, int OM>
__global__ void __launch_bounds__(kLossThreads) dynamics_loss_kernel(const LossArgs a) {
    const DynParams& p = a.p;
    float psum = 0.f, isum = 0.f;
    for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < a.B;
         b += (long long)gridDim.x * blockDim.x) {
        float yp[OM], y[OM], g[OM];
        load_row<OM>(a.yp, b, a.O, yp);
        load_row<OM>(a.y, b, a.O, y);
        const float* xr = a.x + b * a.D;
        if (MODE == 0 || MODE == 1 || MODE == 2) {
#pragma unroll
            for (int i = 0; i < OM; ++i) {
                const float d = yp[i] - y[i];
                if (MODE == 2) { psum += 0.5f * d * d; g[i] = d * a.inv_pred; }
                else { psum += fabsf(d); g[i] = signf(d) * a.inv_pred; }
            }
        } else {
#pragma unroll
            for (int i = 0; i < OM; ++i) g[i] = 0.f;
#pragma unroll
            for (int i = 0; i < OM; i += 5) {   // columns 0 and 5 only (train_nmpc.py:402)
                const float d = yp[i] - y[i];
                psum += fabsf(d);
                g[i] = signf(d) * a.inv_pred;
            }
        }

        if constexpr (MODE == 0) {
            // x_u = [ey, delta, vx, vy, wz, epsi, cur | u]; a = clip(x_u[8]) = u[1], deltv = clip(x_u[9]) =
            // u[2], the index clamped to the last column as JAX does (dynamics.py:326-327)
            const int ia = min(1, a.O - 1), idv = min(2, a.O - 1);
            const float dl = __ldg(xr + 1), vx = __ldg(xr + 2);
            float ua_p = 0.f, ua_a = 0.f, ud_p = 0.f, ud_a = 0.f;
#pragma unroll
            for (int i = 0; i < OM; ++i) {
                if (i == ia) { ua_p = yp[i]; ua_a = y[i]; }
                if (i == idv) { ud_p = yp[i]; ud_a = y[i]; }
            }
            const float d_dl = (dl + clipf(ud_p, -p.sv_max, p.sv_max) * p.dt) - (dl + clipf(ud_a, -p.sv_max, p.sv_max) * p.dt);
            const float d_vx = (vx + clipf(ua_p, -p.a_max, p.a_max) * p.dt) - (vx + clipf(ua_a, -p.a_max, p.a_max) * p.dt);
            isum += fabsf(d_dl) + fabsf(d_vx);   // ey', vy, wz, epsi' do not depend on the controls
            const float g_a = signf(d_vx) * a.inv_int * p.dt * clip_grad(ua_p, -p.a_max, p.a_max);
            const float g_d = signf(d_dl) * a.inv_int * p.dt * clip_grad(ud_p, -p.sv_max, p.sv_max);
#pragma unroll
            for (int i = 0; i < OM; ++i) {
                if (i == ia) g[i] += g_a;
                if (i == idv) g[i] += g_d;
            }
        } else if constexpr (MODE == 1) {
            // initial_state = x[:, [0,0,1,2,3,5,6,7]] -> (s, ey, delta, vx, vy, wz, epsi, cur)
            const float cu = __ldg(xr + 7);
            FrState z0;
            z0.s = z0.ey = __ldg(xr + 0);
            z0.dl = __ldg(xr + 1);
            z0.vx = __ldg(xr + 2);
            z0.ep = __ldg(xr + 6);
            FrState za[5], zp[6];
            zp[0] = z0;
            FrState t = z0;
#pragma unroll
            for (int i = 0; i < 5; ++i) { t = frenet_step(t, cu, y[i], y[5 + i], p); za[i] = t; }
#pragma unroll
            for (int i = 0; i < 5; ++i) zp[i + 1] = frenet_step(zp[i], cu, yp[i], yp[5 + i], p);
            FrState gs = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 4; i >= 0; --i) {
                const float d0 = zp[i + 1].s - za[i].s, d1 = zp[i + 1].ey - za[i].ey, d2 = zp[i + 1].dl - za[i].dl,
                            d3 = zp[i + 1].vx - za[i].vx, d4 = zp[i + 1].ep - za[i].ep;
                isum += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3) + fabsf(d4);
                gs.s += signf(d0) * a.inv_int;
                gs.ey += signf(d1) * a.inv_int;
                gs.dl += signf(d2) * a.inv_int;
                gs.vx += signf(d3) * a.inv_int;
                gs.ep += signf(d4) * a.inv_int;
                float ga, gdv;
                frenet_step_adjoint(zp[i], cu, yp[i], yp[5 + i], p, gs, ga, gdv);
                g[i] += ga;
                g[5 + i] += gdv;
            }
        } else if constexpr (MODE == 2) {
            // initial state zeros except V = x[:,0]; only column 3 (V + ACCL dt) of [0,1,3,4] sees the
            // controls: X' = V dt, Y' = 0, PSI' = 0 on both sides (train_nmpc.py:261-264, dynamics.py:173-186)
            const float v0 = __ldg(xr + 0);
            const float d = (v0 + clipf(yp[0], -p.a_max, p.a_max) * p.dt) - (v0 + clipf(y[0], -p.a_max, p.a_max) * p.dt);
            isum += 0.5f * d * d;
            g[0] += d * a.inv_int * p.dt * clip_grad(yp[0], -p.a_max, p.a_max);
        } else {
            const float DT = 0.1f, WB = 0.33f, MAXV = 7.0f, MINV = 0.0f, MAXS = 0.4189f;
            const float v0 = clipf(__ldg(xr + 0), MINV, MAXV);
            // actual rollout: final state only
            float X = 0.f, Y = 0.f, dl = 0.f, v = v0, yaw = 0.f;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                float sn, cs;
                sincosf(yaw, &sn, &cs);
                X += v * cs * DT;
                Y += v * sn * DT;
                dl = clipf(dl + y[5 + i] * DT, -MAXS, MAXS);
                v = clipf(v + y[i] * DT, MINV, MAXV);
                yaw += (v / WB) * tanf(dl) * DT;
            }
            // predicted rollout, states kept for the adjoint
            float pv[6], pyaw[6], pdl[6], pre_dl[5], pre_v[5];
            float pX = 0.f, pY = 0.f;
            pv[0] = v0; pyaw[0] = 0.f; pdl[0] = 0.f;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                float sn, cs;
                sincosf(pyaw[i], &sn, &cs);
                pX += pv[i] * cs * DT;
                pY += pv[i] * sn * DT;
                pre_dl[i] = pdl[i] + yp[5 + i] * DT;
                pdl[i + 1] = clipf(pre_dl[i], -MAXS, MAXS);
                pre_v[i] = pv[i] + yp[i] * DT;
                pv[i + 1] = clipf(pre_v[i], MINV, MAXV);
                pyaw[i + 1] = pyaw[i] + (pv[i + 1] / WB) * tanf(pdl[i + 1]) * DT;
            }
            const float d0 = pX - X, d1 = pY - Y, d2 = pdl[5] - dl, d3 = pv[5] - v, d4 = pyaw[5] - yaw;
            isum += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3) + fabsf(d4);
            const float gX = signf(d0) * a.inv_int, gY = signf(d1) * a.inv_int;
            float gdl = signf(d2) * a.inv_int, gv = signf(d3) * a.inv_int, gyaw = signf(d4) * a.inv_int;
#pragma unroll
            for (int i = 4; i >= 0; --i) {
                const float td = tanf(pdl[i + 1]);
                gv += gyaw * td / WB * DT;                                  // yaw' reads the NEW v and delta
                gdl += gyaw * (pv[i + 1] / WB) * (1.f + td * td) * DT;
                const float gpv = gv * clip_grad(pre_v[i], MINV, MAXV);
                const float gpd = gdl * clip_grad(pre_dl[i], -MAXS, MAXS);
                g[i] += gpv * DT;
                g[5 + i] += gpd * DT;
                float sn, cs;
                sincosf(pyaw[i], &sn, &cs);
                gv = gpv + (gX * cs + gY * sn) * DT;
                gyaw += (-gX * sn + gY * cs) * pv[i] * DT;
                gdl = gpd;
            }
        }
#pragma unroll
        for (int i = 0; i < OM; ++i)
            if (i < a.O) a.dy[b * a.O + i] = g[i];
    }

    // block partial sums, fixed order -> deterministic
    __shared__ double sp[kLossThreads / 32][2];
    double dp = psum, di = isum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dp += __shfl_down_sync(0xffffffffu, dp, o);
        di += __shfl_down_sync(0xffffffffu, di, o);
    }
    if ((threadIdx.x & 31) == 0) { sp[threadIdx.x >> 5][0] = dp; sp[threadIdx.x >> 5][1] = di; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double tp = 0.0, ti = 0.0;
        for (int w = 0; w < kLossThreads / 32; ++w) { tp += sp[w][0]; ti += sp[w][1]; }
        a.part[2 * blockIdx.x] = tp;
        a.part[2 * blockIdx.x + 1] = ti;
    }
}

// Sum of the per-block (prediction, integration) loss partials: thread t adds blocks t, t+256, ... in
// order, then a fixed shared-memory tree; one block, deterministic.
__global__ void __launch_bounds__(256) dynamics_loss_reduce_kernel(const double* part, int nblocks, double sp,
                                                                   double si, float* losses, int nout) {
    __shared__ double rp[256], ri[256];
    const int t = threadIdx.x;
    double tp = 0.0, ti = 0.0;
    for (int i = t; i < nblocks; i += 256) { tp += part[2 * i]; ti += part[2 * i + 1]; }
    rp[t] = tp; ri[t] = ti;
    __syncthreads();
    for (int h = 128; h > 0; h >>= 1) {
        if (t < h) { rp[t] += rp[t + h]; ri[t] += ri[t + h]; }
        __syncthreads();
    }
    if (t == 0) {
        losses[0] = (float)(rp[0] * sp);
        if (nout > 1) losses[1] = (float)(ri[0] * si);   // the prediction-only loss has ONE output word
    }
}

}  // namespace
}  // namespace irbfn

// ------------------------------------------------------------------------------------------------
// Plain rollouts (what the planners and the loss call): one thread per sample.
namespace irbfn {
namespace {

// integrate_frenet_mult (dynamics.py:284-290): (B,18) -> (B,5,8)
__global__ void integrate_frenet_kernel(const float* xu, float* out, long long B, DynParams p) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float* r = xu + b * 18;
    FrState z;
    z.s = __ldg(r + 0); z.ey = __ldg(r + 1); z.dl = __ldg(r + 2); z.vx = __ldg(r + 3); z.ep = __ldg(r + 6);
    const float vy = __ldg(r + 4), wz = __ldg(r + 5), cu = __ldg(r + 7);
    float* o = out + b * 40;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        z = frenet_step(z, cu, __ldg(r + 8 + i), __ldg(r + 13 + i), p);
        float4* o4 = reinterpret_cast<float4*>(o + 8 * i);
        o4[0] = make_float4(z.s, z.ey, z.dl, z.vx);
        o4[1] = make_float4(vy, wz, z.ep, cu);
    }
}

// dynamic_frenet_onestep_aux (dynamics.py:293-383): (B,W) -> (B,6), W = 7 + O
__global__ void frenet_onestep_aux_kernel(const float* xu, float* out, long long B, int W, DynParams p) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float* r = xu + b * W;
    FrState z;
    z.s = 0.f; z.ey = __ldg(r + 0); z.dl = __ldg(r + 1); z.vx = __ldg(r + 2); z.ep = __ldg(r + 5);
    const float vy = __ldg(r + 3), wz = __ldg(r + 4), cu = __ldg(r + 6);
    const FrState n = frenet_step(z, cu, __ldg(r + min(8, W - 1)), __ldg(r + min(9, W - 1)), p);
    float* o = out + b * 6;
    o[0] = n.ey; o[1] = n.dl; o[2] = n.vx; o[3] = vy; o[4] = wz; o[5] = n.ep;
}

// integrate_st_mult (dynamics.py:94-100): (B,17) -> (B,5,7);  steps == 1, kinematic_only:
// dynamic_st_onestep_aux (dynamics.py:103-187): (B,W>=9) -> (B,7)
__global__ void integrate_st_kernel(const float* xu, float* out, long long B, int W, int steps,
                                    int kinematic_only, DynParams p) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float* r = xu + b * W;
    float x[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) x[i] = __ldg(r + i);
    for (int s = 0; s < steps; ++s) {
        const float a = steps == 1 ? __ldg(r + 7) : __ldg(r + 7 + s);
        const float sv = steps == 1 ? __ldg(r + 8) : __ldg(r + 12 + s);
        st_step(x, a, sv, p, kinematic_only != 0);
        float* o = out + (b * steps + s) * 7;
#pragma unroll
        for (int i = 0; i < 7; ++i) o[i] = x[i];
    }
}

// Prediction term alone, any output width: optax.l2_loss(pred, target).mean() = mean(0.5 d^2)
// (scripts/train_nmpc.py:287) or mean |d|; writes the cotangent and the block partial like the kernels above.
template <int KIND>
__global__ void __launch_bounds__(kLossThreads) prediction_loss_kernel(const float* __restrict__ yp,
                                                                       const float* __restrict__ y,
                                                                       float* __restrict__ dy, double* part,
                                                                       long long n, float inv) {
    float psum = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d = __ldg(yp + i) - __ldg(y + i);
        if (KIND == 0) { psum += 0.5f * d * d; dy[i] = d * inv; }
        else { psum += fabsf(d); dy[i] = signf(d) * inv; }
    }
    __shared__ double sp[kLossThreads / 32];
    double dp = psum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dp += __shfl_down_sync(0xffffffffu, dp, o);
    if ((threadIdx.x & 31) == 0) sp[threadIdx.x >> 5] = dp;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tp = 0.0;
        for (int w = 0; w < kLossThreads / 32; ++w) tp += sp[w];
        part[2 * blockIdx.x] = tp;
        part[2 * blockIdx.x + 1] = 0.0;
    }
}

DynParams make_params(const float* h) {
    DynParams p;
    p.mu = h[0]; p.m = h[1]; p.I = h[2]; p.lf = h[3]; p.lr = h[4]; p.csf = h[5]; p.csr = h[6]; p.h = h[7];
    p.dt = h[8]; p.sv_max = h[9]; p.a_max = h[10]; p.s_max = h[11]; p.v_max = h[12];
    return p;
}

int loss_blocks(long long B) {
    long long nb = (B + kLossThreads - 1) / kLossThreads;
    const long long cap = 148 * 8;   // a multiple of the SM count; grid-stride beyond
    return (int)(nb < 1 ? 1 : (nb > cap ? cap : nb));
}

template <int MODE, int OM>
int launch_loss(const LossArgs& a, int blocks, cudaStream_t st) {
    dynamics_loss_kernel<MODE, OM><<<blocks, kLossThreads, 0, st>>>(a);
    return (int)cudaGetLastError();
}

}  // namespace
}  // namespace irbfn

using namespace irbfn;

extern "C" size_t irbfn_dynamics_loss_workspace_bytes(int64_t B) {
    return B <= 0 ? 0 : (size_t)loss_blocks(B) * 2 * sizeof(double);
}

extern "C" int irbfn_dynamics_loss(int mode, const float* x, const float* y_pred, const float* y,
                                   const float* dyn_params, float* dy, float* losses, int64_t B, int D, int O,
                                   int64_t global_batch, void* workspace, size_t workspace_bytes,
                                   irbfn_stream_t stream) {
    if (B <= 0 || D <= 0 || O <= 0 || global_batch < B) return IRBFN_ERR_BAD_DIMS;
    if (!x || !y_pred || !y || !dyn_params || !dy || !losses) return IRBFN_ERR_NULL_POINTER;
    if (mode < 0 || mode > 3) return IRBFN_ERR_BAD_DIMS;
    if ((mode == IRBFN_LOSS_FRENET_FULLINT || mode == IRBFN_LOSS_ST_FULLINT) && O != 10) return IRBFN_ERR_BAD_DIMS;
    if (O != 2 && O != 10) return IRBFN_ERR_BAD_DIMS;
    if ((mode <= IRBFN_LOSS_FRENET_FULLINT && D != 8) || (mode == IRBFN_LOSS_ST_ONEINT && D < 1) ||
        (mode == IRBFN_LOSS_ST_FULLINT && D < 1))
        return IRBFN_ERR_BAD_DIMS;
    if (workspace_bytes < irbfn_dynamics_loss_workspace_bytes(B) || !workspace) return IRBFN_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    LossArgs a;
    a.x = x; a.yp = y_pred; a.y = y; a.dy = dy; a.part = (double*)workspace; a.B = B; a.D = D; a.O = O;
    a.p = make_params(dyn_params);
    const double gb = (double)global_batch;
    double np_cols = O, ni = 1.0, w = 1.0;
    switch (mode) {
        case IRBFN_LOSS_FRENET_ONEINT: ni = 6.0; w = 100.0; break;
        case IRBFN_LOSS_FRENET_FULLINT: ni = 40.0; break;
        case IRBFN_LOSS_ST_ONEINT: ni = 4.0; break;
        default: ni = 5.0; np_cols = 2.0; break;
    }
    const double sp = 1.0 / (gb * np_cols), si = w / (gb * ni);
    a.inv_pred = (float)sp;
    a.inv_int = (float)si;
    const int blocks = loss_blocks(B);
    int e;
    if (mode == IRBFN_LOSS_FRENET_ONEINT) e = O == 2 ? launch_loss<0, 2>(a, blocks, st) : launch_loss<0, 10>(a, blocks, st);
    else if (mode == IRBFN_LOSS_FRENET_FULLINT) e = launch_loss<1, 10>(a, blocks, st);
    else if (mode == IRBFN_LOSS_ST_ONEINT) e = O == 2 ? launch_loss<2, 2>(a, blocks, st) : launch_loss<2, 10>(a, blocks, st);
    else e = launch_loss<3, 10>(a, blocks, st);
    if (e != 0) return note_cuda_error("dynamics_loss_kernel", e);
    dynamics_loss_reduce_kernel<<<1, 256, 0, st>>>((const double*)workspace, blocks, sp, si, losses, 2);
    e = (int)cudaGetLastError();
    if (e != 0) return note_cuda_error("dynamics_loss_reduce_kernel", e);
    return IRBFN_OK;
}

extern "C" size_t irbfn_prediction_loss_workspace_bytes(int64_t B, int O) {
    return (B <= 0 || O <= 0) ? 0 : (size_t)loss_blocks(B * O) * 2 * sizeof(double);
}

extern "C" int irbfn_prediction_loss(int kind, const float* y_pred, const float* y, float* dy, float* loss,
                                     int64_t B, int O, int64_t global_batch, void* workspace,
                                     size_t workspace_bytes, irbfn_stream_t stream) {
    if (B <= 0 || O <= 0 || global_batch < B) return IRBFN_ERR_BAD_DIMS;
    if (kind != IRBFN_PRED_LOSS_L2 && kind != IRBFN_PRED_LOSS_L1) return IRBFN_ERR_BAD_DIMS;
    if (!y_pred || !y || !dy || !loss) return IRBFN_ERR_NULL_POINTER;
    if (workspace_bytes < irbfn_prediction_loss_workspace_bytes(B, O) || !workspace) return IRBFN_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = (long long)B * O;
    const double sp = 1.0 / ((double)global_batch * (double)O);
    const int blocks = loss_blocks(n);
    if (kind == IRBFN_PRED_LOSS_L2)
        prediction_loss_kernel<0><<<blocks, kLossThreads, 0, st>>>(y_pred, y, dy, (double*)workspace, n, (float)sp);
    else
        prediction_loss_kernel<1><<<blocks, kLossThreads, 0, st>>>(y_pred, y, dy, (double*)workspace, n, (float)sp);
    int e = (int)cudaGetLastError();
    if (e != 0) return note_cuda_error("prediction_loss_kernel", e);
    dynamics_loss_reduce_kernel<<<1, 256, 0, st>>>((const double*)workspace, blocks, sp, 0.0, loss, 1);
    e = (int)cudaGetLastError();
    return e != 0 ? note_cuda_error("dynamics_loss_reduce_kernel", e) : IRBFN_OK;
}

extern "C" int irbfn_integrate(int model, const float* x_and_u, const float* dyn_params, float* states,
                               int64_t B, int width, irbfn_stream_t stream) {
    if (B <= 0) return IRBFN_ERR_BAD_DIMS;
    if (!x_and_u || !dyn_params || !states) return IRBFN_ERR_NULL_POINTER;
    cudaStream_t st = (cudaStream_t)stream;
    const DynParams p = make_params(dyn_params);
    const int nt = 256;
    const int nb = (int)((B + nt - 1) / nt);
    switch (model) {
        case IRBFN_DYN_FRENET_MULT:
            if (width != 18) return IRBFN_ERR_BAD_DIMS;
            integrate_frenet_kernel<<<nb, nt, 0, st>>>(x_and_u, states, B, p);
            break;
        case IRBFN_DYN_FRENET_ONESTEP_AUX:
            if (width < 8) return IRBFN_ERR_BAD_DIMS;
            frenet_onestep_aux_kernel<<<nb, nt, 0, st>>>(x_and_u, states, B, width, p);
            break;
        case IRBFN_DYN_ST_MULT:
            if (width != 17) return IRBFN_ERR_BAD_DIMS;
            integrate_st_kernel<<<nb, nt, 0, st>>>(x_and_u, states, B, width, 5, 0, p);
            break;
        case IRBFN_DYN_ST_ONESTEP_AUX:
            if (width < 9) return IRBFN_ERR_BAD_DIMS;
            integrate_st_kernel<<<nb, nt, 0, st>>>(x_and_u, states, B, width, 1, 1, p);
            break;
        default:
            return IRBFN_ERR_BAD_DIMS;
    }
    const int e = (int)cudaGetLastError();
    return e != 0 ? note_cuda_error("irbfn_integrate", e) : IRBFN_OK;
}
