// XLA FFI handlers over the C ABI of libirbfn_b200.so (include/irbfn_b200.h): what a maintainer of
// hzheng40/irbfn compiles on the reference side so that the Flax modules of src/irbfn_mpc/model.py call the
// sm_100a kernels through jax.ffi.ffi_call (irbfn_flax.py next to this file registers and wraps them).
//
// UNVERIFIED: neither the build container nor the GPU box has jax / jaxlib / the XLA FFI headers,
// so this file has never been compiled.  irbfn_b200/build.py skips it when `import jax` fails;
// tests/test_ffi_source.py checks that every C entry point called here exists in include/irbfn_b200.h with the
// argument count used.
//
//   nvcc|g++ -std=c++17 -shared -fPIC -I$(python -c "import jax.ffi; print(jax.ffi.include_dir())") -Iinclude \
//        irbfn_b200/ffi/irbfn_ffi.cc -Lirbfn_b200 -lirbfn_b200 -lcudart -o irbfn_b200/ffi/libirbfn_ffi.so
//
// Conventions.  Every array is a float32 (or int32) device buffer owned by XLA; outputs are fully overwritten;
// scratch is an extra uint8 RESULT sized on the Python side with the library's *_workspace_bytes (XLA owns it).
// The static gate attributes of a module (model.py:113-125: lower_bounds, upper_bounds, delta, dimension_ranges)
// arrive as five small constant device arrays packed once per module:
//   lower, upper (sum n_d) f32 | offset (S+1) i32 | delta (S) f32 | ranges (R,S) i32.
// Errors: a non-zero IRBFN_ERR_* becomes ffi::Error(kInvalidArgument | kInternal, text); nothing aborts or prints.
#include <cuda_runtime.h>

#include <string>

#include "irbfn_b200.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;
using F32 = ffi::Buffer<ffi::F32>;
using S32 = ffi::Buffer<ffi::S32>;
using RF32 = ffi::ResultBuffer<ffi::F32>;
using RU8 = ffi::ResultBuffer<ffi::U8>;

static ffi::Error Check(int rc) {
  if (rc == IRBFN_OK) return ffi::Error::Success();
  const auto code = rc == IRBFN_ERR_LAUNCH ? ffi::ErrorCode::kInternal : ffi::ErrorCode::kInvalidArgument;
  return ffi::Error(code, std::string(irbfn_error_string(rc)) + " " + irbfn_last_error_detail());
}

static IrbfnGateDesc Gate(const F32& lo, const F32& hi, const S32& off, const F32& delta, const S32& ranges) {
  IrbfnGateDesc g;
  g.num_split = (int32_t)delta.dimensions()[0];
  g.num_ranges = (int32_t)ranges.dimensions()[0];
  g.num_bounds = (int32_t)lo.dimensions()[0];
  g.lower = lo.typed_data();
  g.upper = hi.typed_data();
  g.offset = off.typed_data();
  g.delta = delta.typed_data();
  g.ranges = ranges.typed_data();
  return g;
}

struct Dims { int64_t B; int R, K, D, O; };
static Dims DimsOf(const F32& x, const F32& centers, int O) {
  const auto xd = x.dimensions();
  const auto cd = centers.dimensions();
  return Dims{xd[0], (int)cd[0], (int)cd[1], (int)cd[2], O};
}

// ------------------------------------------------------------------------------------------------ forward
// WCRBFNet.__call__ (model.py:169-198): y (B,O).  ws: irbfn_forward_workspace_bytes(...) bytes (may be 0).
static ffi::Error FwdImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 b, F32 lo, F32 hi, S32 off,
                          F32 delta, S32 ranges, RF32 y, RU8 ws, int32_t basis) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  return Check(irbfn_forward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                             b.typed_data(), nullptr, &g, y->typed_data(), nullptr, d.B, d.R, d.K, d.D, d.O, basis,
                             ws->size_bytes() ? ws->typed_data() : nullptr, ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnFwd, FwdImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()          // x centers log_sigs W b
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()          // gate constants
        .Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis"));

// Training forward: y and the gate gamma (B,R) the backward will read (irbfn_forward_gate).
static ffi::Error FwdGateOutImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 b, F32 lo, F32 hi, S32 off,
                                 F32 delta, S32 ranges, RF32 y, RF32 gate_out, RU8 ws, int32_t basis) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  return Check(irbfn_forward_gate(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                                  b.typed_data(), &g, y->typed_data(), gate_out->typed_data(), d.B, d.R, d.K, d.D, d.O,
                                  basis, ws->size_bytes() ? ws->typed_data() : nullptr, ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnFwdGateOut, FwdGateOutImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()
        .Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis"));

// DeeperWCRBFNet front end (model.py:259-283): hidden (B,K) = sum_r gamma_r phi_r, no readout.
static ffi::Error FwdHiddenImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 lo, F32 hi, S32 off, F32 delta,
                                S32 ranges, RF32 hidden, int32_t basis) {
  const Dims d = DimsOf(x, centers, 0);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  return Check(irbfn_forward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), nullptr, nullptr, nullptr, &g,
                             nullptr, hidden->typed_data(), d.B, d.R, d.K, d.D, 0, basis, nullptr, 0, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnFwdHidden, FwdHiddenImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>()
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()
        .Ret<F32>().Attr<int32_t>("basis"));

// ClusterWCRBFNet blend + readout (model.py:395-414) with the gate (B,R) computed by the caller's softmax.
static ffi::Error FwdExtGateImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 b, F32 gate, RF32 y,
                                 RU8 ws, int32_t basis) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  return Check(irbfn_forward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                             b.typed_data(), gate.typed_data(), nullptr, y->typed_data(), nullptr, d.B, d.R, d.K, d.D,
                             d.O, basis, ws->size_bytes() ? ws->typed_data() : nullptr, ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnFwdExtGate, FwdExtGateImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis"));

// softmax(x Wc + bc) and the logits (model.py:402-406).
static ffi::Error SoftmaxGateImpl(cudaStream_t s, F32 x, F32 Wc, F32 bc, RF32 gate, RF32 logits) {
  const auto xd = x.dimensions();
  return Check(irbfn_softmax_gate(x.typed_data(), Wc.typed_data(), bc.typed_data(), gate->typed_data(),
                                  logits->typed_data(), xd[0], (int)xd[1], (int)Wc.dimensions()[1], s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnSoftmaxGate, SoftmaxGateImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>().Ret<F32>());

// Gate-sparse inference forward (irbfn_forward_sparse): |y - dense| <= tol_abs per output.
static ffi::Error FwdSparseImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 b, F32 lo, F32 hi, S32 off,
                                F32 delta, S32 ranges, RF32 y, RU8 ws, int32_t basis, float tol_abs) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  return Check(irbfn_forward_sparse(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                                    b.typed_data(), &g, y->typed_data(), d.B, d.R, d.K, d.D, d.O, basis, tol_abs, nullptr,
                                    ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnFwdSparse, FwdSparseImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()
        .Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis").Attr<float>("tol_abs"));

// ------------------------------------------------------------------------------------------------ backward
// VJP of WCRBFNet.apply, tanh gate: (dC, dLS, dW, db, dx).  ws: irbfn_backward_workspace_bytes(B,R,K,D,O,mask).
static ffi::Error BwdImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 lo, F32 hi, S32 off, F32 delta,
                          S32 ranges, F32 dy, RF32 dC, RF32 dLS, RF32 dW, RF32 db, RF32 dx, RU8 ws, int32_t basis,
                          int32_t grad_mask, float gate_tol) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  if (gate_tol >= 0.0f)   // gate-sparse VJP: terms of (query, region) pairs with gamma <= gate_tol are left out
    return Check(irbfn_backward_sparse(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(), nullptr,
                                       &g, dy.typed_data(), nullptr, dC->typed_data(), dLS->typed_data(), dW->typed_data(),
                                       db->typed_data(), dx->typed_data(), nullptr, d.B, d.R, d.K, d.D, d.O, basis,
                                       (unsigned)grad_mask, gate_tol, ws->typed_data(), ws->size_bytes(), s));
  return Check(irbfn_backward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(), nullptr, &g,
                              dy.typed_data(), nullptr, dC->typed_data(), dLS->typed_data(), dW->typed_data(),
                              db->typed_data(), dx->typed_data(), nullptr, d.B, d.R, d.K, d.D, d.O, basis,
                              (unsigned)grad_mask, ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnBwd, BwdImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()                     // x centers log_sigs W
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()          // gate constants
        .Arg<F32>()                                                      // dy
        .Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>()
        .Attr<int32_t>("basis").Attr<int32_t>("grad_mask").Attr<float>("gate_tol"));

// Parameter gradients with the gate saved by IrbfnFwdGateOut (training step; no dx): (dC, dLS, dW, db).
static ffi::Error BwdParamsImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 gate, F32 dy, RF32 dC,
                                RF32 dLS, RF32 dW, RF32 db, RU8 ws, int32_t basis, float gate_tol) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  if (gate_tol >= 0.0f)
    return Check(irbfn_backward_sparse(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                                       gate.typed_data(), nullptr, dy.typed_data(), nullptr, dC->typed_data(),
                                       dLS->typed_data(), dW->typed_data(), db->typed_data(), nullptr, nullptr, d.B, d.R,
                                       d.K, d.D, d.O, basis, (unsigned)IRBFN_GRAD_PARAMS, gate_tol, ws->typed_data(),
                                       ws->size_bytes(), s));
  return Check(irbfn_backward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                              gate.typed_data(), nullptr, dy.typed_data(), nullptr, dC->typed_data(), dLS->typed_data(),
                              dW->typed_data(), db->typed_data(), nullptr, nullptr, d.B, d.R, d.K, d.D, d.O, basis,
                              (unsigned)IRBFN_GRAD_PARAMS, ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnBwdParams, BwdParamsImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis")
        .Attr<float>("gate_tol"));

// DeeperWCRBFNet: cotangent of hidden (B,K) -> (dC, dLS, dx).
static ffi::Error BwdHiddenImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 lo, F32 hi, S32 off, F32 delta,
                                S32 ranges, F32 dhidden, RF32 dC, RF32 dLS, RF32 dx, RU8 ws, int32_t basis) {
  const Dims d = DimsOf(x, centers, 0);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  const unsigned mask = IRBFN_GRAD_CENTERS | IRBFN_GRAD_LOG_SIGS | IRBFN_GRAD_X;
  return Check(irbfn_backward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), nullptr, nullptr, &g, nullptr,
                              dhidden.typed_data(), dC->typed_data(), dLS->typed_data(), nullptr, nullptr,
                              dx->typed_data(), nullptr, d.B, d.R, d.K, d.D, 0, basis, mask, ws->typed_data(),
                              ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnBwdHidden, BwdHiddenImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>()
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()
        .Arg<F32>()
        .Ret<F32>().Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int32_t>("basis"));

// ClusterWCRBFNet: external gate -> (dC, dLS, dW, db, dx through the RBF layers, dgate); the softmax / Dense(D->R)
// chain rule stays in JAX on the caller's side.
static ffi::Error BwdExtGateImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 gate, F32 dy, RF32 dC,
                                 RF32 dLS, RF32 dW, RF32 db, RF32 dx, RF32 dgate, RU8 ws, int32_t basis) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const unsigned mask = IRBFN_GRAD_PARAMS | IRBFN_GRAD_X | IRBFN_GRAD_GATE;
  return Check(irbfn_backward(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                              gate.typed_data(), nullptr, dy.typed_data(), nullptr, dC->typed_data(), dLS->typed_data(),
                              dW->typed_data(), db->typed_data(), dx->typed_data(), dgate->typed_data(), d.B, d.R, d.K,
                              d.D, d.O, basis, mask, ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnBwdExtGate, BwdExtGateImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>()
        .Attr<int32_t>("basis"));

// ------------------------------------------------------------------------------------------------ Jacobian
// jax.vmap(jax.jacrev(lambda q: apply(params, q[None])[0]))(x): jac (B,O,D) and y (B,O), one pass.
static ffi::Error JacobianImpl(cudaStream_t s, F32 x, F32 centers, F32 log_sigs, F32 W, F32 b, F32 lo, F32 hi, S32 off,
                               F32 delta, S32 ranges, RF32 jac, RF32 y, int32_t basis) {
  const Dims d = DimsOf(x, centers, (int)W.dimensions()[1]);
  const IrbfnGateDesc g = Gate(lo, hi, off, delta, ranges);
  return Check(irbfn_jacobian(x.typed_data(), centers.typed_data(), log_sigs.typed_data(), W.typed_data(),
                              b.typed_data(), nullptr, &g, jac->typed_data(), y->typed_data(), nullptr, d.B, d.R, d.K,
                              d.D, d.O, basis, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnJacobian, JacobianImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
        .Arg<F32>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<S32>()
        .Ret<F32>().Ret<F32>().Attr<int32_t>("basis"));

// ------------------------------------------------------------------------------------------------ losses
// Tail of the four reference loss_fns after wcrbf.apply (train_nmpc_frenet.py:343-384,399-413; train_nmpc.py:260-293,
// 306-413): losses (2,) = [pred_loss, int_loss] of this shard and the cotangent dy (B,O), one launch.
// dyn_params: the 13 values of scripts/train_nmpc_frenet.py:322-338 (a host array for the C entry point).
static ffi::Error DynamicsLossImpl(cudaStream_t s, F32 x, F32 y_pred, F32 y, RF32 losses, RF32 dy, RU8 ws, int32_t mode,
                                   int64_t global_batch, ffi::Span<const float> dyn_params) {
  if (dyn_params.size() != 13) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "dyn_params must hold 13 values");
  const auto xd = x.dimensions();
  const auto yd = y.dimensions();
  return Check(irbfn_dynamics_loss(mode, x.typed_data(), y_pred.typed_data(), y.typed_data(), dyn_params.begin(),
                                   dy->typed_data(), losses->typed_data(), xd[0], (int)xd[1], (int)yd[1], global_batch,
                                   ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnDynamicsLoss, DynamicsLossImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>().Arg<F32>()
        .Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>()
        .Attr<int32_t>("mode").Attr<int64_t>("global_batch").Attr<ffi::Span<const float>>("dyn_params"));

// optax.l2_loss(pred, target).mean() (kind 0) or jnp.abs(pred - target).mean() (kind 1) with its cotangent.
static ffi::Error PredictionLossImpl(cudaStream_t s, F32 y_pred, F32 y, RF32 loss, RF32 dy, RU8 ws, int32_t kind,
                                     int64_t global_batch) {
  const auto yd = y.dimensions();
  return Check(irbfn_prediction_loss(kind, y_pred.typed_data(), y.typed_data(), dy->typed_data(), loss->typed_data(),
                                     yd[0], (int)yd[1], global_batch, ws->typed_data(), ws->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnPredictionLoss, PredictionLossImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Arg<F32>()
        .Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>()
        .Attr<int32_t>("kind").Attr<int64_t>("global_batch"));

// integrate_frenet_mult / dynamic_frenet_onestep_aux / integrate_st_mult / dynamic_st_onestep_aux (dynamics.py).
static ffi::Error IntegrateImpl(cudaStream_t s, F32 x_and_u, RF32 states, int32_t model, ffi::Span<const float> dyn_params) {
  if (dyn_params.size() != 13) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "dyn_params must hold 13 values");
  const auto d = x_and_u.dimensions();
  return Check(irbfn_integrate(model, x_and_u.typed_data(), dyn_params.begin(), states->typed_data(), d[0], (int)d[1], s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(IrbfnIntegrate, IntegrateImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Arg<F32>().Ret<F32>().Attr<int32_t>("model").Attr<ffi::Span<const float>>("dyn_params"));
