"""Flax / JAX side of the drop-in: the reference's modules (src/irbfn_mpc/model.py) backed by the sm_100a
kernels through XLA FFI custom calls with custom VJPs.

UNVERIFIED -- neither the build container nor the GPU box has jax / jaxlib / flax, so this
module has never been imported.  It is complete source for the binding BASELINE.json asks for, written against
the public jax.ffi API (jax >= 0.4.38); importing it without JAX raises ImportError, and nothing else in this
repository imports it.  The tested host mirror of the same interface is irbfn_b200/model.py (torch).

What a maintainer of hzheng40/irbfn does:
  1. build libirbfn_b200.so (python -m irbfn_b200.build) and libirbfn_ffi.so (irbfn_ffi.cc, command in its header);
  2. `from irbfn_b200.ffi.irbfn_flax import WCRBFNet, DeeperWCRBFNet, ClusterWCRBFNet` instead of
     `from irbfn_mpc.model import ...` (scripts/train_nmpc.py:24-34, irbfn_planner.py:17-27): same constructor
     arguments, same `init` / `apply`, same parameter tree, so `TrainState.create(apply_fn=wcrbf.apply, ...)`,
     `jax.value_and_grad(loss_fn)`, `jax.jit`, `checkpoints.restore_checkpoint(ckpt_dir, target=state)` and the
     shipped checkpoints keep working.

Parameter tree (must equal what the reference's double nn.vmap creates, model.py:143-167):
    {'params': {'rbf_list': {'centers': (R,K,D), 'log_sigs': (R,K)}, 'linear': {'kernel': (K,O), 'bias': (O,)}}}
  DeeperWCRBFNet adds linear_pre1 (K,64), linear_pre2 (64,64), linear (64,O) (model.py:254-256);
  ClusterWCRBFNet adds cluster {'kernel': (D,R), 'bias': (R,)} (model.py:393).
"""
from __future__ import annotations

import ctypes
import functools
import os
from typing import Callable, Optional, Sequence

import numpy as np

import jax                       # noqa: E402  (ImportError here = no JAX on this machine: use irbfn_b200.model)
import jax.numpy as jnp          # noqa: E402
from flax import linen as nn     # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
_core = ctypes.CDLL(os.path.join(HERE, "..", "libirbfn_b200.so"))
_ffi = ctypes.CDLL(os.path.join(HERE, "libirbfn_ffi.so"))

for _name in ("IrbfnFwd", "IrbfnFwdGateOut", "IrbfnFwdHidden", "IrbfnFwdExtGate", "IrbfnSoftmaxGate", "IrbfnFwdSparse",
              "IrbfnBwd", "IrbfnBwdParams", "IrbfnBwdHidden", "IrbfnBwdExtGate", "IrbfnJacobian", "IrbfnDynamicsLoss",
              "IrbfnPredictionLoss", "IrbfnIntegrate"):
    jax.ffi.register_ffi_target(_name, jax.ffi.pycapsule(getattr(_ffi, _name)), platform="CUDA")

_core.irbfn_forward_workspace_bytes.restype = ctypes.c_size_t
_core.irbfn_forward_workspace_bytes.argtypes = [ctypes.c_int64] + [ctypes.c_int] * 5
_core.irbfn_backward_workspace_bytes.restype = ctypes.c_size_t
_core.irbfn_backward_workspace_bytes.argtypes = [ctypes.c_int64] + [ctypes.c_int] * 4 + [ctypes.c_uint]
_core.irbfn_forward_sparse_workspace_bytes.restype = ctypes.c_size_t
_core.irbfn_forward_sparse_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int]
_core.irbfn_dynamics_loss_workspace_bytes.restype = ctypes.c_size_t
_core.irbfn_dynamics_loss_workspace_bytes.argtypes = [ctypes.c_int64]
_core.irbfn_prediction_loss_workspace_bytes.restype = ctypes.c_size_t
_core.irbfn_prediction_loss_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int]

BASIS = {"gaussian": 0, "inverse_quadratic": 1}          # flax_rbf.py:34-37, 49-52; others: no kernel -> KeyError
GRAD_CENTERS, GRAD_LOG_SIGS, GRAD_KERNEL, GRAD_X, GRAD_GATE = 0x01, 0x02, 0x04, 0x08, 0x10
GRAD_ALL = GRAD_CENTERS | GRAD_LOG_SIGS | GRAD_KERNEL | GRAD_X
_S = jax.ShapeDtypeStruct
_f32 = lambda a: jnp.asarray(a, jnp.float32)             # cast at the call boundary; the tree is never mutated


def _scratch(nbytes: int):
    return _S((max(int(nbytes), 1),), jnp.uint8)         # XLA owns the scratch: an extra uint8 result


def gate_constants(lower_bounds, upper_bounds, delta, dimension_ranges, activation_idx):
    """Pack a module's static gate attributes (model.py:113-125) once: five small constant arrays.
    activation_idx is only used for its length (model.py:74-81, 128: the first S input columns are gated)."""
    S = len(activation_idx)
    lb, ub = list(lower_bounds)[:S], list(upper_bounds)[:S]
    offs = np.cumsum([0] + [len(b) for b in lb]).astype(np.int32)
    return (jnp.asarray(np.concatenate([np.asarray(b, np.float32) for b in lb])),
            jnp.asarray(np.concatenate([np.asarray(b, np.float32) for b in ub])),
            jnp.asarray(offs), jnp.asarray(np.asarray(delta, np.float32)[:S]),
            jnp.asarray(np.asarray(dimension_ranges, np.int32)[:, :S]))


# ------------------------------------------------------------------------------------------------------------------
# WCRBFNet: y = irbfn_apply(x, centers, log_sigs, W, b)            custom VJP: residuals are x + params (phi recomputed)
# ------------------------------------------------------------------------------------------------------------------
@functools.partial(jax.custom_vjp, nondiff_argnums=(5, 6, 7))
def irbfn_apply(x, centers, log_sigs, W, b, gate_consts, basis, gate_tol=-1.0):
    # gate_tol >= 0: jax.grad through this call uses the gate-sparse VJP (irbfn_backward_sparse); < 0: dense
    B, D = x.shape
    R, K, _ = centers.shape
    O = W.shape[1]
    y, _ = jax.ffi.ffi_call("IrbfnFwd", (_S((B, O), jnp.float32),
                                         _scratch(_core.irbfn_forward_workspace_bytes(B, R, K, D, O, 1))),
                            vmap_method="broadcast_all")(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), _f32(b), *gate_consts, basis=np.int32(basis))
    return y


def _apply_fwd(x, centers, log_sigs, W, b, gate_consts, basis, gate_tol=-1.0):
    return irbfn_apply(x, centers, log_sigs, W, b, gate_consts, basis, gate_tol), (x, centers, log_sigs, W)


def _apply_bwd(gate_consts, basis, gate_tol, res, dy):
    x, centers, log_sigs, W = res
    B, D = x.shape
    R, K, _ = centers.shape
    O = W.shape[1]
    outs = (_S(centers.shape, jnp.float32), _S(log_sigs.shape, jnp.float32), _S(W.shape, jnp.float32),
            _S((O,), jnp.float32), _S(x.shape, jnp.float32),
            _scratch(_core.irbfn_backward_workspace_bytes(B, R, K, D, O, GRAD_ALL)))
    dC, dLS, dW, db, dx, _ = jax.ffi.ffi_call("IrbfnBwd", outs)(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), *gate_consts, _f32(dy),
        basis=np.int32(basis), grad_mask=np.int32(GRAD_ALL), gate_tol=np.float32(gate_tol))
    return dx, dC, dLS, dW, db


irbfn_apply.defvjp(_apply_fwd, _apply_bwd)


def irbfn_apply_sparse(x, centers, log_sigs, W, b, gate_consts, basis, tol_abs):
    """Inference-only gate-sparse forward (|y - dense| <= tol_abs per output); not differentiable."""
    B, D = x.shape
    O = W.shape[1]
    y, _ = jax.ffi.ffi_call("IrbfnFwdSparse", (_S((B, O), jnp.float32),
                                               _scratch(_core.irbfn_forward_sparse_workspace_bytes(B, int(gate_consts[0].shape[0])))))(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), _f32(b), *gate_consts, basis=np.int32(basis),
        tol_abs=np.float32(tol_abs))
    return y


# ------------------------------------------------------------------------------------------------------------------
# DeeperWCRBFNet front end: hidden (B,K) = sum_r gamma_r phi_r     (model.py:259-283)
# ------------------------------------------------------------------------------------------------------------------
@functools.partial(jax.custom_vjp, nondiff_argnums=(3, 4))
def irbfn_hidden(x, centers, log_sigs, gate_consts, basis):
    B = x.shape[0]
    K = centers.shape[1]
    return jax.ffi.ffi_call("IrbfnFwdHidden", _S((B, K), jnp.float32), vmap_method="broadcast_all")(
        _f32(x), _f32(centers), _f32(log_sigs), *gate_consts, basis=np.int32(basis))


def _hidden_fwd(x, centers, log_sigs, gate_consts, basis):
    return irbfn_hidden(x, centers, log_sigs, gate_consts, basis), (x, centers, log_sigs)


def _hidden_bwd(gate_consts, basis, res, dhidden):
    x, centers, log_sigs = res
    B, D = x.shape
    R, K, _ = centers.shape
    mask = GRAD_CENTERS | GRAD_LOG_SIGS | GRAD_X
    outs = (_S(centers.shape, jnp.float32), _S(log_sigs.shape, jnp.float32), _S(x.shape, jnp.float32),
            _scratch(_core.irbfn_backward_workspace_bytes(B, R, K, D, 0, mask)))
    dC, dLS, dx, _ = jax.ffi.ffi_call("IrbfnBwdHidden", outs)(
        _f32(x), _f32(centers), _f32(log_sigs), *gate_consts, _f32(dhidden), basis=np.int32(basis))
    return dx, dC, dLS


irbfn_hidden.defvjp(_hidden_fwd, _hidden_bwd)


# ------------------------------------------------------------------------------------------------------------------
# ClusterWCRBFNet: blend + readout with an externally computed gate (B,R)   (model.py:395-414)
# ------------------------------------------------------------------------------------------------------------------
@functools.partial(jax.custom_vjp, nondiff_argnums=(6,))
def irbfn_apply_gated(x, centers, log_sigs, W, b, gate, basis):
    B, D = x.shape
    R, K, _ = centers.shape
    O = W.shape[1]
    y, _ = jax.ffi.ffi_call("IrbfnFwdExtGate", (_S((B, O), jnp.float32),
                                                _scratch(_core.irbfn_forward_workspace_bytes(B, R, K, D, O, 0))),
                            vmap_method="broadcast_all")(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), _f32(b), _f32(gate), basis=np.int32(basis))
    return y


def _gated_fwd(x, centers, log_sigs, W, b, gate, basis):
    return irbfn_apply_gated(x, centers, log_sigs, W, b, gate, basis), (x, centers, log_sigs, W, gate)


def _gated_bwd(basis, res, dy):
    x, centers, log_sigs, W, gate = res
    B, D = x.shape
    R, K, _ = centers.shape
    O = W.shape[1]
    mask = GRAD_ALL | GRAD_GATE
    outs = (_S(centers.shape, jnp.float32), _S(log_sigs.shape, jnp.float32), _S(W.shape, jnp.float32),
            _S((O,), jnp.float32), _S(x.shape, jnp.float32), _S(gate.shape, jnp.float32),
            _scratch(_core.irbfn_backward_workspace_bytes(B, R, K, D, O, mask)))
    dC, dLS, dW, db, dx, dgate, _ = jax.ffi.ffi_call("IrbfnBwdExtGate", outs)(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), _f32(gate), _f32(dy), basis=np.int32(basis))
    return dx, dC, dLS, dW, db, dgate        # dx: through the RBF layers; the gate's own x-dependence flows via dgate


irbfn_apply_gated.defvjp(_gated_fwd, _gated_bwd)


# ------------------------------------------------------------------------------------------------------------------
# Jacobian, losses, rollouts
# ------------------------------------------------------------------------------------------------------------------
def irbfn_jacobian(x, centers, log_sigs, W, b, gate_consts, basis):
    """(jac (B,O,D), y (B,O)) == jax.vmap(jax.jacrev(lambda q: apply(params, q[None])[0]))(x), one pass."""
    B, D = x.shape
    O = W.shape[1]
    return jax.ffi.ffi_call("IrbfnJacobian", (_S((B, O, D), jnp.float32), _S((B, O), jnp.float32)))(
        _f32(x), _f32(centers), _f32(log_sigs), _f32(W), _f32(b), *gate_consts, basis=np.int32(basis))


DYN_PARAMS = (1.0489, 1.0489, 0.04712, 0.15875, 0.17145, 4.718, 4.718, 0.074, 0.1, 3.2, 9.51, 0.4189, 7.0)
LOSS_MODES = {"frenet_oneint": 0, "frenet_fullint": 1, "st_oneint": 2, "st_fullint": 3}


def _dynamics_loss_call(mode, x, y_pred, y, dyn_params, global_batch):
    B, O = y.shape
    losses, dy, _ = jax.ffi.ffi_call("IrbfnDynamicsLoss", (_S((2,), jnp.float32), _S((B, O), jnp.float32),
                                                           _scratch(_core.irbfn_dynamics_loss_workspace_bytes(B))))(
        _f32(x), _f32(y_pred), _f32(y), mode=np.int32(mode), global_batch=np.int64(global_batch),
        dyn_params=np.asarray(dyn_params, np.float32))
    return losses, dy


@functools.partial(jax.custom_vjp, nondiff_argnums=(0, 4, 5))
def dynamics_loss(mode, x, y_pred, y, dyn_params=DYN_PARAMS, global_batch=None):
    """(loss, (pred_loss, int_loss)) of the reference's training steps after `wcrbf.apply`
    (scripts/train_nmpc_frenet.py:343-384, 399-413; scripts/train_nmpc.py:260-293, 306-413), differentiable in
    y_pred.  Use inside loss_fn:  `loss, aux = dynamics_loss(LOSS_MODES['frenet_fullint'], x, wcrbf.apply(p, x), y)`."""
    losses, _ = _dynamics_loss_call(mode, x, y_pred, y, dyn_params, global_batch or y.shape[0])
    return losses.sum(), (losses[0], losses[1])


def _dl_fwd(mode, x, y_pred, y, dyn_params, global_batch):
    losses, dy = _dynamics_loss_call(mode, x, y_pred, y, dyn_params, global_batch or y.shape[0])
    return (losses.sum(), (losses[0], losses[1])), dy


def _dl_bwd(mode, dyn_params, global_batch, dy, ct):
    return None, ct[0] * dy, None           # x and y are data; the aux terms carry no cotangent in the reference


dynamics_loss.defvjp(_dl_fwd, _dl_bwd)


def integrate(model, x_and_u, dyn_params=DYN_PARAMS):
    """integrate_frenet_mult (0) / dynamic_frenet_onestep_aux (1) / integrate_st_mult (2) / dynamic_st_onestep_aux (3)
    of src/irbfn_mpc/dynamics.py, forward only (what the planners call, irbfn_planner.py:212-214)."""
    B = x_and_u.shape[0]
    shape = {0: (B, 5, 8), 1: (B, 6), 2: (B, 5, 7), 3: (B, 7)}[int(model)]
    return jax.ffi.ffi_call("IrbfnIntegrate", _S(shape, jnp.float32))(
        _f32(x_and_u), model=np.int32(model), dyn_params=np.asarray(dyn_params, np.float32))


# ------------------------------------------------------------------------------------------------------------------
# Flax modules: same attributes, defaults and parameter tree as src/irbfn_mpc/model.py
# ------------------------------------------------------------------------------------------------------------------
class _RBFParams(nn.Module):
    """Holds what the reference's double nn.vmap of RBFLayer creates under the scope name it is assigned to
    (`self.rbf_list = ...`, model.py:161): centers (R,K,D) ~ normal(1.0), log_sigs (R,K) = 0 (flax_rbf.py:247-256).
    Fixed-centre modules have no `centers` leaf, fixed-width modules no leaves at all."""
    num_regions: int
    num_kernels: int
    in_features: int
    centers_values: Optional[jnp.ndarray] = None
    fixed_centers: bool = False
    fixed_width: bool = False

    @nn.compact
    def __call__(self):
        R, K, D = self.num_regions, self.num_kernels, self.in_features
        if self.centers_values is None:
            centers = self.param("centers", nn.initializers.normal(1.0), (R, K, D))
        else:
            cv = jnp.broadcast_to(jnp.asarray(self.centers_values, jnp.float32), (R, K, D))
            centers = cv if self.fixed_centers else self.param("centers", lambda key, shape: cv, (R, K, D))
        if self.centers_values is not None and self.fixed_centers and self.fixed_width:
            log_sigs = jnp.zeros((R, K), jnp.float32)
        else:
            log_sigs = self.param("log_sigs", nn.initializers.constant(0.0), (R, K))
        return centers, log_sigs


class _DenseParams(nn.Module):
    """The leaves nn.Dense(features) creates (kernel lecun-normal, bias zeros) without applying it: the readout
    runs inside the fused kernel."""
    features: int
    in_features: int

    @nn.compact
    def __call__(self):
        kernel = self.param("kernel", nn.initializers.lecun_normal(), (self.in_features, self.features))
        bias = self.param("bias", nn.initializers.zeros_init(), (self.features,))
        return kernel, bias


class WCRBFNet(nn.Module):
    """Drop-in for src/irbfn_mpc/model.py:98-198."""
    in_features: int
    out_features: int
    num_kernels: int
    basis_func: Callable
    num_regions: int
    lower_bounds: Sequence[Sequence[float]]
    upper_bounds: Sequence[Sequence[float]]
    dimension_ranges: Sequence[Sequence[float]]
    activation_idx: Sequence[int]
    delta: Sequence[float]
    centers: Optional[jax.Array] = None
    fixed_centers: Optional[bool] = False
    fixed_width: Optional[bool] = False
    gate_tol: float = -1.0     # not in the reference: >= 0 selects the gate-sparse VJP (irbfn_backward_sparse)

    def setup(self):
        self.num_split_dimensions = len(self.activation_idx)
        self.rbf_list = _RBFParams(self.num_regions, self.num_kernels, self.in_features, self.centers,
                                   bool(self.fixed_centers), bool(self.fixed_width))
        self.linear = _DenseParams(self.out_features, self.num_kernels)

    def __call__(self, x):
        centers, log_sigs = self.rbf_list()
        kernel, bias = self.linear()
        consts = gate_constants(self.lower_bounds, self.upper_bounds, self.delta, self.dimension_ranges,
                                self.activation_idx)
        return irbfn_apply(x, centers, log_sigs, kernel, bias, consts, BASIS[self.basis_func.__name__],
                           float(self.gate_tol))


class DeeperWCRBFNet(nn.Module):
    """Drop-in for src/irbfn_mpc/model.py:201-289: fused front end, the 64-wide tail stays in XLA."""
    in_features: int
    out_features: int
    num_kernels: int
    basis_func: Callable
    num_regions: int
    lower_bounds: Sequence[Sequence[float]]
    upper_bounds: Sequence[Sequence[float]]
    dimension_ranges: Sequence[Sequence[float]]
    activation_idx: Sequence[int]
    delta: Sequence[float]

    def setup(self):
        self.num_split_dimensions = len(self.activation_idx)
        self.rbf_list = _RBFParams(self.num_regions, self.num_kernels, self.in_features)
        self.linear_pre1 = nn.Dense(64)
        self.linear_pre2 = nn.Dense(64)
        self.linear = nn.Dense(self.out_features)

    def __call__(self, x):
        centers, log_sigs = self.rbf_list()
        consts = gate_constants(self.lower_bounds, self.upper_bounds, self.delta, self.dimension_ranges,
                                self.activation_idx)
        rbf_out = irbfn_hidden(x, centers, log_sigs, consts, BASIS[self.basis_func.__name__])
        out_pre1 = self.linear_pre1(rbf_out)                       # model.py:285
        out_pre2 = self.linear_pre2(nn.relu(out_pre1))             # model.py:286
        return self.linear(nn.relu(out_pre2))                      # model.py:287


class ClusterWCRBFNet(nn.Module):
    """Drop-in for src/irbfn_mpc/model.py:341-414: returns (out, logits)."""
    in_features: int
    out_features: int
    num_kernels: int
    basis_func: Callable
    num_regions: int

    def setup(self):
        self.rbf_list = _RBFParams(self.num_regions, self.num_kernels, self.in_features)
        self.linear = _DenseParams(self.out_features, self.num_kernels)
        self.cluster = nn.Dense(self.num_regions)

    def __call__(self, x):
        centers, log_sigs = self.rbf_list()
        kernel, bias = self.linear()
        logits = self.cluster(x)                                   # model.py:402
        cluster_ind = nn.softmax(logits)                           # model.py:403 (XLA; its VJP receives dgate)
        out = irbfn_apply_gated(x, centers, log_sigs, kernel, bias, cluster_ind, BASIS[self.basis_func.__name__])
        return out, logits
