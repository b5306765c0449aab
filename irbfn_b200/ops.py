"""Device-level operators: thin wrappers that hand torch CUDA buffers to the C ABI.

torch is used for device memory and streams only; all arithmetic happens in
libirbfn_b200.so.  Every function requires contiguous fp32 CUDA tensors and raises on
anything else -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import (BASIS_GAUSSIAN, BASIS_INVERSE_QUADRATIC, GRAD_CENTERS, GRAD_GATE, GRAD_KERNEL,
                   GRAD_LOG_SIGS, GRAD_PARAMS, GRAD_X)

__all__ = [
    "GatePack", "forward", "forward_sparse", "backward", "region_gate", "softmax_gate", 
    "dynamics_loss", "integrate", "default_dyn_params", "jacobian", "fused_update_", "fused_update_max_n", "lattice_queries",
    "gather_rows", "GridKeys", "grid_index", "nearest_row", "forward_path",
    "BASIS_GAUSSIAN", "BASIS_INVERSE_QUADRATIC", "GRAD_CENTERS", "GRAD_LOG_SIGS", "GRAD_KERNEL",
    "GRAD_X", "GRAD_GATE", "GRAD_PARAMS",
]


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk(t: Optional[torch.Tensor], name: str, shape=None, dtype=torch.float32):
    if t is None:
        return
    if not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError("%s has shape %s, expected %s" % (name, tuple(t.shape), tuple(shape)))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class GatePack:
    """The static tanh-gate attributes of a module (lower_bounds, upper_bounds, delta,
    dimension_ranges; reference model.py:113-125) packed flat on one device, plus the
    ``IrbfnGateDesc`` that points at them."""

    def __init__(self, lower_bounds: Sequence[Sequence[float]],
                 upper_bounds: Sequence[Sequence[float]], delta: Sequence[float],
                 dimension_ranges: Sequence[Sequence[int]], device):
        S = len(lower_bounds)
        if S == 0 or len(upper_bounds) != S or len(delta) < S:
            raise ValueError("lower_bounds, upper_bounds and delta must describe the same split dims")
        if S > _lib.MAX_S:
            raise ValueError("at most %d split dimensions are supported" % _lib.MAX_S)
        offs = [0]
        for d in range(S):
            if len(lower_bounds[d]) != len(upper_bounds[d]) or len(lower_bounds[d]) == 0:
                raise ValueError("bounds of split dim %d are ragged or empty" % d)
            offs.append(offs[-1] + len(lower_bounds[d]))
        rng = np.asarray(dimension_ranges, dtype=np.int64).reshape(-1, S) if len(dimension_ranges) \
            else np.zeros((0, S), np.int64)
        for d in range(S):
            if rng.size and (rng[:, d].min() < 0 or rng[:, d].max() >= len(lower_bounds[d])):
                raise IndexError("dimension_ranges indexes outside the bounds of split dim %d" % d)
        self.num_split = S
        self.num_ranges = int(rng.shape[0])
        self.num_bounds = offs[-1]
        self.device = torch.device(device)
        f32 = lambda v: torch.tensor(np.asarray(v, dtype=np.float32), device=self.device)
        i32 = lambda v: torch.tensor(np.asarray(v, dtype=np.int32), device=self.device)
        self.lower = f32(np.concatenate([np.asarray(b, np.float64) for b in lower_bounds]))
        self.upper = f32(np.concatenate([np.asarray(b, np.float64) for b in upper_bounds]))
        self.offset = i32(offs)
        self.delta = f32(list(delta)[:S])
        self.ranges = i32(rng.reshape(-1) if rng.size else np.zeros(1))
        self.desc = _lib.GateDesc(S, self.num_ranges, self.num_bounds, self.lower.data_ptr(),
                                  self.upper.data_ptr(), self.offset.data_ptr(),
                                  self.delta.data_ptr(), self.ranges.data_ptr())

    def ref(self):
        return C.byref(self.desc)


def forward_path(B: int, R: int, K: int, D: int, O: int, y_only: bool = True) -> str:
    """Which kernel ``forward`` runs for a shape (irbfn_forward_path): "tiled" (throughput kernel), "split" (the same
    with the kernels split over CTAs, small batches) or "small" (the pair-parallel latency kernel of pred_step)."""
    rc = _lib.load().irbfn_forward_path(int(B), int(R), int(K), int(D), int(O), 1 if y_only else 0)
    _lib.check(rc if rc < 0 else 0, "irbfn_forward_path")
    return ("tiled", "split", "small")[rc]


def forward(x, centers, log_sigs, kernel=None, bias=None, *, gate=None, gd: GatePack = None,
            basis: int = BASIS_GAUSSIAN, want_y: bool = True, want_hidden: bool = False,
            out=None, split: bool = True, gate_out=None):
    """y (B,O) and/or hidden (B,K) of the fused IRBFN forward (irbfn_forward).  ``split=False`` withholds
    the workspace, so a small batch runs as ONE CTA per query tile instead of spreading its kernels over
    the device -- what a caller wants when it keeps the device busy with several launches at once.
    ``gate_out`` (B,R): the tanh gate is written there too (irbfn_forward_gate; training steps hand it to
    ``backward`` as its external gate)."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    _chk(centers, "centers")
    R, K, Dc = centers.shape
    if Dc != D:
        raise ValueError("centers feature count %d != x feature count %d" % (Dc, D))
    _chk(log_sigs, "log_sigs", (R, K))
    O = 0
    y = hidden = None
    if want_y:
        _chk(kernel, "kernel")
        if kernel is None or bias is None:
            raise ValueError("kernel and bias are required for y")
        O = kernel.shape[1]
        _chk(kernel, "kernel", (K, O))
        _chk(bias, "bias", (O,))
        y = out if out is not None else torch.empty((B, O), device=x.device, dtype=torch.float32)
        _chk(y, "out", (B, O))
    if want_hidden:
        hidden = torch.empty((B, K), device=x.device, dtype=torch.float32)
    if gate is not None:
        _chk(gate, "gate", (B, R))
    if basis not in (BASIS_GAUSSIAN, BASIS_INVERSE_QUADRATIC):
        _lib.check(-3, "irbfn_forward")
    if (gate is None) == (gd is None):
        _lib.check(-5, "irbfn_forward")
    if B == 0:  # empty tensors have no storage to point at
        return (y, hidden) if want_y and want_hidden else (y if want_y else hidden)
    with torch.cuda.device(x.device):
        nbytes = lib.irbfn_forward_workspace_bytes(B, R, K, D, O, 1 if gd is not None else 0) if split else 0
        ws = _workspace(nbytes, x.device)       # partial readouts of the small-batch split-K path
        if gate_out is not None:
            if gd is None or not want_y or want_hidden:
                raise ValueError("gate_out needs the tanh gate (gd) and the plain y output")
            _chk(gate_out, "gate_out", (B, R))
            rc = lib.irbfn_forward_gate(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias), gd.ref(),
                                        _ptr(y), _ptr(gate_out), B, R, K, D, O, basis, _ptr(ws), nbytes, _stream())
        else:
            rc = lib.irbfn_forward(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias),
                                   _ptr(gate), gd.ref() if gd is not None else None, _ptr(y),
                                   _ptr(hidden), B, R, K, D, O, basis, _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_forward")
    if want_y and want_hidden:
        return y, hidden
    return y if want_y else hidden


def forward_sparse(x, centers, log_sigs, kernel, bias, *, gd: GatePack, basis: int = BASIS_GAUSSIAN,
                   tol_abs: float = 0.0, out=None, counters=None, gate_tol: Optional[float] = None):
    """Gate-sparse, region-binned forward (irbfn_forward_sparse): y within ``tol_abs`` (absolute, per output) of
    ``forward``.  ``counters``: optional int64 CUDA tensor of 2 elements receiving the number of (warp, region)
    evaluations skipped / done.  ``gate_tol`` instead of ``tol_abs``: the gate threshold itself
    (irbfn_forward_sparse_gate: a local network is skipped for a query when its gate is <= gate_tol)."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    _chk(centers, "centers")
    R, K, Dc = centers.shape
    if Dc != D:
        raise ValueError("centers feature count %d != x feature count %d" % (Dc, D))
    _chk(log_sigs, "log_sigs", (R, K))
    O = kernel.shape[1]
    _chk(kernel, "kernel", (K, O))
    _chk(bias, "bias", (O,))
    y = out if out is not None else torch.empty((B, O), device=x.device, dtype=torch.float32)
    _chk(y, "out", (B, O))
    if gd is None:
        _lib.check(-5, "irbfn_forward_sparse")
    if counters is not None and (counters.dtype != torch.int64 or counters.numel() != 2 or not counters.is_cuda):
        raise ValueError("counters must be an int64 CUDA tensor with 2 elements")
    if B == 0:
        return y
    with torch.cuda.device(x.device):
        nbytes = lib.irbfn_forward_sparse_workspace_bytes(B, gd.desc.num_bounds)
        ws = _workspace(nbytes, x.device)
        fn = lib.irbfn_forward_sparse if gate_tol is None else lib.irbfn_forward_sparse_gate
        rc = fn(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias), gd.ref(), _ptr(y), B, R, K, D, O, basis,
                float(tol_abs if gate_tol is None else gate_tol), _ptr(counters), _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_forward_sparse")
    return y


def forward_streamed(x, centers, log_sigs, kernel, bias, y, ready, queries_per_flag: int, *, gd: GatePack = None,
                     gate=None, basis: int = BASIS_GAUSSIAN, status=None):
    """Forward over rows that are still being uploaded (irbfn_forward_streamed).  ``x`` (B,D) is the DEVICE buffer the
    caller's copy streams are filling piece by piece; ``ready`` a zeroed int32/uint32 CUDA tensor with one flag per
    piece of ``queries_per_flag`` rows (a multiple of 1,024) plus ONE more word (the kernel's tile counter); a flag is
    set behind its piece's upload (``upload_streamed``).  ``y`` (B,O): a CUDA tensor or a PINNED CPU tensor (the kernel
    then stores straight into host memory)."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    _chk(centers, "centers")
    R, K, _ = centers.shape
    _chk(log_sigs, "log_sigs", (R, K))
    O = kernel.shape[1]
    _chk(kernel, "kernel", (K, O))
    _chk(bias, "bias", (O,))
    if not (isinstance(y, torch.Tensor) and y.dtype == torch.float32 and y.is_contiguous() and tuple(y.shape) == (B, O)
            and (y.is_cuda or y.is_pinned())):
        raise ValueError("y must be a contiguous float32 (B,O) CUDA tensor or pinned CPU tensor")
    if not (ready.is_cuda and ready.dtype in (torch.int32, torch.uint32) and ready.is_contiguous()):
        raise ValueError("ready must be a contiguous 32-bit integer CUDA tensor")
    if (ready.numel() - 1) * int(queries_per_flag) < B:
        raise ValueError("ready has %d flags of %d rows (+ the counter word): fewer than B = %d"
                         % (ready.numel() - 1, queries_per_flag, B))
    if B == 0:
        return y
    with torch.cuda.device(x.device):
        rc = lib.irbfn_forward_streamed(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias), _ptr(gate),
                                        gd.ref() if gd is not None else None, C.c_void_p(y.data_ptr()), B, R, K, D, O,
                                        basis, _ptr(ready), int(queries_per_flag), _ptr(status), _stream())
    _lib.check(rc, "irbfn_forward_streamed")
    return y


def upload_streamed_available() -> bool:
    """Whether the driver offers the stream-ordered 32-bit store ``upload_streamed`` sets its flags with."""
    return bool(_lib.load().irbfn_upload_streamed_available())


def upload_streamed(x_host, x_dev, ready, queries_per_flag: int, copy_streams):
    """The upload side of ``forward_streamed`` (irbfn_upload_streamed): pinned ``x_host`` (B,D) -> ``x_dev`` in pieces
    of ``queries_per_flag`` rows alternating over ``copy_streams`` (torch.cuda.Stream objects), each followed on its
    stream by the store that sets the piece's flag."""
    lib = _lib.load()
    if not (x_host.is_pinned() and x_host.dtype == torch.float32 and x_host.is_contiguous()):
        raise ValueError("x_host must be a contiguous pinned float32 CPU tensor")
    _chk(x_dev, "x_dev", tuple(x_host.shape))
    B, D = x_host.shape
    arr = (C.c_void_p * len(copy_streams))(*[s.cuda_stream for s in copy_streams])
    with torch.cuda.device(x_dev.device):
        rc = lib.irbfn_upload_streamed(C.c_void_p(x_host.data_ptr()), _ptr(x_dev), B, D, _ptr(ready),
                                       int(queries_per_flag), arr, len(copy_streams))
    _lib.check(rc, "irbfn_upload_streamed")


_ws_cache = {}


def _workspace(nbytes: int, device) -> Optional[torch.Tensor]:
    if nbytes == 0:
        return None
    if torch.cuda.is_current_stream_capturing():
        # under CUDA-graph capture the scratch comes from the capturing graph's own memory pool and is NOT cached:
        # every captured graph keeps its own workspace (torch's captures share one capture stream, so a cache keyed by
        # stream would hand two graphs the same buffer -- a race if they are ever replayed on different streams)
        return torch.empty(nbytes, device=device, dtype=torch.uint8)
    key = (torch.device(device), torch.cuda.current_stream(device).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(nbytes, device=device, dtype=torch.uint8)
        _ws_cache[key] = ws
    return ws


def backward(x, centers, log_sigs, kernel, dy=None, *, dhidden=None, gate=None,
             gd: GatePack = None, basis: int = BASIS_GAUSSIAN, grad_mask: int = GRAD_PARAMS,
             flat_out: Optional[torch.Tensor] = None, gate_tol: Optional[float] = None):
    """VJP of the forward (irbfn_backward).  Returns a dict with the requested leaves:
    centers, log_sigs, kernel, bias, x, gate.  If ``flat_out`` is given the parameter
    gradients are written into that flat buffer in the order [centers, log_sigs, kernel,
    bias] (the all-reduce payload) and the dict holds views of it.  ``gate_tol`` (>= 0): gate-sparse parameter
    gradients (irbfn_backward_sparse) -- terms of (query, region) pairs with gamma <= gate_tol are left out."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    _chk(centers, "centers")
    R, K, _ = centers.shape
    _chk(log_sigs, "log_sigs", (R, K))
    O = 0
    if dy is not None:
        _chk(kernel, "kernel")
        O = kernel.shape[1]
        _chk(dy, "dy", (B, O))
    if dhidden is not None:
        _chk(dhidden, "dhidden", (B, K))
    if gate is not None:
        _chk(gate, "gate", (B, R))
    dev = x.device
    out = {}
    sizes = [("centers", GRAD_CENTERS, (R, K, D)), ("log_sigs", GRAD_LOG_SIGS, (R, K)),
             ("kernel", GRAD_KERNEL, (K, O)), ("bias", GRAD_KERNEL, (O,))]
    if flat_out is not None:
        _chk(flat_out, "flat_out")
        off = 0
        for name, bit, shp in sizes:
            if grad_mask & bit:
                n = int(np.prod(shp))
                out[name] = flat_out[off:off + n].view(shp)
                off += n
        if off > flat_out.numel():
            raise ValueError("flat_out too small: need %d floats" % off)
    else:
        for name, bit, shp in sizes:
            if grad_mask & bit:
                out[name] = torch.empty(shp, device=dev, dtype=torch.float32)
    if grad_mask & GRAD_X:
        out["x"] = torch.empty((B, D), device=dev, dtype=torch.float32)
    if grad_mask & GRAD_GATE:
        out["gate"] = torch.empty((B, R), device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        nbytes = lib.irbfn_backward_workspace_bytes(B, R, K, D, O, grad_mask)
        ws = _workspace(nbytes, dev)
        head = (_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(gate),
                gd.ref() if gd is not None else None, _ptr(dy), _ptr(dhidden),
                _ptr(out.get("centers")), _ptr(out.get("log_sigs")),
                _ptr(out.get("kernel")), _ptr(out.get("bias")), _ptr(out.get("x")),
                _ptr(out.get("gate")), B, R, K, D, O, basis, grad_mask)
        if gate_tol is None:
            rc = lib.irbfn_backward(*head, _ptr(ws), nbytes, _stream())
        else:
            rc = lib.irbfn_backward_sparse(*head, float(gate_tol), _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_backward" if gate_tol is None else "irbfn_backward_sparse")
    return out


def region_gate(x, gd: GatePack, num_regions: int):
    """gamma (B,R) of the tanh region gate (irbfn_region_gate)."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    gamma = torch.empty((B, num_regions), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        rc = lib.irbfn_region_gate(_ptr(x), gd.ref(), _ptr(gamma), B, num_regions, D, _stream())
    _lib.check(rc, "irbfn_region_gate")
    return gamma


def softmax_gate(x, cluster_kernel, cluster_bias, want_logits: bool = True):
    """(gate, logits) of the ClusterWCRBFNet gate (irbfn_softmax_gate)."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    R = cluster_kernel.shape[1]
    _chk(cluster_kernel, "cluster kernel", (D, R))
    _chk(cluster_bias, "cluster bias", (R,))
    gate = torch.empty((B, R), device=x.device, dtype=torch.float32)
    logits = torch.empty((B, R), device=x.device, dtype=torch.float32) if want_logits else None
    with torch.cuda.device(x.device):
        rc = lib.irbfn_softmax_gate(_ptr(x), _ptr(cluster_kernel), _ptr(cluster_bias), _ptr(gate),
                                    _ptr(logits), B, D, R, _stream())
    _lib.check(rc, "irbfn_softmax_gate")
    return gate, logits


# --------------------------------------------------------------------------- loss-side dynamics
def default_dyn_params(mu: float = 1.0489, cs: float = 4.718):
    """dyn_params of the reference's training scripts (scripts/train_nmpc_frenet.py:322-338):
    [mu, m, I, lf, lr, C_Sf, C_Sr, h, dt, sv_max, a_max, s_max, v_max]."""
    return [mu, 1.0489, 0.04712, 0.15875, 0.17145, cs, cs, 0.074, 0.1, 3.2, 9.51, 0.4189, 7.0]


def _dyn_array(dyn_params):
    v = [float(t) for t in dyn_params]
    if len(v) != 13:
        raise ValueError("dyn_params must hold 13 values")
    return (C.c_float * 13)(*v)


def dynamics_loss(mode: int, x, y_pred, y, dyn_params, global_batch: Optional[int] = None,
                  dy: Optional[torch.Tensor] = None, losses: Optional[torch.Tensor] = None):
    """Fused loss tail of the reference's training steps (irbfn_dynamics_loss): returns
    (losses, dy) with losses = device tensor [pred_loss, int_loss] of this shard (means over
    ``global_batch``) and dy = dloss/dy_pred (B,O)."""
    lib = _lib.load()
    _chk(x, "x")
    _chk(y_pred, "y_pred")
    B, D = x.shape
    O = y_pred.shape[1]
    _chk(y_pred, "y_pred", (B, O))
    _chk(y, "y", (B, O))
    gb = int(global_batch) if global_batch is not None else B
    dy = dy if dy is not None else torch.empty((B, O), device=x.device, dtype=torch.float32)
    losses = losses if losses is not None else torch.empty(2, device=x.device, dtype=torch.float32)
    _chk(dy, "dy", (B, O))
    _chk(losses, "losses", (2,))
    with torch.cuda.device(x.device):
        nbytes = lib.irbfn_dynamics_loss_workspace_bytes(B)
        ws = _workspace(nbytes, x.device)
        rc = lib.irbfn_dynamics_loss(mode, _ptr(x), _ptr(y_pred), _ptr(y), _dyn_array(dyn_params), _ptr(dy),
                                     _ptr(losses), B, D, O, gb, _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_dynamics_loss")
    return losses, dy


def prediction_loss(y_pred, y, global_batch: Optional[int] = None, kind: str = "l2",
                    dy: Optional[torch.Tensor] = None, loss: Optional[torch.Tensor] = None):
    """Prediction loss and its cotangent in one launch (irbfn_prediction_loss): ``kind`` "l2" is
    optax.l2_loss(...).mean() = mean(0.5 d^2), "l1" is mean |d|; the mean runs over ``global_batch`` rows.
    Returns (loss, dy): a one-element device tensor (this shard's part) and dloss/dy_pred (B,O)."""
    lib = _lib.load()
    _chk(y_pred, "y_pred")
    B, O = y_pred.shape
    _chk(y, "y", (B, O))
    kinds = {"l2": 0, "l1": 1}
    if kind not in kinds:
        raise ValueError("loss must be 'l2' or 'l1'")
    gb = int(global_batch) if global_batch is not None else B
    dy = dy if dy is not None else torch.empty((B, O), device=y_pred.device, dtype=torch.float32)
    loss = loss if loss is not None else torch.empty(1, device=y_pred.device, dtype=torch.float32)
    _chk(dy, "dy", (B, O))
    _chk(loss, "loss", (1,))
    with torch.cuda.device(y_pred.device):
        nbytes = lib.irbfn_prediction_loss_workspace_bytes(B, O)
        ws = _workspace(nbytes, y_pred.device)
        rc = lib.irbfn_prediction_loss(kinds[kind], _ptr(y_pred), _ptr(y), _ptr(dy), _ptr(loss), B, O, gb,
                                       _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_prediction_loss")
    return loss, dy


def clip_adam_(params, grad, mu, nu, step_count, lr: float, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8,
               max_norm: float = 0.0, grad_norm: Optional[torch.Tensor] = None):
    """optax.chain(clip_by_global_norm(max_norm), adam(lr)) on flat fp32 device buffers, in place
    (irbfn_clip_adam: two launches).  ``step_count`` is a float64 device scalar, incremented by the call.
    Returns the pre-clip global gradient norm (one-element device tensor)."""
    lib = _lib.load()
    n = params.numel()
    for t, name in ((params, "params"), (grad, "grad"), (mu, "mu"), (nu, "nu")):
        _chk(t, name, (n,))
    _chk(step_count, "step_count", dtype=torch.float64)
    if step_count.numel() != 1:
        raise ValueError("step_count must be a one-element float64 CUDA tensor")
    grad_norm = grad_norm if grad_norm is not None else torch.empty(1, device=params.device, dtype=torch.float32)
    _chk(grad_norm, "grad_norm", (1,))
    with torch.cuda.device(params.device):
        nbytes = lib.irbfn_clip_adam_workspace_bytes(n)
        ws = _workspace(nbytes, params.device)
        rc = lib.irbfn_clip_adam(_ptr(params), _ptr(grad), _ptr(mu), _ptr(nu), n, _ptr(step_count), lr, b1, b2, eps,
                                 max_norm, _ptr(grad_norm), _ptr(ws), nbytes, _stream())
    _lib.check(rc, "irbfn_clip_adam")
    return grad_norm


def lattice_queries(axes, fixed, device=None):
    """(prod n_i, D) fp32 CUDA tensor: np.meshgrid(*axes.values(), indexing="ij") over the columns that are the
    keys of ``axes`` (insertion order, first varies slowest), the other columns held at ``fixed`` -- built on the
    device by irbfn_lattice_queries from a few hundred bytes of axis values."""
    import numpy as np
    lib = _lib.load()
    device = torch.device(device if device is not None else "cuda")
    cols = [int(c) for c in axes.keys()]
    vals = [np.asarray(axes[c], np.float32).ravel() for c in axes.keys()]
    D = len(fixed)
    if any(c < 0 or c >= D for c in cols) or len(set(cols)) != len(cols):
        raise ValueError("axis columns must be distinct and inside [0, %d)" % D)
    n = int(np.prod([len(v) for v in vals])) if vals else 1
    off = np.cumsum([0] + [len(v) for v in vals]).astype(np.int32)
    v_d = torch.from_numpy(np.concatenate(vals) if vals else np.zeros(1, np.float32)).to(device)
    o_d, c_d = torch.from_numpy(off).to(device), torch.from_numpy(np.asarray(cols or [0], np.int32)).to(device)
    f_d = torch.from_numpy(np.asarray(fixed, np.float32)).to(device)
    x = torch.empty((n, D), device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        rc = lib.irbfn_lattice_queries(_ptr(v_d), _ptr(o_d), _ptr(c_d), len(cols), _ptr(f_d), _ptr(x), n, D, _stream())
    _lib.check(rc, "irbfn_lattice_queries")
    return x


def fused_update_max_n() -> int:
    return int(_lib.load().irbfn_fused_update_max_n())


def fused_update_(params, mu, nu, reduced, n: int, n_extra: int, step_count, lr: float, b1: float = 0.9,
                  b2: float = 0.999, eps: float = 1e-8, max_norm: float = 0.0, grad_norm=None, peers=None,
                  rank: int = 0, world: int = 1, apply_update: bool = True, status=None):
    """Tail of a data-parallel training step in ONE launch (irbfn_fused_update): optional one-shot all-reduce
    over NVLink peer memory (``peers``: int64 device tensor of the ranks' symmetric-buffer addresses, see
    parallel.PeerBuffer), then clip_by_global_norm + Adam on the flat buffers.  ``reduced`` (n + n_extra floats)
    receives the summed gradient; with ``peers=None`` it already holds it."""
    lib = _lib.load()
    _chk(reduced, "reduced")
    if reduced.numel() < n + n_extra:
        raise ValueError("reduced holds %d floats, need %d" % (reduced.numel(), n + n_extra))
    if apply_update:
        for t, name in ((params, "params"), (mu, "mu"), (nu, "nu")):
            _chk(t, name, (n,))
        _chk(step_count, "step_count", dtype=torch.float64)
    if peers is not None and (peers.dtype != torch.int64 or peers.numel() != world or not peers.is_cuda):
        raise ValueError("peers must be an int64 CUDA tensor with one address per rank")
    with torch.cuda.device(reduced.device):
        rc = lib.irbfn_fused_update(_ptr(params), _ptr(mu), _ptr(nu), _ptr(reduced), n, n_extra, _ptr(peers), rank, world,
                                    1 if apply_update else 0, _ptr(step_count), lr, b1, b2, eps, max_norm,
                                    _ptr(grad_norm), _ptr(status), _stream())
    _lib.check(rc, "irbfn_fused_update")
    return grad_norm


_DYN_OUT = {_lib.DYN_FRENET_MULT: (5, 8), _lib.DYN_FRENET_ONESTEP_AUX: (6,), _lib.DYN_ST_MULT: (5, 7),
            _lib.DYN_ST_ONESTEP_AUX: (7,)}


def integrate(model: int, x_and_u, dyn_params):
    """integrate_frenet_mult / dynamic_frenet_onestep_aux / integrate_st_mult /
    dynamic_st_onestep_aux of the reference's dynamics.py on the device (irbfn_integrate)."""
    lib = _lib.load()
    _chk(x_and_u, "x_and_u")
    if model not in _DYN_OUT:
        raise ValueError("unknown dynamics model %r" % (model,))
    B, W = x_and_u.shape
    out = torch.empty((B,) + _DYN_OUT[model], device=x_and_u.device, dtype=torch.float32)
    if B == 0:
        return out
    with torch.cuda.device(x_and_u.device):
        rc = lib.irbfn_integrate(model, _ptr(x_and_u), _dyn_array(dyn_params), _ptr(out), B, W, _stream())
    _lib.check(rc, "irbfn_integrate")
    return out


# --------------------------------------------------------------------------- one-pass Jacobian
def jacobian(x, centers, log_sigs, kernel, bias=None, *, gate=None, gd: GatePack = None,
             basis: int = BASIS_GAUSSIAN, want_y: bool = False, want_region_y: bool = False,
             gate_tol: Optional[float] = None):
    """dy/dx (B,O,D) of the IRBFN forward in one launch per 4 output columns (irbfn_jacobian).
    Returns jac, or (jac, y) / (jac, y, region_y) when asked.  ``gate_tol`` (>= 0, tanh gate): the gate-sparse kernel
    (irbfn_jacobian_sparse) -- (query, region) pairs with an interval factor <= gate_tol are left out."""
    lib = _lib.load()
    _chk(x, "x")
    B, D = x.shape
    _chk(centers, "centers")
    R, K, Dc = centers.shape
    if Dc != D:
        raise ValueError("centers feature count %d != x feature count %d" % (Dc, D))
    _chk(log_sigs, "log_sigs", (R, K))
    _chk(kernel, "kernel")
    O = kernel.shape[1]
    _chk(kernel, "kernel", (K, O))
    if want_y:
        if bias is None:
            raise ValueError("bias is required for y")
        _chk(bias, "bias", (O,))
    if gate is not None:
        _chk(gate, "gate", (B, R))
    jac = torch.empty((B, O, D), device=x.device, dtype=torch.float32)
    y = torch.empty((B, O), device=x.device, dtype=torch.float32) if want_y else None
    ry = torch.empty((B, R, O), device=x.device, dtype=torch.float32) if want_region_y else None
    if want_region_y and gate is None and gd is not None and gd.num_ranges < R:
        ry.zero_()          # regions without a range row have gamma == 0 and are skipped
    if B > 0 and gate_tol is not None:
        if gd is None or want_region_y:
            raise ValueError("gate_tol needs the tanh gate (gd) and no region_y")
        with torch.cuda.device(x.device):
            nbytes = lib.irbfn_jacobian_sparse_workspace_bytes(B, gd.desc.num_bounds)
            ws = _workspace(nbytes, x.device)
            rc = lib.irbfn_jacobian_sparse(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel),
                                           _ptr(bias) if want_y else None, gd.ref(), _ptr(jac), _ptr(y), B, R, K, D, O,
                                           basis, float(gate_tol), _ptr(ws), nbytes, _stream())
        _lib.check(rc, "irbfn_jacobian_sparse")
    elif B > 0:
        with torch.cuda.device(x.device):
            rc = lib.irbfn_jacobian(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias) if want_y else None,
                                    _ptr(gate), gd.ref() if gd is not None else None, _ptr(jac), _ptr(y),
                                    _ptr(ry), B, R, K, D, O, basis, _stream())
        _lib.check(rc, "irbfn_jacobian")
    out = (jac,) + ((y,) if want_y else ()) + ((ry,) if want_region_y else ())
    return out[0] if len(out) == 1 else out


# --------------------------------------------------------------------------- table-side kernels
def gather_rows(X, Y, index, x_out=None, y_out=None, check: bool = False):
    """(X[index], Y[index]) in one launch (irbfn_gather_rows); ``index`` is an int64 CUDA tensor.
    ``check=True`` synchronises and raises IndexError if any index was outside [0, N)."""
    lib = _lib.load()
    _chk(X, "X")
    N, D = X.shape
    O = 0
    if Y is not None:
        _chk(Y, "Y")
        if Y.shape[0] != N:
            raise ValueError("X and Y must have the same number of rows")
        O = Y.shape[1]
    _chk(index, "index", dtype=torch.int64)
    n = index.shape[0]
    xb = x_out if x_out is not None else torch.empty((n, D), device=X.device, dtype=torch.float32)
    yb = None
    if Y is not None:
        yb = y_out if y_out is not None else torch.empty((n, O), device=X.device, dtype=torch.float32)
        _chk(yb, "y_out", (n, O))
    _chk(xb, "x_out", (n, D))
    flag = torch.zeros(1, device=X.device, dtype=torch.int32)
    if n > 0:
        with torch.cuda.device(X.device):
            rc = lib.irbfn_gather_rows(_ptr(X), _ptr(Y), _ptr(index), _ptr(xb), _ptr(yb), n, N, D, O, _ptr(flag),
                                       _stream())
        _lib.check(rc, "irbfn_gather_rows")
        if check and int(flag.item()) != 0:
            raise IndexError("gather_rows: an index is outside [0, %d)" % N)
    return (xb, yb) if Y is not None else xb


class GridKeys:
    """Sorted per-dimension key arrays of an explicit table (ExplicitPlanner.input_keys,
    explicit_planner.py:165-175), packed flat on one device."""

    def __init__(self, keys: Sequence[Sequence[float]], device, dtype=torch.float32):
        """dtype=torch.float64 keeps the reference's float64 keys (the table's NumPy arrays): grid_index on float64
        queries is then index-exact for every value, also those within float32 rounding of a key."""
        self.device = torch.device(device)
        self.lengths = [len(k) for k in keys]
        if not keys or min(self.lengths) == 0:
            raise ValueError("every dimension needs at least one key")
        if dtype not in (torch.float32, torch.float64):
            raise ValueError("GridKeys dtype must be torch.float32 or torch.float64")
        npdt = np.float32 if dtype == torch.float32 else np.float64
        offs = np.concatenate([[0], np.cumsum(self.lengths)]).astype(np.int32)
        self.keys = torch.tensor(np.concatenate([np.asarray(k, npdt) for k in keys]), device=self.device)
        self.offset = torch.tensor(offs, device=self.device)
        self.num_dims = len(keys)


def grid_index(q, gk: GridKeys):
    """(B, num_dims) int32 grid indices: min(len_d - 1, searchsorted(keys_d, q[:, d], 'right')).  q has the dtype of
    the keys: float32 (irbfn_grid_index) or float64 (irbfn_grid_index_f64, the reference's own precision)."""
    lib = _lib.load()
    if not (isinstance(q, torch.Tensor) and q.is_cuda and q.is_contiguous() and q.dim() == 2
            and q.dtype == gk.keys.dtype):
        raise ValueError("q must be a contiguous 2-D CUDA tensor of the keys' dtype (%s)" % gk.keys.dtype)
    B, nd = q.shape
    if nd != gk.num_dims:
        raise ValueError("q has %d columns, the grid has %d dimensions" % (nd, gk.num_dims))
    out = torch.empty((B, nd), device=q.device, dtype=torch.int32)
    if B > 0:
        fn = lib.irbfn_grid_index if q.dtype == torch.float32 else lib.irbfn_grid_index_f64
        with torch.cuda.device(q.device):
            rc = fn(_ptr(q), _ptr(gk.keys), _ptr(gk.offset), _ptr(out), B, nd, _stream())
        _lib.check(rc, "irbfn_grid_index")
    return out


def nearest_row(q, table, want_distance: bool = True):
    """Brute-force nearest table row per query (irbfn_nearest_row): (index int64 (B,), distance (B,))."""
    lib = _lib.load()
    _chk(q, "q")
    _chk(table, "table")
    B, D = q.shape
    N, Dt = table.shape
    if Dt != D:
        raise ValueError("table has %d columns, queries have %d" % (Dt, D))
    idx = torch.empty(B, device=q.device, dtype=torch.int64)
    dist = torch.empty(B, device=q.device, dtype=torch.float32) if want_distance else None
    if B > 0:
        with torch.cuda.device(q.device):
            nbytes = lib.irbfn_nearest_row_workspace_bytes(B, N)
            ws = _workspace(nbytes, q.device)
            rc = lib.irbfn_nearest_row(_ptr(q), _ptr(table), _ptr(idx), _ptr(dist), B, N, D, _ptr(ws), nbytes,
                                       _stream())
        _lib.check(rc, "irbfn_nearest_row")
    return (idx, dist) if want_distance else idx
