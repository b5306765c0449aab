"""Host-side mirror of the reference's Flax modules for the IRBFN hot path.

Same class names, constructor attributes, parameter tree and call shapes as
``src/irbfn_mpc/model.py`` (WCRBFNet :98-198, DeeperWCRBFNet :201-289, ClusterWCRBFNet :341-414):

    net = WCRBFNet(in_features=..., out_features=..., num_kernels=..., basis_func=gaussian,
                   num_regions=..., lower_bounds=..., upper_bounds=..., dimension_ranges=...,
                   activation_idx=..., delta=...)
    params = net.init(rng, x_example)              # {'params': {'rbf_list': {...}, 'linear': {...}}}
    y = net.apply(params, x)                        # (B, out_features)

JAX / Flax are not installable in this environment (probed on the GPU box too), so
device memory, streams and the autodiff tape come from torch instead: parameters and queries
are fp32 CUDA tensors, ``apply`` is differentiable through a ``torch.autograd.Function`` whose
backward is the hand-written VJP (the role ``jax.custom_vjp`` plays in INTEGRATION.md's XLA FFI
binding), and ``torch.autograd.grad`` stands where ``jax.grad`` does.  All arithmetic of the
path runs in libirbfn_b200.so through the C ABI; there is no CPU fallback -- CPU inputs to
``apply`` are staged to the device by a pinned, pipelined host path and the result is copied
back, nothing is computed on the host.
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence

import numpy as np
import torch

from . import ops
from .basis import basis_id, log_sig_shift

__all__ = ["WCRBFNet", "DeeperWCRBFNet", "ClusterWCRBFNet", "device_params"]


def _default_device():
    if not torch.cuda.is_available():
        raise RuntimeError("irbfn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _leaf(v, device):
    """fp32 contiguous tensor on ``device`` (cast at the call boundary; the tree is untouched)."""
    if isinstance(v, torch.Tensor):
        if v.device == device and v.dtype == torch.float32 and v.is_contiguous():
            return v
        return v.to(device=device, dtype=torch.float32).contiguous()
    return torch.as_tensor(np.ascontiguousarray(np.asarray(v), dtype=np.float32)).to(device)


def device_params(params: dict, device=None) -> dict:
    """Copy a parameter tree (NumPy / torch leaves, e.g. from checkpoint.restore_params) to the
    device once, as fp32 tensors, so that repeated ``apply`` calls do no host work."""
    device = torch.device(device) if device is not None else _default_device()

    def conv(node):
        if isinstance(node, dict):
            return {k: conv(v) for k, v in node.items()}
        return _leaf(node, device)

    return conv(params)


def _unwrap(params: dict) -> dict:
    return params["params"] if "params" in params else params


class _IrbfnFunction(torch.autograd.Function):
    """y / hidden = irbfn_forward(...); backward = irbfn_backward (the custom VJP)."""

    @staticmethod
    def forward(ctx, x, centers, log_sigs, kernel, bias, gate, gd, basis, want_hidden):
        ctx.gd, ctx.basis, ctx.want_hidden = gd, basis, want_hidden
        ctx.save_for_backward(x, centers, log_sigs, kernel, gate)
        if want_hidden:
            return ops.forward(x, centers, log_sigs, gate=gate, gd=gd, basis=basis, want_y=False,
                               want_hidden=True)
        return ops.forward(x, centers, log_sigs, kernel, bias, gate=gate, gd=gd, basis=basis)

    @staticmethod
    def backward(ctx, dout):
        x, centers, log_sigs, kernel, gate = ctx.saved_tensors
        need = ctx.needs_input_grad  # x, centers, log_sigs, kernel, bias, gate
        mask = 0
        if need[0]:
            mask |= ops.GRAD_X
        if need[1]:
            mask |= ops.GRAD_CENTERS
        if need[2]:
            mask |= ops.GRAD_LOG_SIGS
        if (need[3] or need[4]) and not ctx.want_hidden:
            mask |= ops.GRAD_KERNEL
        if gate is not None and need[5]:
            mask |= ops.GRAD_GATE
        if mask == 0:
            return (None,) * 9
        dout = dout.contiguous()
        if ctx.want_hidden:
            g = ops.backward(x, centers, log_sigs, None, None, dhidden=dout, gate=gate, gd=ctx.gd,
                             basis=ctx.basis, grad_mask=mask, gate_tol=getattr(ctx.gd, "gate_tol", None))
        else:
            g = ops.backward(x, centers, log_sigs, kernel, dout, gate=gate, gd=ctx.gd, basis=ctx.basis,
                             grad_mask=mask, gate_tol=getattr(ctx.gd, "gate_tol", None))
        return (g.get("x") if need[0] else None, g.get("centers") if need[1] else None,
                g.get("log_sigs") if need[2] else None, g.get("kernel") if need[3] else None,
                g.get("bias") if need[4] else None, g.get("gate") if need[5] else None, None, None, None)


class _SoftmaxGateFunction(torch.autograd.Function):
    """gate, logits = softmax(x Wc + bc), x Wc + bc  (model.py:402-406) via irbfn_softmax_gate."""

    @staticmethod
    def forward(ctx, x, Wc, bc):
        gate, logits = ops.softmax_gate(x, Wc, bc)
        ctx.save_for_backward(x, Wc, gate)
        return gate, logits

    @staticmethod
    def backward(ctx, dgate, dlogits_direct):
        x, Wc, gate = ctx.saved_tensors
        dlogits = gate * (dgate - (dgate * gate).sum(-1, keepdim=True))
        if dlogits_direct is not None:
            dlogits = dlogits + dlogits_direct
        return dlogits @ Wc.T, x.T @ dlogits, dlogits.sum(0)


def _lecun_normal(gen, fan_in, fan_out, device):
    # flax nn.Dense default kernel_init: variance_scaling(1.0, "fan_in", "truncated_normal")
    std = (1.0 / fan_in) ** 0.5 / 0.87962566103423978
    w = torch.empty((fan_in, fan_out), dtype=torch.float32)
    torch.nn.init.trunc_normal_(w, mean=0.0, std=1.0, a=-2.0, b=2.0, generator=gen)
    return (w * std).to(device)


class _IrbfnModule:
    """Attributes and helpers shared by the three module classes (model.py:113-125)."""

    def __init__(self, in_features: int, out_features: int, num_kernels: int, basis_func: Callable,
                 num_regions: int, lower_bounds: Sequence[Sequence[float]] = (),
                 upper_bounds: Sequence[Sequence[float]] = (),
                 dimension_ranges: Sequence[Sequence[int]] = (), activation_idx: Sequence[int] = (),
                 delta: Sequence[float] = (), centers=None, fixed_centers: Optional[bool] = False,
                 fixed_width: Optional[bool] = False, fixed_log_sigs=None):
        self.in_features = int(in_features)
        self.out_features = int(out_features)
        self.num_kernels = int(num_kernels)
        self.basis_func = basis_func
        self.num_regions = int(num_regions)
        # Flax freezes list attributes to tuples; do the same so the module is hashable/static
        self.lower_bounds = tuple(tuple(float(v) for v in b) for b in lower_bounds)
        self.upper_bounds = tuple(tuple(float(v) for v in b) for b in upper_bounds)
        self.dimension_ranges = tuple(tuple(int(v) for v in r) for r in dimension_ranges)
        self.activation_idx = tuple(int(v) for v in activation_idx)
        self.delta = tuple(float(v) for v in delta)
        self.centers = None if centers is None else np.asarray(centers, dtype=np.float32)
        self.fixed_centers = bool(fixed_centers)
        self.fixed_width = bool(fixed_width)
        # FixedWidthCenteredRBFLayer's source is absent from the reference; its
        # checkpoints hold no rbf_list leaves.  The constant width is therefore an explicit
        # attribute here (default log_sig = 0, the RBFLayer initial value) -- parity unpinned.
        self.fixed_log_sigs = fixed_log_sigs
        self.num_split_dimensions = len(self.activation_idx)   # model.py:128 (F11: only the length)
        self._basis = basis_id(basis_func)                     # raises for unsupported functions
        self._ls_shift = log_sig_shift(basis_func)             # gaussian_wide / gaussian_wider: width shift
        if self.in_features > ops._lib.MAX_D:
            raise ValueError("in_features %d exceeds the compiled maximum %d" % (in_features, ops._lib.MAX_D))
        if self.centers is not None:
            c = self.centers
            if c.ndim == 2:
                c = np.broadcast_to(c[None], (self.num_regions,) + c.shape)
            if c.shape != (self.num_regions, self.num_kernels, self.in_features):
                raise ValueError("centers must be (K,D) or (R,K,D); got %s" % (self.centers.shape,))
            self.centers = np.ascontiguousarray(c)
        self._packs = {}
        self._const = {}

    # ---- static gate attributes, packed once per device (constraint 2) ----
    def _gate_pack(self, device):
        key = str(device)
        if key not in self._packs:
            S = self.num_split_dimensions
            if S == 0:
                raise ValueError("activation_idx is empty: the tanh gate needs at least one split dimension")
            self._packs[key] = ops.GatePack(self.lower_bounds[:S], self.upper_bounds[:S], self.delta[:S],
                                            [r[:S] for r in self.dimension_ranges], device)
        # None: dense VJP.  >= 0: autograd through apply() uses the gate-sparse VJP (irbfn_backward_sparse): terms of
        # (query, region) pairs whose gate is <= gate_tol are left out (set ``module.gate_tol = 2.0 ** -24``)
        self._packs[key].gate_tol = getattr(self, "gate_tol", None)
        return self._packs[key]

    def _fixed(self, name, device):
        key = (name, str(device))
        if key not in self._const:
            if name == "centers":
                v = self.centers
            else:
                v = np.zeros((self.num_regions, self.num_kernels), np.float32) if self.fixed_log_sigs is None \
                    else np.broadcast_to(np.asarray(self.fixed_log_sigs, np.float32),
                                         (self.num_regions, self.num_kernels))
            self._const[key] = _leaf(v, device)
        return self._const[key]

    def _rbf_leaves(self, p, device):
        """(centers, log_sigs) from the tree or, for the fixed-centre / fixed-width layers, from
        the module attributes (their checkpoints lack those leaves)."""
        rb = p.get("rbf_list", {})
        centers = _leaf(rb["centers"], device) if "centers" in rb else None
        log_sigs = _leaf(rb["log_sigs"], device) if "log_sigs" in rb else None
        if centers is None:
            if self.centers is None:
                raise KeyError("params have no rbf_list/centers and the module has no `centers` attribute")
            centers = self._fixed("centers", device)
        if log_sigs is None:
            if not self.fixed_width:
                raise KeyError("params have no rbf_list/log_sigs and the module is not fixed_width")
            log_sigs = self._fixed("log_sigs", device)
        if self._ls_shift != 0.0:
            # gaussian_wide / _wider on the gaussian kernel (basis.py).  The shifted widths are kept until the source
            # tensor is replaced or written to (its version counter), so repeated calls on a device-resident tree
            # launch no extra kernel and hand the same buffer to captured graphs (hostpath: graph keys are addresses).
            if log_sigs.requires_grad:
                return centers, log_sigs + self._ls_shift        # autograd needs the fresh node
            key = ("ls_shift", str(device))
            hit = self._const.get(key)
            # the entry holds the source tensor itself (identity, not address: a freed tensor's address is reused)
            if hit is None or hit[0] is not log_sigs or hit[1] != log_sigs._version:
                self._const[key] = hit = (log_sigs, log_sigs._version, log_sigs + self._ls_shift)
            log_sigs = hit[2]
        return centers, log_sigs

    # ---- init: same tree, shapes and distributions as the reference -------------------------
    def _init_rbf(self, gen, device):
        R, K, D = self.num_regions, self.num_kernels, self.in_features
        rb = {}
        if self.centers is None:
            rb["centers"] = torch.randn((R, K, D), generator=gen).to(device)   # normal(1.0), flax_rbf.py:247-251
        elif not self.fixed_centers:
            rb["centers"] = _leaf(self.centers, device).clone()                # warm start, trainable
        if not (self.centers is not None and self.fixed_centers and self.fixed_width):
            rb["log_sigs"] = torch.zeros((R, K), device=device)                # constant 0, flax_rbf.py:252-256
        return rb

    @staticmethod
    def _generator(rng):
        if isinstance(rng, torch.Generator):
            return rng
        seed = int(np.asarray(rng).ravel()[-1]) if not isinstance(rng, (int, np.integer)) else int(rng)
        return torch.Generator().manual_seed(seed)

    def _device_of(self, x):
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.device
        return _default_device()

    def _prep_x(self, x, device):
        if isinstance(x, torch.Tensor) and x.is_cuda:
            if x.dim() != 2 or x.shape[1] != self.in_features:
                raise ValueError("x must be (B, %d); got %s" % (self.in_features, tuple(x.shape)))
            return x if (x.dtype == torch.float32 and x.is_contiguous()) else x.float().contiguous()
        return None

    def __call__(self, params, x):
        return self.apply(params, x)


class WCRBFNet(_IrbfnModule):
    """Interpolating RBF network: tanh region gate -> per-region RBF layers -> gated blend ->
    one shared Dense readout (model.py:98-198)."""

    def init(self, rng, x, device=None):
        device = torch.device(device) if device is not None else self._device_of(x)
        gen = self._generator(rng)
        p = {}
        rb = self._init_rbf(gen, device)
        if rb:
            p["rbf_list"] = rb
        p["linear"] = {"kernel": _lecun_normal(gen, self.num_kernels, self.out_features, device),
                       "bias": torch.zeros(self.out_features, device=device)}
        return {"params": p}

    def apply(self, params, x):
        device = self._device_of(x)
        p = _unwrap(params)
        xd = self._prep_x(x, device)
        if xd is None:
            return self.apply_host(params, x)
        centers, log_sigs = self._rbf_leaves(p, device)
        kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
        return _IrbfnFunction.apply(xd, centers, log_sigs, kernel, bias, None, self._gate_pack(device),
                                    self._basis, False)

    def apply_lattice(self, params, axes, fixed, tol_abs: Optional[float] = None, out=None):
        """The sweep of irbfn_vs_explicit*.ipynb cells 8-10 without moving the lattice over the host link:
        queries = meshgrid(axes, indexing="ij") with the other columns held at ``fixed``, generated on the device
        (``ops.lattice_queries``), then ``apply`` (or ``apply_sparse`` when ``tol_abs`` is given).
        Returns (x, y), both CUDA tensors -- or, with ``out`` (a pinned CPU tensor (B,O), O <= 16, dense evaluation),
        y is stored by the kernel straight into that host buffer while it computes (no download pass; valid once the
        current stream has been synchronised) and (x, out) is returned."""
        device = _default_device()
        x = ops.lattice_queries(axes, fixed, device)
        if out is not None:
            if tol_abs is not None or self.out_features > 16:
                raise ValueError("out= needs the dense evaluation and at most 16 outputs")
            p = _unwrap(params)
            centers, log_sigs = self._rbf_leaves(p, device)
            kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
            key = ("ready1", str(device))
            if key not in self._const:
                self._const[key] = torch.zeros(2, dtype=torch.int32, device=device)
            ready = self._const[key]          # [flag of the one piece = the whole batch, tile counter]
            ready[0:1].fill_(1)
            ready[1:2].zero_()
            qpf = ((x.shape[0] + 1023) // 1024) * 1024
            ops.forward_streamed(x, centers, log_sigs, kernel, bias, out, ready, qpf, gd=self._gate_pack(device),
                                 basis=self._basis)
            return x, out
        y = self.apply(params, x) if tol_abs is None else self.apply_sparse(params, x, tol_abs)
        return x, y

    def apply_sparse(self, params, x, tol_abs: float, counters=None):
        """``apply`` for inference sweeps with gate-sparse evaluation (irbfn_forward_sparse): local networks whose
        gate cannot move any output by more than ``tol_abs`` in total are skipped, queries are binned so that
        warps skip uniformly.  |result - apply(params, x)| <= tol_abs per output.  Not differentiable."""
        device = self._device_of(x)
        p = _unwrap(params)
        xd = self._prep_x(x, device)
        if xd is None:
            xd = _leaf(x, device)
        centers, log_sigs = self._rbf_leaves(p, device)
        kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
        return ops.forward_sparse(xd, centers, log_sigs, kernel, bias, gd=self._gate_pack(device), basis=self._basis,
                                  tol_abs=tol_abs, counters=counters)

    def jacobian(self, params, x, with_value: bool = False):
        """jax.vmap(jax.jacrev(lambda q: self.apply(params, q[None])[0]))(x): (B, O, D), one pass
        (SURVEY.md 8f rank 2).  ``with_value`` also returns y (B, O) from the same launch."""
        device = self._device_of(x)
        p = _unwrap(params)
        xd = self._prep_x(x, device)
        if xd is None:
            xd = _leaf(x, device)
        centers, log_sigs = self._rbf_leaves(p, device)
        kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
        return ops.jacobian(xd, centers, log_sigs, kernel, bias, gd=self._gate_pack(device), basis=self._basis,
                            want_y=with_value)

    def apply_host(self, params, x, out=None, chunk: Optional[int] = None, copy_streams: int = 1,
                   graph: bool = False, run_streams: int = 1, streamed: Optional[bool] = None):
        """Host-buffer entry point: x is a NumPy array or CPU tensor (B,D).  Returns a NumPy array (or fills ``out``, a
        pinned CPU tensor, and returns it).
        Default: the batch is cut into chunks whose host->device copy, kernel and device->host copy overlap on three
        streams (hostpath.HostPipeline).  ``streamed=True`` (outputs of at most 16 columns): ONE persistent launch whose
        query tiles wait for their piece of the upload and which stores y straight into the pinned host buffer
        (hostpath.StreamedPipeline) -- 8 % faster than the chunk pipeline on some boxes of this pool, 10 % slower on
        others (the per-copy set-up cost of the host link decides), so it is opt-in; ``streamed="auto"`` times both
        once per (device, batch size) on the caller's buffers (``out`` must be given) and keeps the faster."""
        from .hostpath import HostPipeline, PredStepPipeline, StreamedPipeline
        device = _default_device()
        p = _unwrap(params)
        centers, log_sigs = self._rbf_leaves(p, device)
        kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
        gd = self._gate_pack(device)
        rows = int(x.shape[0]) if getattr(x, "ndim", 0) == 2 else 0
        if (out is None and not streamed and not graph and chunk is None and copy_streams == 1 and run_streams == 1
                and 0 < rows <= PredStepPipeline.MAX_ROWS and not (isinstance(x, torch.Tensor) and x.is_pinned())):
            # the planner's call (pred_step): a few queries, the caller waits -- one captured graph per batch size
            key = ("pred", str(device))
            if key not in self._const:
                self._const[key] = PredStepPipeline(device, self.in_features, self.out_features)
            return self._const[key].run(x, lambda xs, ys: ops.forward(xs, centers, log_sigs, kernel, bias, gd=gd,
                                                                      basis=self._basis, out=ys),
                                        (centers.data_ptr(), log_sigs.data_ptr(), kernel.data_ptr(), bias.data_ptr()))
        if streamed and (self.out_features > 16 or not ops.upload_streamed_available()):
            streamed = False
        if streamed == "auto":
            # which pipeline wins depends on the box's host link (per-copy set-up cost): time both ONCE per (device,
            # batch size) on the caller's own buffers and keep the faster.  Small batches take the chunk pipeline.
            B = int(x.shape[0])
            akey = ("auto", str(device), B)
            if B < (1 << 18) or out is None:
                streamed = False
            elif akey in self._const:
                streamed = self._const[akey]
            else:
                def _time(flag):
                    for _ in range(3):
                        self.apply_host(params, x, out=out, chunk=chunk, streamed=flag)
                    torch.cuda.synchronize(device)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(10):
                        self.apply_host(params, x, out=out, chunk=chunk, streamed=flag)
                    e1.record()
                    torch.cuda.synchronize(device)
                    return e0.elapsed_time(e1)
                # first use of either pipeline: allocations, lazy loading.  A streamed call whose upload never arrived
                # (status word: a profiler serialising the kernel against the copy streams) rules that path out.
                self.apply_host(params, x, out=out, chunk=chunk, streamed=True)
                try:
                    self._const[("spipe", str(device), chunk)].check()
                    usable = True
                except RuntimeError:
                    usable = False
                if not usable:
                    streamed = self._const[akey] = False
                else:
                    _time(True), _time(False)
                    t_s, t_c = min(_time(True), _time(True)), min(_time(False), _time(False))
                    streamed = self._const[akey] = bool(t_s < 0.98 * t_c)
        if streamed:
            key = ("spipe", str(device), chunk)
            if key not in self._const:
                self._const[key] = StreamedPipeline(device, self.in_features, self.out_features,
                                                    chunk if chunk is not None else 1 << 15, 2)
            return self._const[key].run(x, out, lambda xd, yh, ready, qpf, status: ops.forward_streamed(
                xd, centers, log_sigs, kernel, bias, yh, ready, qpf, gd=gd, basis=self._basis, status=status))
        if chunk is None:
            # one full wave of the persistent forward per chunk: (SM count) query tiles of 1,024, so no
            # chunk leaves SMs idle (2^17 queries = 128 tiles on 148 SMs cost 13 % of the kernel time)
            chunk = torch.cuda.get_device_properties(device).multi_processor_count * 1024
        key = ("pipe", str(device), chunk, copy_streams, run_streams)
        if key not in self._const:
            self._const[key] = HostPipeline(device, self.in_features, self.out_features, chunk, copy_streams,
                                            run_streams)
        pipe = self._const[key]
        return pipe.run(x, out, lambda xs, ys: ops.forward(xs, centers, log_sigs, kernel, bias, gd=gd,
                                                           basis=self._basis, out=ys, split=run_streams == 1),
                        graph=graph,
                        graph_key=(centers.data_ptr(), log_sigs.data_ptr(), kernel.data_ptr(), bias.data_ptr()))


class DeeperWCRBFNet(_IrbfnModule):
    """Same front end, then Dense(64) relu Dense(64) relu Dense(O) (model.py:201-289).  The
    front end (gate, RBF layers, blend) is the fused kernel; the 64-wide tail is three library
    GEMMs, as it stays in XLA on the reference side of the custom call."""

    hidden_features = 64

    def init(self, rng, x, device=None):
        device = torch.device(device) if device is not None else self._device_of(x)
        gen = self._generator(rng)
        H = self.hidden_features
        p = {}
        rb = self._init_rbf(gen, device)
        if rb:
            p["rbf_list"] = rb
        p["linear_pre1"] = {"kernel": _lecun_normal(gen, self.num_kernels, H, device), "bias": torch.zeros(H, device=device)}
        p["linear_pre2"] = {"kernel": _lecun_normal(gen, H, H, device), "bias": torch.zeros(H, device=device)}
        p["linear"] = {"kernel": _lecun_normal(gen, H, self.out_features, device),
                       "bias": torch.zeros(self.out_features, device=device)}
        return {"params": p}

    def apply(self, params, x):
        device = self._device_of(x)
        p = _unwrap(params)
        xd = self._prep_x(x, device)
        host = xd is None
        if host:
            xd = _leaf(x, device)
        centers, log_sigs = self._rbf_leaves(p, device)
        h = _IrbfnFunction.apply(xd, centers, log_sigs, None, None, None, self._gate_pack(device), self._basis, True)
        lin = lambda name, t: t @ _leaf(p[name]["kernel"], device) + _leaf(p[name]["bias"], device)
        h = torch.relu(lin("linear_pre1", h))      # model.py:285
        h = torch.relu(lin("linear_pre2", h))      # model.py:286
        y = lin("linear", h)                       # model.py:287
        return y.cpu().numpy() if host else y


class ClusterWCRBFNet(_IrbfnModule):
    """Gate = softmax(Dense(D -> R)(x)) instead of the tanh regions; returns (out, logits)
    (model.py:341-414)."""

    def init(self, rng, x, device=None):
        device = torch.device(device) if device is not None else self._device_of(x)
        gen = self._generator(rng)
        p = {}
        rb = self._init_rbf(gen, device)
        if rb:
            p["rbf_list"] = rb
        p["cluster"] = {"kernel": _lecun_normal(gen, self.in_features, self.num_regions, device),
                        "bias": torch.zeros(self.num_regions, device=device)}
        p["linear"] = {"kernel": _lecun_normal(gen, self.num_kernels, self.out_features, device),
                       "bias": torch.zeros(self.out_features, device=device)}
        return {"params": p}

    def apply(self, params, x):
        device = self._device_of(x)
        p = _unwrap(params)
        xd = self._prep_x(x, device)
        host = xd is None
        if host:
            xd = _leaf(x, device)
        centers, log_sigs = self._rbf_leaves(p, device)
        kernel, bias = _leaf(p["linear"]["kernel"], device), _leaf(p["linear"]["bias"], device)
        gate, logits = _SoftmaxGateFunction.apply(xd, _leaf(p["cluster"]["kernel"], device),
                                                  _leaf(p["cluster"]["bias"], device))
        y = _IrbfnFunction.apply(xd, centers, log_sigs, kernel, bias, gate, None, self._basis, False)
        if host:
            return y.cpu().numpy(), logits.cpu().numpy()
        return y, logits
