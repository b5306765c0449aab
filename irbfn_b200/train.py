"""Data-parallel training step for WCRBFNet: forward -> loss cotangent -> hand-written VJP into
one flat gradient buffer -> all-reduce -> clip_by_global_norm -> Adam.

Mirrors the reference's step (scripts/train_nmpc.py:258-300: ``value_and_grad(loss_fn)`` then
``state.apply_gradients``; optimiser = ``optax.chain(clip_by_global_norm(g), adam(lr))`` at
:231-233).  The reference's loss also rolls the prediction through the vehicle dynamics
(dynamics.py): ``loss="frenet_oneint" | "frenet_fullint" | "st_oneint" | "st_fullint"`` selects
the reference's four loss_fn tails (train_nmpc_frenet.py:343-384, 399-413; train_nmpc.py:260-293,
306-413), evaluated together with their cotangent by ONE fused launch (irbfn_dynamics_loss)
instead of the reference's hstack + two vmapped lax.scan rollouts + their autodiff transposes.
``loss="l2" | "l1"`` is the prediction term alone.

Parameters, gradients and both Adam moments live in flat fp32 device buffers laid out
[centers | log_sigs | kernel | bias]; the leaves of the params tree are views, so the whole
update is a handful of element-wise launches over one buffer and the all-reduce is one call.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import ops, parallel
from .model import WCRBFNet, _leaf, _unwrap


class TrainState:
    """step / params / opt_state holder (the reference's flax TrainState, scripts/train_nmpc.py:236)."""

    def __init__(self, module: WCRBFNet, params: dict, lr: float, max_grad_norm: float = 1.0,
                 b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8, device=None,
                 gate_tol: Optional[float] = None, fwd_tol_abs: Optional[float] = None):
        if not isinstance(module, WCRBFNet):
            raise TypeError("TrainState drives WCRBFNet (the shape train_nmpc*.py trains)")
        p = _unwrap(params)
        device = torch.device(device) if device is not None else _leaf_device(p)
        self.module = module
        self.apply_fn = module.apply
        self.lr, self.max_grad_norm, self.b1, self.b2, self.eps = lr, max_grad_norm, b1, b2, eps
        self.step = 0
        self.step_t = None            # device copy of the Adam count (float64), created below
        R, K, D, O = module.num_regions, module.num_kernels, module.in_features, module.out_features
        rb = p.get("rbf_list", {})
        self.train_centers = "centers" in rb
        self.train_log_sigs = "log_sigs" in rb
        shapes = []
        if self.train_centers:
            shapes.append(("centers", (R, K, D)))
        if self.train_log_sigs:
            shapes.append(("log_sigs", (R, K)))
        shapes += [("kernel", (K, O)), ("bias", (O,))]
        n = sum(int(np.prod(s)) for _, s in shapes)
        self.flat = torch.empty(n, device=device, dtype=torch.float32)
        self._grad_ext = torch.zeros(n + 1, device=device, dtype=torch.float32)   # [grad | loss]: ONE collective
        self.grad = self._grad_ext[:n]
        self.mu = torch.zeros_like(self.flat)
        self.nu = torch.zeros_like(self.flat)
        self.views, self.gviews = {}, {}
        off = 0
        for name, shp in shapes:
            m = int(np.prod(shp))
            self.views[name] = self.flat[off:off + m].view(shp)
            self.gviews[name] = self.grad[off:off + m].view(shp)
            off += m
        src = {"centers": rb.get("centers"), "log_sigs": rb.get("log_sigs"),
               "kernel": p["linear"]["kernel"], "bias": p["linear"]["bias"]}
        for name in self.views:
            self.views[name].copy_(_leaf(src[name], device))
        self.mask = (ops.GRAD_CENTERS if self.train_centers else 0) | \
                    (ops.GRAD_LOG_SIGS if self.train_log_sigs else 0) | ops.GRAD_KERNEL
        self.device = device
        self.step_t = torch.zeros((), device=device, dtype=torch.float64)
        self.grad_norm = torch.zeros(1, device=device, dtype=torch.float32)
        self.dyn_params = ops.default_dyn_params()   # scripts/train_nmpc_frenet.py:322-338
        self.loss_terms = None
        # Where the backward writes this rank's LOCAL gradient (+ the loss behind it).  One process: the reduced
        # buffer itself.  Several ranks with NVLink peer memory: this rank's symmetric buffer, which the fused
        # all-reduce + update kernel of every rank reads (enable_peer_allreduce); otherwise NCCL reduces in place.
        self._grad_src = self._grad_ext
        self.peer = None
        self.peer_status = None
        self.gate_buf = None          # (B,R) gate written by the forward, read by the backward
        # None: dense parameter gradients.  >= 0: gate-sparse (irbfn_backward_sparse): the terms of (sample, region)
        # pairs whose gate is <= gate_tol are left out of the parameter gradients (2^-24 changes them by ~4e-8)
        self.gate_tol = gate_tol
        # fwd_tol_abs (>= 0): the step's FORWARD is gate-sparse too (irbfn_forward_sparse: every output within
        # fwd_tol_abs of the dense forward -- an ABSOLUTE output tolerance, because the threshold on the gate that keeps
        # the outputs accurate depends on the size of the readout weights); the backward then takes the gate description
        # instead of the gate the dense forward would have written
        self.fwd_tol_abs = fwd_tol_abs
        self.sparse_forward = fwd_tol_abs is not None and module.num_regions > 1
        self.fused_tail = n + 1 <= ops.fused_update_max_n() if device.type == "cuda" else False

    def enable_peer_allreduce(self) -> str:
        """Collective.  Switch the gradient all-reduce of this state from NCCL to the one-shot kernel over
        NVLink peer memory fused with clip + Adam (``ops.fused_update_``), when the gradient is small enough
        and peer memory can be mapped.  Returns how the gradient is reduced afterwards: 'peer:<mechanism>' or
        'nccl:<reason>'."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return "single-process"
        if self.peer is not None:
            return "peer:" + self.peer.how
        if not self.fused_tail:
            return "nccl:gradient larger than the fused kernel's limit"
        try:
            pb = parallel.PeerBuffer(self._grad_ext.numel(), self.device)
        except Exception as e:        # noqa: BLE001
            return "nccl:%s" % (str(e)[:120],)
        self.peer = pb
        self._grad_src = pb.local[: self._grad_ext.numel()]
        self.peer_status = torch.zeros(1, device=self.device, dtype=torch.int32)
        return "peer:" + pb.how

    @property
    def params(self) -> dict:
        rb = {k: self.views[k] for k in ("centers", "log_sigs") if k in self.views}
        p = {"linear": {"kernel": self.views["kernel"], "bias": self.views["bias"]}}
        if rb:
            p["rbf_list"] = rb
        return {"params": p}

    def opt_state(self) -> dict:
        return {"count": self.step, "mu": self.mu, "nu": self.nu}

    # ---- the reference's checkpoint layout -------------------------------------------------------
    def _tree_of(self, flat: torch.Tensor) -> dict:
        """A flat buffer laid out like ``self.flat`` -> {'params': {rbf_list: .., linear: ..}} of NumPy copies."""
        out, off = {}, 0
        for name, view in self.views.items():
            m = view.numel()
            out[name] = flat[off:off + m].view(view.shape).detach().cpu().numpy().copy()
            off += m
        p = {"linear": {"bias": out["bias"], "kernel": out["kernel"]}}
        rb = {k: out[k] for k in ("centers", "log_sigs") if k in out}
        if rb:
            p["rbf_list"] = rb
        return {"params": p}

    def state_dict(self) -> dict:
        """The tree ``flax.serialization.to_state_dict(TrainState)`` gives for the reference's
        ``tx = optax.chain(clip_by_global_norm(g), adam(lr))`` (scripts/train_nmpc.py:231-236): what
        ``checkpoints.save_checkpoint(CKPT_DIR, state, step)`` writes (:497-502) and
        ``restore_checkpoint(ckpt_dir, target=state)`` expects (irbfn_planner.py:88-90):

            {step, params: {params: tree},
             opt_state: {'0': {},                                   # clip_by_global_norm: EmptyState
                         '1': {'0': {count, mu: {params: tree}, nu: {params: tree}},   # scale_by_adam
                               '1': {}}}}                           # scale(-lr): EmptyState

        Pass it to ``checkpoint.save_checkpoint``."""
        return {"step": np.int64(self.step),
                "params": self._tree_of(self.flat),
                "opt_state": {"0": {}, "1": {"0": {"count": np.int32(self.step), "mu": self._tree_of(self.mu),
                                                   "nu": self._tree_of(self.nu)}, "1": {}}}}

    def load_state_dict(self, tree: dict) -> "TrainState":
        """Resume from a checkpoint tree in the reference's layout (``checkpoint.load_checkpoint``): parameters,
        both Adam moments and the step count.  Also accepts the bare-Adam layout of the oldest shipped
        checkpoints (``opt_state['0'] = {count, mu, nu}``, e.g. scripts/ckpts/nmpc_4_regions)."""
        def fill(flat, sub):
            p = sub["params"] if "params" in sub else sub
            src = {"centers": p.get("rbf_list", {}).get("centers"), "log_sigs": p.get("rbf_list", {}).get("log_sigs"),
                   "kernel": p["linear"]["kernel"], "bias": p["linear"]["bias"]}
            off = 0
            for name, view in self.views.items():
                if src[name] is None:
                    raise KeyError("checkpoint has no leaf for '%s'" % name)
                leaf = _leaf(src[name], self.device)
                if tuple(leaf.shape) != tuple(view.shape):
                    raise ValueError("leaf '%s': checkpoint %s, module %s" % (name, tuple(leaf.shape), tuple(view.shape)))
                flat[off:off + view.numel()].copy_(leaf.reshape(-1))
                off += view.numel()

        fill(self.flat, tree["params"])
        opt = tree.get("opt_state") or {}
        adam = None
        for node in (opt.get("1", {}).get("0") if isinstance(opt.get("1"), dict) else None, opt.get("0"), opt):
            if isinstance(node, dict) and "mu" in node and "nu" in node:
                adam = node
                break
        if adam is not None:
            fill(self.mu, adam["mu"])
            fill(self.nu, adam["nu"])
            self.step = int(np.asarray(adam.get("count", tree.get("step", 0))))
        else:
            self.mu.zero_(); self.nu.zero_()
            self.step = int(np.asarray(tree.get("step", 0)))
        self.step_t.fill_(float(self.step))
        return self


def _leaf_device(p):
    for v in p.values():
        if isinstance(v, dict):
            d = _leaf_device(v)
            if d is not None:
                return d
        elif isinstance(v, torch.Tensor) and v.is_cuda:
            return v.device
    return torch.device("cuda", torch.cuda.current_device())


def loss_and_cotangent(y_pred: torch.Tensor, y: torch.Tensor, global_batch: int, loss: str = "l2"):
    """(sum of this shard's loss terms / global element count, dL/dy_pred) for the mean over
    the GLOBAL batch, so that summing shard gradients over ranks gives the one-device gradient."""
    n = float(global_batch * y.shape[1])
    diff = y_pred - y
    if loss == "l2":      # optax.l2_loss = 0.5 * (pred - target)^2
        return 0.5 * (diff * diff).sum() / n, diff / n
    if loss == "l1":
        return diff.abs().sum() / n, torch.sign(diff) / n
    raise ValueError("loss must be 'l2' or 'l1'")


DYNAMICS_LOSSES = {"frenet_oneint": 0, "frenet_fullint": 1, "st_oneint": 2, "st_fullint": 3}


def clip_and_adam_(state: TrainState):
    """optax.chain(clip_by_global_norm(max_norm), adam(lr)) on the flat buffers, in place: two launches
    (irbfn_clip_adam).  The Adam count lives on the device (``state.step_t``), so the call can sit inside a
    CUDA graph; ``state.step`` is its host mirror.  Returns the pre-clip global gradient norm (device)."""
    state.step += 1
    ops.clip_adam_(state.flat, state.grad, state.mu, state.nu, state.step_t.view(1), state.lr, state.b1, state.b2,
                   state.eps, state.max_grad_norm, state.grad_norm)
    return state.grad_norm[0]


def clip_and_adam_torch_(flat, grad, mu, nu, step: int, lr, b1=0.9, b2=0.999, eps=1e-8, max_norm=1.0):
    """The same update spelled with torch element-wise ops on any device (host-side tests compare the
    kernel against this and against oracle/optim_oracle.py); ``step`` is the count AFTER the increment."""
    gnorm = torch.linalg.vector_norm(grad)
    if max_norm > 0 and not bool(gnorm < max_norm):
        grad.div_(gnorm).mul_(max_norm)
    mu.mul_(b1).add_(grad, alpha=1 - b1)
    nu.mul_(b2).addcmul_(grad, grad, value=1 - b2)
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    flat.addcdiv_(mu / bc1, (nu / bc2).sqrt_().add_(eps), value=-lr)
    return gnorm


def _local_grad(state: TrainState, x: torch.Tensor, y: torch.Tensor, gb: int, loss: str):
    """Forward + fused loss + VJP on this rank's shard: the local gradient and, behind it, the local loss term
    land in ``state._grad_src``.  The forward also writes the gate it computes, so the backward does not
    recompute it.  Returns the local loss (device scalar view)."""
    m = state.module
    dev = x.device
    centers, log_sigs = m._rbf_leaves(_unwrap(state.params), dev)
    v = state.views
    gd = m._gate_pack(dev)
    B = int(x.shape[0])
    if state.gate_buf is None or state.gate_buf.shape[0] != B:
        state.gate_buf = torch.empty((B, m.num_regions), device=dev, dtype=torch.float32)
    # the gate-sparse forward pays its binning pass and the backward's own gate pass back only on large shards
    # (27 x 500 x 8 -> 10, scripts/dev_o10_shard.py: 3.97 vs 4.26 ms at 262,144 samples, 1.46 vs 0.66 ms at 32,768)
    sparse_fwd = state.sparse_forward and (1 << 18) <= B < (1 << 31)
    if sparse_fwd:
        y_pred = ops.forward_sparse(x, centers, log_sigs, v["kernel"], v["bias"], gd=gd, basis=m._basis,
                                    tol_abs=state.fwd_tol_abs)
        gate_kw = dict(gd=gd)
    else:
        y_pred = ops.forward(x, centers, log_sigs, v["kernel"], v["bias"], gd=gd, basis=m._basis,
                             gate_out=state.gate_buf)
        gate_kw = dict(gate=state.gate_buf)
    if loss in DYNAMICS_LOSSES:
        terms, dy = ops.dynamics_loss(DYNAMICS_LOSSES[loss], x, y_pred, y, state.dyn_params, gb)
        state.loss_terms = terms          # [pred_loss, int_loss] of this shard (device)
        lval = terms.sum()
    else:
        # one launch; loss_and_cotangent is its torch spelling.  The loss lands directly behind the gradient in the
        # buffer the collective reduces: it rides along, still one collective, no concat, no copy
        n = state.grad.numel()
        lv, dy = ops.prediction_loss(y_pred, y, gb, loss, loss=state._grad_src[n:n + 1])
        ops.backward(x, centers, log_sigs, v["kernel"], dy, basis=m._basis, grad_mask=state.mask,
                     flat_out=state._grad_src[:n], gate_tol=state.gate_tol, **gate_kw)
        return lv[0]
    n = state.grad.numel()
    ops.backward(x, centers, log_sigs, v["kernel"], dy.contiguous(), basis=m._basis, grad_mask=state.mask,
                 flat_out=state._grad_src[:n], gate_tol=state.gate_tol, **gate_kw)
    state._grad_src[n:n + 1].copy_(lval.reshape(1))      # the loss rides along: still one collective, no concat
    return lval


def _world():
    import torch.distributed as dist
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def grad_step(state: TrainState, x: torch.Tensor, y: torch.Tensor, global_batch: Optional[int] = None,
              loss: str = "l2"):
    """Forward + VJP on this rank's shard, then the gradient all-reduce.  Leaves the reduced
    gradient in ``state.grad`` and returns the (global) loss as a device scalar."""
    gb = int(global_batch) if global_batch is not None else int(x.shape[0])
    lval = _local_grad(state, x, y, gb, loss)
    n = state.grad.numel()
    if state.peer is not None:
        ops.fused_update_(None, None, None, state._grad_ext, n, 1, None, 0.0, peers=state.peer.ptrs,
                          rank=state.peer.rank, world=state.peer.world, apply_update=False, status=state.peer_status)
        return state._grad_ext[-1].clone()
    if gb != x.shape[0]:
        parallel.allreduce_sum_(state._grad_ext)
        lval = state._grad_ext[-1].clone()
    return lval


def train_step(state: TrainState, x: torch.Tensor, y: torch.Tensor, global_batch: Optional[int] = None,
               loss: str = "l2"):
    """One optimiser step (scripts/train_nmpc.py:258-300).  ``x``/``y`` are this rank's shard; ``global_batch``
    the total over ranks.  Small gradients (every shipped NMPC shape) finish in ONE launch after the backward:
    one-shot all-reduce over NVLink peer memory (if enabled) + clip + Adam; larger ones take the NCCL
    all-reduce and the two-launch update.  Returns the loss (device scalar; no host sync -- the reference's
    per-step wandb device_get is the caller's business)."""
    gb = int(global_batch) if global_batch is not None else int(x.shape[0])
    if not state.fused_tail or (state.peer is None and gb != x.shape[0] and _world() > 1):
        lval = grad_step(state, x, y, global_batch, loss)
        clip_and_adam_(state)
        return lval
    _local_grad(state, x, y, gb, loss)
    n = state.grad.numel()
    state.step += 1
    pb = state.peer
    ops.fused_update_(state.flat, state.mu, state.nu, state._grad_ext, n, 1, state.step_t.view(1), state.lr, state.b1,
                      state.b2, state.eps, state.max_grad_norm, grad_norm=state.grad_norm,
                      peers=pb.ptrs if pb is not None else None, rank=pb.rank if pb is not None else 0,
                      world=pb.world if pb is not None else 1, apply_update=True, status=state.peer_status)
    return state._grad_ext[-1].clone()


class GraphedTrainStep:
    """One optimiser step captured in a CUDA graph and replayed (forward, fused loss, VJP, the NCCL
    all-reduce, clip + Adam): at 8 GPUs the step holds ~0.1 ms of kernels, less than the host needs to
    launch them one by one ("capture the step in a CUDA graph").

    ``x`` / ``y`` are static device buffers of this rank's shard: fill them (``load`` or the on-device
    gather of pipeline.DeviceTable writing straight into them) and call the object.  Capture warms the
    step up on a side stream first; parameters and optimiser state are restored afterwards, so capturing
    does not advance training."""

    def __init__(self, state: TrainState, shard_rows: int, global_batch: Optional[int] = None, loss: str = "l2",
                 warmup: int = 3):
        m = state.module
        dev = state.device
        self.state, self.loss_name = state, loss
        self.global_batch = int(global_batch) if global_batch is not None else int(shard_rows)
        self.x = torch.zeros((shard_rows, m.in_features), device=dev)
        self.y = torch.zeros((shard_rows, m.out_features), device=dev)
        snap = (state.flat.clone(), state.mu.clone(), state.nu.clone(), state.step_t.clone(), state.step)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):             # allocates workspaces, initialises NCCL channels
                train_step(state, self.x, self.y, self.global_batch, loss)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = train_step(state, self.x, self.y, self.global_batch, loss)
        torch.cuda.synchronize(dev)
        state.flat.copy_(snap[0]); state.mu.copy_(snap[1]); state.nu.copy_(snap[2]); state.step_t.copy_(snap[3])
        state.step = snap[4]

    def load(self, x: torch.Tensor, y: torch.Tensor):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        return self

    def __call__(self):
        self.graph.replay()
        self.state.step += 1
        return self.loss
