"""TEST INFRASTRUCTURE -- torch-CPU stand-ins for the slice of jax / jax.numpy / jax.lax / chex / optax
that the reference's hot-path source touches (see oracle/refshim/__init__.py).

Library semantics reproduced here (JAX's documented behaviour, not reference code):
  * arrays are immutable; ``a.at[idx].set(v)`` returns an updated copy;
  * an out-of-range STATIC integer index clamps (``x_u[9]`` on a 9-vector reads ``x_u[8]`` -- JAX's gather
    clamps; dynamics.py:327 hits this for 2-output networks);
  * ``reshape(..., order="F")``;
  * ``jnp.array([...])`` of scalars / 0-d arrays stacks them;
  * ``jax.vmap(f, in_axes)``: per-example call over the mapped axis, outputs stacked on axis 0;
  * ``jax.lax.scan(f, init, xs)``: ``carry, y = f(carry, xs[i])``, ys stacked;
  * ``jax.lax.select(pred, a, b)`` = where;
  * ``jax.jit`` / ``chex.assert_max_traces``: identity decorators;
  * ``jax.value_and_grad(f, has_aux)``: torch.autograd.grad over the leaves of the first argument;
  * ``optax.l2_loss(p, t) = 0.5 (p - t)^2``; ``optax.softmax_cross_entropy(logits, labels) =
    -sum(labels * log_softmax(logits), -1)`` (optax's published definitions).
"""
from __future__ import annotations

import types

import numpy as np
import torch

_DTYPE = torch.float64


def set_dtype(name: str) -> None:
    """'float64' = jax_enable_x64 (training with --use_float64); 'float32' = JAX's default."""
    global _DTYPE
    _DTYPE = {"float64": torch.float64, "float32": torch.float32}[name]


def dtype():
    return _DTYPE


class _AtIndex:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def _apply(self, fn):
        out = self.arr.clone()
        out[self.idx] = fn(out[self.idx])
        return out

    def set(self, v):
        return self._apply(lambda cur: _as_plain(v) if isinstance(v, torch.Tensor) else v)

    def add(self, v):
        return self._apply(lambda cur: cur + v)

    def multiply(self, v):
        return self._apply(lambda cur: cur * v)


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIndex(self.arr, idx)


def _as_plain(t):
    return t.as_subclass(torch.Tensor) if isinstance(t, Array) else t


class Array(torch.Tensor):
    """A torch tensor with the few jax.Array behaviours the reference relies on."""

    @property
    def at(self):
        return _At(self)

    def __getitem__(self, idx):
        # JAX clamps out-of-range static integer indices on reads
        if isinstance(idx, int) and self.dim() >= 1:
            n = self.shape[0]
            if idx >= n:
                idx = n - 1
            elif idx < -n:
                idx = 0
        return super().__getitem__(idx)

    def reshape(self, *shape, order="C"):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        if order == "C":
            return super().reshape(*shape)
        if order != "F":
            raise ValueError(order)
        rev = tuple(reversed(range(self.dim())))
        t = self.permute(*rev) if self.dim() > 1 else self
        t = torch.Tensor.reshape(t.contiguous(), *tuple(reversed(shape)))
        return t.permute(*tuple(reversed(range(len(shape))))) if len(shape) > 1 else t

    def astype(self, dt):
        return self.to(_to_torch_dtype(dt))

    def block_until_ready(self):
        return self


def _to_torch_dtype(dt):
    if isinstance(dt, torch.dtype):
        return dt
    return {"float32": torch.float32, "float64": torch.float64, "int32": torch.int32,
            "int64": torch.int64}[np.dtype(dt).name]


def asarray(v, dt=None):
    """-> Array.  Tensors keep their identity (and autograd history); numbers / lists / ndarrays are
    converted to the active float dtype (integers stay integers)."""
    if isinstance(v, Array):
        return v if dt is None else v.to(dt)
    if isinstance(v, torch.Tensor):
        return v.as_subclass(Array) if dt is None else v.to(dt).as_subclass(Array)
    if isinstance(v, (list, tuple)) and any(isinstance(e, torch.Tensor) for e in _flat(v)):
        return torch.stack([asarray(e, dt or _DTYPE) for e in v]).as_subclass(Array)
    a = np.asarray(v)
    if dt is None:
        dt = _DTYPE if a.dtype.kind == "f" else _to_torch_dtype(a.dtype) if a.dtype.kind in "iu" else None
        if a.dtype.kind == "b":
            dt = torch.bool
    return torch.as_tensor(a, dtype=dt).as_subclass(Array)


def _flat(v):
    for e in v:
        if isinstance(e, (list, tuple)):
            yield from _flat(e)
        else:
            yield e


# ----------------------------------------------------------------------------------------- pytrees
def tree_map(fn, tree, *rest):
    if isinstance(tree, dict):
        return {k: tree_map(fn, tree[k], *[r[k] for r in rest]) for k in tree}
    if isinstance(tree, (list, tuple)):
        return type(tree)(tree_map(fn, t, *[r[i] for r in rest]) for i, t in enumerate(tree))
    return fn(tree, *rest)


def tree_leaves(tree):
    out = []
    tree_map(out.append, tree)
    return out


def _stack_tree(items, axis=0):
    first = items[0]
    if isinstance(first, dict):
        return {k: _stack_tree([it[k] for it in items], axis) for k in first}
    if isinstance(first, (list, tuple)):
        return type(first)(_stack_tree([it[i] for it in items], axis) for i in range(len(first)))
    return torch.stack([asarray(it) for it in items], dim=axis).as_subclass(Array)


# ----------------------------------------------------------------------------------------- jax.*
def jit(fn=None, **kw):
    if fn is None:
        return lambda f: f
    return fn


def vmap(fn, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = None
        for a, ax in zip(args, axes):
            if ax is not None:
                n = asarray(a).shape[ax]
                break
        outs = []
        for i in range(n):
            call = [a if ax is None else asarray(a).select(ax, i).as_subclass(Array) for a, ax in zip(args, axes)]
            outs.append(fn(*call))
        return _stack_tree(outs, out_axes)
    return mapped


def scan(f, init, xs, length=None):
    carry = init
    ys = []
    n = asarray(xs).shape[0] if xs is not None else length
    for i in range(n):
        carry, y = f(carry, asarray(xs)[i] if xs is not None else None)
        ys.append(y)
    return carry, _stack_tree(ys, 0)


def select(pred, on_true, on_false):
    return torch.where(asarray(pred), asarray(on_true), asarray(on_false)).as_subclass(Array)


def value_and_grad(fn, argnums=0, has_aux=False):
    def run(*args):
        args = list(args)
        tracked = tree_map(lambda l: asarray(l).detach().clone().requires_grad_(True).as_subclass(Array),
                           args[argnums])
        args[argnums] = tracked
        out = fn(*args)
        value, aux = (out if has_aux else (out, None))
        leaves = tree_leaves(tracked)
        grads = torch.autograd.grad(value, leaves, allow_unused=True)
        it = iter(grads)
        gtree = tree_map(lambda l: (lambda g: torch.zeros_like(l) if g is None else g)(next(it)).as_subclass(Array),
                         tracked)
        detach = lambda t: t.detach() if isinstance(t, torch.Tensor) else t
        value = detach(value)
        if has_aux:
            return (value, tree_map(detach, aux)), gtree
        return value, gtree
    return run


def grad(fn, argnums=0, has_aux=False):
    vg = value_and_grad(fn, argnums, has_aux)

    def run(*args):
        v, g = vg(*args)
        return (g, v[1]) if has_aux else g
    return run


# ----------------------------------------------------------------------------------------- jax.numpy
def _wrap(fn):
    def call(*a, **k):
        a = [asarray(v) if isinstance(v, (torch.Tensor, np.ndarray, list, tuple)) else v for v in a]
        r = fn(*a, **k)
        return r.as_subclass(Array) if isinstance(r, torch.Tensor) else r
    return call


def _clip(x, a_min=None, a_max=None, min=None, max=None):
    lo = a_min if a_min is not None else min
    hi = a_max if a_max is not None else max
    return torch.clamp(asarray(x), min=lo, max=hi).as_subclass(Array)


def _sum(a, axis=None, keepdims=False):
    a = asarray(a)
    return (a.sum() if axis is None else a.sum(dim=axis, keepdim=keepdims)).as_subclass(Array)


def _mean(a, axis=None):
    a = asarray(a)
    return (a.mean() if axis is None else a.mean(dim=axis)).as_subclass(Array)


def _repeat(a, repeats, axis=None):
    return torch.repeat_interleave(asarray(a), repeats, dim=axis).as_subclass(Array)


def _shape_arg(shape):
    return (shape,) if isinstance(shape, int) else tuple(shape)


def _cat(parts, dim):
    parts = [asarray(p) for p in parts]
    return torch.cat(parts, dim=dim).as_subclass(Array)


def _hstack(parts):
    parts = [asarray(p) for p in parts]
    return _cat(parts, 0 if parts[0].dim() == 1 else 1)


def _vstack(parts):
    parts = [asarray(p) for p in parts]
    return _cat([p[None] if p.dim() == 1 else p for p in parts], 0)


def _atan2(a, b):
    a, b = asarray(a), asarray(b)
    a, b = torch.broadcast_tensors(a, b)
    return torch.atan2(a, b).as_subclass(Array)


def build_modules() -> dict:
    jnp = types.ModuleType("jax.numpy")
    jnp.array = lambda v, dtype=None: asarray(v, _to_torch_dtype(dtype) if dtype is not None else None)
    jnp.asarray = jnp.array
    jnp.zeros = lambda shape, dtype=None: torch.zeros(_shape_arg(shape), dtype=_DTYPE).as_subclass(Array)
    jnp.ones = lambda shape, dtype=None: torch.ones(_shape_arg(shape), dtype=_DTYPE).as_subclass(Array)
    jnp.ones_like = _wrap(torch.ones_like)
    jnp.zeros_like = _wrap(torch.zeros_like)
    jnp.broadcast_to = lambda a, shape: torch.broadcast_to(asarray(a), tuple(shape)).as_subclass(Array)
    jnp.expand_dims = lambda a, axis: asarray(a).unsqueeze(axis).as_subclass(Array)
    jnp.repeat = _repeat
    jnp.sum = _sum
    jnp.mean = _mean
    jnp.clip = _clip
    jnp.hstack = _hstack
    jnp.vstack = _vstack
    jnp.stack = lambda parts, axis=0: torch.stack([asarray(p) for p in parts], dim=axis).as_subclass(Array)
    jnp.concatenate = lambda parts, axis=0: _cat(parts, axis)
    jnp.shape = lambda a: tuple(asarray(a).shape)
    jnp.atan2 = jnp.arctan2 = _atan2
    jnp.atan = jnp.arctan = _wrap(torch.atan)
    for name in ("tanh", "exp", "log", "sin", "cos", "tan", "abs", "sqrt", "square", "isnan", "any", "all"):
        setattr(jnp, name, _wrap(getattr(torch, name)))
    jnp.where = lambda c, a, b: select(c, a, b)
    jnp.float32, jnp.float64, jnp.int32 = np.float32, np.float64, np.int32
    jnp.ndarray = jnp.DeviceArray = Array
    linalg = types.ModuleType("jax.numpy.linalg")
    linalg.norm = lambda a, axis=None: (torch.linalg.vector_norm(asarray(a)) if axis is None
                                        else torch.linalg.vector_norm(asarray(a), dim=axis)).as_subclass(Array)
    jnp.linalg = linalg

    lax = types.ModuleType("jax.lax")
    lax.scan, lax.select = scan, select

    tree_util = types.ModuleType("jax.tree_util")
    tree_util.tree_map, tree_util.tree_leaves = tree_map, tree_leaves

    random = types.ModuleType("jax.random")
    random.PRNGKey = lambda seed: int(seed)
    random.split = lambda key, num=2: tuple(int(key) * 1000003 + i + 1 for i in range(num))

    jax = types.ModuleType("jax")
    jax.numpy, jax.lax, jax.tree_util, jax.random = jnp, lax, tree_util, random
    jax.jit, jax.vmap, jax.value_and_grad, jax.grad = jit, vmap, value_and_grad, grad
    jax.Array = Array
    jax.tree = types.SimpleNamespace(map=tree_map, leaves=tree_leaves)
    jax.__path__ = []

    chex = types.ModuleType("chex")
    chex.assert_max_traces = lambda *a, **k: (lambda f: f)

    optax = types.ModuleType("optax")
    optax.l2_loss = lambda predictions, targets=None: (
        0.5 * (asarray(predictions) - (0 if targets is None else asarray(targets))) ** 2).as_subclass(Array)
    optax.softmax_cross_entropy = lambda logits, labels: (
        -(asarray(labels) * torch.log_softmax(asarray(logits), dim=-1)).sum(-1)).as_subclass(Array)

    return {"jax": jax, "jax.numpy": jnp, "jax.lax": lax, "jax.tree_util": tree_util, "jax.random": random,
            "chex": chex, "optax": optax}
