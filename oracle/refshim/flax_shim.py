"""TEST INFRASTRUCTURE -- a minimal stand-in for flax.linen, enough to run the reference's modules
(src/irbfn_mpc/model.py, the flax_rbf snapshot) unmodified.  See oracle/refshim/__init__.py.

flax.linen semantics reproduced (library behaviour, not reference code):
  * a Module subclass is a dataclass of its annotated fields;
  * ``setup()`` runs when the module is bound; a Module assigned to an attribute there becomes a
    sub-module whose parameters live under that attribute's NAME in the parent's tree
    (``self.linear = nn.Dense(..)`` -> ``params['linear']``);
  * ``self.param(name, init_fn, shape)`` declares a leaf: created by ``init_fn(key, shape)`` under
    ``init``, read from the variables under ``apply``;
  * ``nn.vmap(target, in_axes, out_axes, axis_size, variable_axes={'params': 0 | None}, split_rngs)``:
    a lifted per-example loop over the mapped axis whose parameters are stacked on axis 0
    (``variable_axes 0``) or shared (``None``); the lifted module is transparent in the tree, which is
    why shipped checkpoints hold ``rbf_list/centers`` as (R, K, D) (model.py:143-166);
  * ``nn.Dense(features)``: ``x @ kernel + bias``, kernel (in, features) lecun-normal, bias zeros;
  * ``module.init(rng, x)`` -> ``{'params': tree}``; ``module.apply(variables, x)``.

Constructor keywords the snapshot's RBFLayer predates: model.py:161-166 passes ``centers_values=`` to
whichever layer it picked; current flax_rbf gives RBFLayer that field, the 2022 snapshot does not.  A
keyword that is not a field is accepted only when its value is None (i.e. unused), else TypeError.
"""
from __future__ import annotations

import copy
import dataclasses
import types

import numpy as np
import torch

from . import jax_shim as J


class _Scope:
    """A node of the variables tree.  ``rng`` is not None under ``init``."""

    def __init__(self, tree: dict, rng=None):
        self.tree, self.rng = tree, rng

    def child(self, name: str) -> "_Scope":
        if self.rng is None and name not in self.tree:
            raise KeyError("no parameters named '%s' (have %s)" % (name, sorted(self.tree)))
        sub = self.tree.setdefault(name, {})
        rng = None if self.rng is None else np.random.default_rng(self.rng.integers(2 ** 63))
        return _Scope(sub, rng)


class Module:
    name = None
    _scope = None

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        dataclasses.dataclass(cls, eq=False, repr=False)
        generated = cls.__init__
        names = {f.name for f in dataclasses.fields(cls)}

        def __init__(self, *a, **k):
            extra = {n: v for n, v in k.items() if n not in names}
            for n, v in extra.items():
                if v is not None:
                    raise TypeError("%s has no field '%s'" % (cls.__name__, n))
            generated(self, *a, **{n: v for n, v in k.items() if n in names})

        cls.__init__ = __init__

    # --- binding -------------------------------------------------------------------------------
    def _bind(self, scope: _Scope):
        m = copy.copy(self)
        object.__setattr__(m, "_scope", scope)
        object.__setattr__(m, "_in_setup", True)
        if hasattr(m, "setup"):
            m.setup()
        object.__setattr__(m, "_in_setup", False)
        return m

    def __setattr__(self, key, value):
        if getattr(self, "_in_setup", False) and isinstance(value, (Module, _Lifted)):
            value = value._bind(self._scope.child(key))
        object.__setattr__(self, key, value)

    def param(self, name, init_fn, *init_args):
        tree = self._scope.tree
        if name not in tree:
            if self._scope.rng is None:
                raise KeyError("parameter '%s' missing from the variables (have %s)" % (name, sorted(tree)))
            tree[name] = J.asarray(init_fn(self._scope.rng, *init_args))
        leaf = J.asarray(tree[name])
        if init_args and tuple(leaf.shape) != tuple(init_args[0]):
            raise ValueError("parameter '%s' has shape %s, module expects %s"
                             % (name, tuple(leaf.shape), tuple(init_args[0])))
        return leaf

    # --- public API ----------------------------------------------------------------------------
    def apply(self, variables, *args, **kw):
        tree = J.tree_map(J.asarray, variables["params"])
        return self._bind(_Scope(tree))(*args, **kw)

    def init(self, rng, *args, **kw):
        seed = rng if isinstance(rng, (int, np.integer)) else 0
        tree = {}
        self._bind(_Scope(tree, np.random.default_rng(seed)))(*args, **kw)
        return {"params": J.tree_map(lambda t: t.detach(), tree)}


class _Lifted:
    """nn.vmap(target, ...)(**kwargs): a per-example loop; parameters stacked (axis 0) or shared (None)."""

    def __init__(self, target, kwargs, in_axes, out_axes, axis_size, param_axis):
        self.target, self.kwargs = target, kwargs
        self.in_axes, self.out_axes, self.axis_size, self.param_axis = in_axes, out_axes, axis_size, param_axis
        self._scope = None

    def _bind(self, scope):
        m = copy.copy(self)
        m._scope = scope
        return m

    def __call__(self, x):
        x = J.asarray(x)
        scope = self._scope
        n = self.axis_size if self.in_axes is None else x.shape[self.in_axes]
        inner = self.target(**self.kwargs)
        outs = []
        if self.param_axis is None:                       # shared parameters (the batch lift)
            for i in range(n):
                xi = x if self.in_axes is None else x.select(self.in_axes, i).as_subclass(J.Array)
                outs.append(inner._bind(scope)(xi))
        else:                                             # parameters stacked on axis 0 (the region lift)
            if self.param_axis != 0:
                raise NotImplementedError("variable_axes other than 0 / None")
            fresh = scope.rng is not None and not scope.tree
            slices = []
            for i in range(n):
                if fresh:
                    sub = _Scope({}, np.random.default_rng(scope.rng.integers(2 ** 63)))
                else:
                    sub = _Scope(J.tree_map(lambda a: J.asarray(a)[i], scope.tree))
                xi = x if self.in_axes is None else x.select(self.in_axes, i).as_subclass(J.Array)
                outs.append(inner._bind(sub)(xi))
                slices.append(sub.tree)
            if fresh:
                scope.tree.update(J._stack_tree(slices, 0))
        return J._stack_tree(outs, self.out_axes)


def vmap(target, in_axes=0, out_axes=0, axis_size=None, variable_axes=None, split_rngs=None):
    param_axis = (variable_axes or {}).get("params", None)

    def construct(**kwargs):
        return _Lifted(target, kwargs, in_axes, out_axes, axis_size, param_axis)
    return construct


def _lecun_normal(rng, shape):
    return rng.normal(size=shape) * np.sqrt(1.0 / shape[0])


class Dense(Module):
    features: int

    def __call__(self, x):
        x = J.asarray(x)
        kernel = self.param("kernel", _lecun_normal, (x.shape[-1], self.features))
        bias = self.param("bias", lambda rng, shape: np.zeros(shape), (self.features,))
        return (x @ kernel + bias).as_subclass(J.Array)


def absent_layer(name):
    class _Absent:
        def __init__(self, *a, **k):
            raise NotImplementedError(
                "%s: source absent from the reference checkout (flax_rbf is un-vendored; SURVEY.md F2)" % name)
    _Absent.__name__ = name
    return _Absent


def build_modules() -> dict:
    init = types.ModuleType("flax.linen.initializers")
    init.normal = lambda stddev=1e-2: (lambda rng, shape, dtype=None: rng.normal(size=shape) * stddev)
    init.constant = lambda value: (lambda rng, shape, dtype=None: np.full(shape, float(value)))
    init.zeros = lambda rng, shape, dtype=None: np.zeros(shape)
    init.lecun_normal = lambda: _lecun_normal

    linen = types.ModuleType("flax.linen")
    linen.Module, linen.Dense, linen.vmap, linen.initializers = Module, Dense, vmap, init
    linen.relu = lambda a: torch.relu(J.asarray(a)).as_subclass(J.Array)
    linen.softmax = lambda a, axis=-1: torch.softmax(J.asarray(a), dim=axis).as_subclass(J.Array)
    linen.tanh = lambda a: torch.tanh(J.asarray(a)).as_subclass(J.Array)

    flax = types.ModuleType("flax")
    flax.linen = linen
    flax.__path__ = []
    return {"flax": flax, "flax.linen": linen, "flax.linen.initializers": init}
