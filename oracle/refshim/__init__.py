"""TEST INFRASTRUCTURE -- executes the reference's OWN source files, unmodified, on the CPU.

JAX, Flax, chex, optax and flax_rbf cannot be installed here (no wheels, no network),
so the reference cannot be imported as it stands.  This package provides small stand-ins for exactly
the surface of those libraries that the hot path's source touches, backed by torch-CPU tensors
(float64 by default; autograd plays jax.grad), and registers them in ``sys.modules`` so that

    src/irbfn_mpc/model.py                    (_region_activation, WCRBFNet, DeeperWCRBFNet, ClusterWCRBFNet)
    src/irbfn_mpc/dynamics.py                 (integrate_*_mult, dynamic_*_onestep_aux)
    deprecated/.../flax_rbf/flax_rbf.py       (RBFLayer, gaussian, inverse_quadratic, ...)
    scripts/train_nmpc.py, train_nmpc_frenet.py   (the nested train_step_* functions, cut out with ``ast``)

are imported / compiled from where they lie under /root/reference and RUN.  The values they produce
are what ``tests/golden/*.npz`` holds (``tests/golden/make_golden.py``); ``oracle/irbfn_oracle.py`` and
``oracle/dynamics_oracle.py`` are checked against them (``tests/test_oracle_vs_ref.py``).

What is reference code and what is not:
  * every arithmetic statement on the path is the reference's (its files are imported with
    ``importlib`` from /root/reference; nothing is copied into this repo);
  * the stand-ins implement LIBRARY semantics only: ``jnp.*`` element-wise functions map one-to-one to
    torch functions, ``jax.vmap`` / ``jax.lax.scan`` / ``nn.vmap`` are python loops that stack,
    ``nn.Dense`` is ``x @ kernel + bias`` (flax.linen.Dense's published definition),
    ``optax.l2_loss`` is ``0.5 (p - t)^2``, ``jax.value_and_grad`` is ``torch.autograd.grad``.
  * ``flax_rbf``: the reference's dependency is un-vendored and un-pinned (pyproject.toml:18); the 2022
    snapshot under deprecated/ is loaded AS the module ``flax_rbf.flax_rbf``.  The three layer classes
    the snapshot lacks are registered as placeholders that raise when constructed.

Only ``tests/`` (and the fixture generator under ``tests/golden/``) may import this package; it needs
/root/reference, which exists in the build container only.
"""
from __future__ import annotations

import ast
import importlib
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("IRBFN_REFERENCE_ROOT", "/root/reference")
FLAX_RBF_SNAPSHOT = "deprecated/f1tenth_gym/examples/flax_rbf/flax_rbf/flax_rbf.py"

_SHIM_NAMES = ("jax", "jax.numpy", "jax.lax", "jax.tree_util", "jax.random", "chex", "optax",
               "flax", "flax.linen", "flax.linen.initializers", "flax_rbf", "flax_rbf.flax_rbf",
               "irbfn_mpc", "irbfn_mpc.model", "irbfn_mpc.dynamics")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "src", "irbfn_mpc", "model.py"))


def _load_file_as(name: str, path: str) -> types.ModuleType:
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class Reference:
    """Handle on the reference modules loaded under the stand-ins."""

    def __init__(self, model, dynamics, flax_rbf, jax, jnp, optax):
        self.model, self.dynamics, self.flax_rbf = model, dynamics, flax_rbf
        self.jax, self.jnp, self.optax = jax, jnp, optax

    def basis(self, name: str):
        """The reference's basis function object, resolved the way the scripts do (eval of the config's
        name, scripts/train_nmpc.py:207)."""
        return getattr(self.flax_rbf, name)


_loaded: Reference | None = None


def load(dtype: str = "float64") -> Reference:
    """Install the stand-ins and import the reference's modules (once per process)."""
    global _loaded
    from . import jax_shim, flax_shim
    jax_shim.set_dtype(dtype)
    if _loaded is not None:
        return _loaded
    if not available():
        raise FileNotFoundError("%s not found: the reference can only be executed in the build container"
                                % REFERENCE_ROOT)
    clash = [n for n in _SHIM_NAMES if n in sys.modules]
    if clash:
        raise RuntimeError("refuse to shadow already-imported modules: %s" % clash)
    jax = jax_shim.build_modules()              # jax, jax.numpy, jax.lax, jax.tree_util, jax.random, chex, optax
    sys.modules.update(jax)
    sys.modules.update(flax_shim.build_modules())
    # flax_rbf: the vendored snapshot, loaded under the name the reference imports it by
    pkg = types.ModuleType("flax_rbf")
    pkg.__path__ = []
    sys.modules["flax_rbf"] = pkg
    frbf = _load_file_as("flax_rbf.flax_rbf", os.path.join(REFERENCE_ROOT, FLAX_RBF_SNAPSHOT))
    pkg.flax_rbf = frbf
    for missing in ("WarmstartCenteredRBFLayer", "FixedCenteredRBFLayer", "FixedWidthCenteredRBFLayer"):
        if not hasattr(frbf, missing):
            setattr(frbf, missing, flax_shim.absent_layer(missing))
    src = os.path.join(REFERENCE_ROOT, "src", "irbfn_mpc")
    pkg2 = types.ModuleType("irbfn_mpc")
    pkg2.__path__ = [src]
    sys.modules["irbfn_mpc"] = pkg2
    model = _load_file_as("irbfn_mpc.model", os.path.join(src, "model.py"))
    dynamics = _load_file_as("irbfn_mpc.dynamics", os.path.join(src, "dynamics.py"))
    _loaded = Reference(model, dynamics, frbf, jax["jax"], jax["jax.numpy"], jax["optax"])
    return _loaded


def nested_function(script_rel: str, outer: str, name: str, namespace: dict):
    """Compile the nested function ``name`` defined inside ``outer`` of a reference script, unmodified,
    with ``namespace`` as its globals (the closure variables the script's main() would have held:
    wcrbf, dyn_params, ...).  The script itself is never run: its main() loads absent data tables."""
    path = os.path.join(REFERENCE_ROOT, script_rel)
    with open(path) as f:
        tree = ast.parse(f.read(), filename=path)
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == outer:
            for sub in ast.walk(node):
                if isinstance(sub, ast.FunctionDef) and sub.name == name:
                    mod = ast.Module(body=[sub], type_ignores=[])
                    code = compile(mod, path, "exec")       # keeps the script's own line numbers
                    exec(code, namespace)
                    return namespace[name], (sub.lineno, sub.end_lineno)
    raise LookupError("%s.%s not found in %s" % (outer, name, script_rel))
