"""The oracle is pinned to values produced by the REFERENCE'S OWN SOURCE.

tests/golden/*.npz were written by tests/golden/make_golden.py, which imports
``src/irbfn_mpc/model.py``, ``src/irbfn_mpc/dynamics.py``, the ``flax_rbf`` snapshot and the nested
``train_step_*`` functions of the two training scripts unmodified from /root/reference and runs them
under the torch-CPU library stand-ins of ``oracle/refshim`` (JAX itself is not installable).

Part 1 (runs anywhere, including the GPU box): the NumPy / torch oracles reproduce every committed
reference-executed value: outputs to 1e-12 (norm-wise), gradients to 1e-9 per leaf.
Part 2 (only where /root/reference exists, i.e. the build container): the reference is executed AGAIN,
live, and must reproduce the committed fixtures -- so the fixtures cannot drift from the source they
claim to come from; the parameter tree its own ``init`` creates must be the tree the product's
``irbfn_b200.model`` creates.

Gradients: the reference has no analytic gradient code -- they are autodiff of its forward (F7); the
fixtures hold torch autograd through the reference's executed forward, which is that same definition.
"""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, golden_paths, load_golden, rel_l2, rel_max
from oracle import dynamics_oracle as dyn
from oracle import irbfn_oracle as orc
from oracle import refshim

SOURCE = "reference code executed under oracle/refshim"
LEAF = {"rbf_list/centers": "centers", "rbf_list/log_sigs": "log_sigs", "linear/kernel": "kernel", "linear/bias": "bias"}


def _ids(paths):
    return [os.path.basename(p)[:-4] for p in paths]


# ------------------------------------------------------------------------------------------------
# Part 1: oracle vs committed reference-executed values
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_paths(), ids=_ids(golden_paths()))
def test_fixture_is_reference_executed_and_oracle_matches(path):
    cfg, params, arr = load_golden(path)
    assert cfg["source"] == SOURCE
    assert set(cfg["reference_sha256"]) >= {"src/irbfn_mpc/model.py", refshim.FLAX_RBF_SNAPSHOT}
    x = arr["x"].astype(np.float64)
    kind = cfg["kind"]
    if kind == "deeper":
        assert rel_max(orc.deeper_forward(x, params, cfg), arr["y64"]) < 1e-12
        return
    if kind == "cluster":
        y, logits = orc.cluster_forward(x, params, cfg)
        assert np.abs(logits - arr["logits64"]).max() < 1e-10
        gate, _ = orc.softmax_gate(x, params["cluster"]["kernel"], params["cluster"]["bias"])
        rb = params["rbf_list"]
        g = orc.backward(x, rb["centers"], rb["log_sigs"], params["linear"]["kernel"], params["linear"]["bias"],
                         gate, 2 * y / y.size, orc.BASIS_IDS[cfg["basis_func"]])
        dlogits = gate * (g["gamma"] - (g["gamma"] * gate).sum(-1, keepdims=True))
        assert rel_l2(x.T @ dlogits, arr["g/cluster/kernel"]) < 1e-9
        assert rel_l2(dlogits.sum(0), arr["g/cluster/bias"]) < 1e-9
        g["x"] = g["x"] + dlogits @ np.asarray(params["cluster"]["kernel"], np.float64).T
    else:
        y = orc.wcrbf_forward(x, params, cfg)
        g = orc.wcrbf_vjp(x, params, cfg, 2 * y / y.size)
    assert rel_max(y, arr["y64"]) < 1e-12
    for key, leaf in LEAF.items():
        assert rel_l2(g[leaf], arr["g/" + key]) < 1e-9, key
    assert rel_l2(g["x"], arr["g/x"]) < 1e-9


def test_dynamics_oracle_matches_reference_executed_values():
    z = np.load(os.path.join(GOLDEN_DIR, "dynamics_ref.npz"))
    meta = json.loads(str(z["meta_json"]))
    assert meta["source"] == SOURCE and "src/irbfn_mpc/dynamics.py" in meta["reference_sha256"]
    P = z["dyn_params"]
    assert np.array_equal(P, dyn.default_dyn_params())
    assert np.abs(dyn.integrate_frenet_mult(z["frenet_mult_in"], P).numpy() - z["frenet_mult_out"]).max() < 1e-12
    assert np.abs(dyn.integrate_st_mult(z["st_mult_in"], P).numpy() - z["st_mult_out"]).max() < 1e-12
    assert np.abs(dyn.dynamic_st_onestep_aux(z["st_aux_in"], P).numpy() - z["st_aux_out"]).max() < 1e-12
    for O in (2, 10):      # O = 2 exercises JAX's clamped static index x_u[9] (dynamics.py:327)
        out = dyn.dynamic_frenet_onestep_aux(z["frenet_aux_in_o%d" % O], P).numpy()
        assert np.abs(out - z["frenet_aux_out_o%d" % O]).max() < 1e-12
    for mode in range(4):
        assert "train_step" in meta["lines"]["mode%d" % mode]
        x, y, yp = z["step%d_x" % mode], z["step%d_y" % mode], z["step%d_ypred" % mode]
        loss, pl, il, g = dyn.training_loss_and_grad(mode, x, yp, y, P)
        ref = z["step%d_loss" % mode]
        assert abs(loss - ref[0]) <= 1e-12 * max(1.0, abs(ref[0]))
        if len(ref) == 3:
            assert abs(pl - ref[1]) <= 1e-12 and abs(il - ref[2]) <= 1e-10
        assert rel_l2(g, z["step%d_dypred" % mode]) < 1e-10


def test_full_training_step_gradients_match_reference_executed_values():
    """network -> hstack -> dynamics -> |.|.mean() -> value_and_grad, as the scripts' steps do it."""
    z = np.load(os.path.join(GOLDEN_DIR, "train_step_ref.npz"))
    meta = json.loads(str(z["meta_json"]))
    assert meta["source"] == SOURCE
    P = z["dyn_params"]
    for tag, case in meta["cases"].items():
        mode, cfg = int(tag[4:]), case["cfg"]
        params = {"rbf_list": {"centers": z[tag + "/p/rbf_list/centers"], "log_sigs": z[tag + "/p/rbf_list/log_sigs"]},
                  "linear": {"kernel": z[tag + "/p/linear/kernel"], "bias": z[tag + "/p/linear/bias"]}}
        x, y = z[tag + "/x"].astype(np.float64), z[tag + "/y"]
        yp = orc.wcrbf_forward(x, params, cfg)
        loss, _, _, dyp = dyn.training_loss_and_grad(mode, x, yp, y, P)
        assert abs(loss - z[tag + "/loss"][0]) <= 1e-11 * max(1.0, abs(loss))
        g = orc.wcrbf_vjp(x, params, cfg, dyp)
        for key, leaf in LEAF.items():
            assert rel_l2(g[leaf], z[tag + "/g/" + key]) < 1e-9, (tag, key)


# ------------------------------------------------------------------------------------------------
# Part 2: execute the reference again, live (build container only)
# ------------------------------------------------------------------------------------------------
needs_reference = pytest.mark.skipif(not refshim.available(), reason="/root/reference not present on this box")


@pytest.fixture(scope="module")
def ref():
    return refshim.load("float64")


def _module(ref, cfg):
    import sys
    sys.path.insert(0, GOLDEN_DIR)
    import make_golden
    return make_golden, make_golden.build_reference_module(ref, cfg, cfg["kind"])


@needs_reference
@pytest.mark.parametrize("name", ["c1b_nmpc_4_regions", "c2_frenet_12x100", "c2b_frenet_1x1000_o10", "deeper_1x100",
                                  "cluster_500x10", "warmstart_27x500_o10"])
def test_live_reference_execution_reproduces_fixture(ref, name):
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, name + ".npz"))
    mg, module = _module(ref, cfg)
    assert cfg["reference_sha256"] == mg.ref_digests(), "reference sources changed since the fixtures were made"
    n = 8 if cfg["num_regions"] > 100 else 48
    p64 = ref.jax.tree_util.tree_map(lambda a: np.asarray(a, np.float64), params)
    out = module.apply({"params": p64}, ref.jnp.array(arr["x"][:n].astype(np.float64)))
    y = (out[0] if cfg["kind"] == "cluster" else out).numpy()
    assert np.abs(y - arr["y64"][:n]).max() <= 1e-13 * np.abs(arr["y64"]).max()


@needs_reference
def test_reference_region_activation_is_what_the_oracle_restates(ref):
    """_region_activation (model.py:42-95) called directly, overlapping 4-region shipped bounds."""
    cfg, _, arr = load_golden(os.path.join(GOLDEN_DIR, "c1b_nmpc_4_regions.npz"))
    x = arr["x"][:256].astype(np.float64)
    g_ref = ref.model._region_activation(ref.jnp.array(x), cfg["num_regions"], len(cfg["activation_idx"]),
                                         cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                         cfg["dimension_ranges"]).numpy()
    g = orc.region_activation(x, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                              cfg["dimension_ranges"])
    assert g_ref.shape == g.shape and np.abs(g - g_ref).max() < 1e-15
    assert (g.sum(1) > 1.0 + 1e-6).any(), "shipped regions overlap: the gate is not normalised (F4)"


@needs_reference
def test_reference_rbf_layer_and_bases(ref):
    """RBFLayer.__call__ + gaussian / inverse_quadratic (flax_rbf.py:34-37,49-52,258-285) on one query."""
    rng = np.random.default_rng(3)
    c, ls, x = rng.normal(size=(1, 7, 5)), rng.normal(size=(1, 7)) * 0.5, rng.normal(size=(5,))
    for name, bid in (("gaussian", orc.GAUSSIAN), ("inverse_quadratic", orc.INVERSE_QUADRATIC)):
        layer = ref.flax_rbf.RBFLayer(in_features=5, num_kernels=7, basis_func=ref.basis(name))
        phi = layer.apply({"params": {"centers": c[0], "log_sigs": ls[0]}}, ref.jnp.array(x)).numpy()
        assert np.abs(orc.rbf_activations(x[None], c, ls, bid)[0, 0] - phi).max() < 1e-15


@needs_reference
@pytest.mark.parametrize("name", ["c2_frenet_12x100", "deeper_1x100", "cluster_500x10"])
def test_reference_init_tree_is_the_products_tree(ref, name):
    """The nested scopes the reference's own setup() creates (model.py:143-167) == irbfn_b200.model's init."""
    import torch
    from irbfn_b200 import basis as pbasis
    from irbfn_b200 import model as pm
    cfg, _, _ = load_golden(os.path.join(GOLDEN_DIR, name + ".npz"))
    if cfg["num_regions"] > 100:
        cfg = dict(cfg, num_regions=6)
    _, module = _module(ref, cfg)
    tree_ref = module.init(0, ref.jnp.ones((2, cfg["in_features"])))
    shapes_ref = ref.jax.tree_util.tree_map(lambda a: tuple(a.shape), tree_ref)
    bf = getattr(pbasis, cfg["basis_func"])
    common = dict(in_features=cfg["in_features"], out_features=cfg["out_features"], num_kernels=cfg["num_kernels"],
                  basis_func=bf, num_regions=cfg["num_regions"])
    if cfg["kind"] == "cluster":
        net = pm.ClusterWCRBFNet(**common)
    else:
        cls = pm.DeeperWCRBFNet if cfg["kind"] == "deeper" else pm.WCRBFNet
        net = cls(**common, lower_bounds=cfg["lower_bounds"], upper_bounds=cfg["upper_bounds"],
                  dimension_ranges=cfg["dimension_ranges"], activation_idx=cfg["activation_idx"], delta=cfg["delta"])
    tree = net.init(0, torch.ones((2, cfg["in_features"])), device="cpu")
    shapes = ref.jax.tree_util.tree_map(lambda a: tuple(a.shape), tree)
    assert shapes == shapes_ref
    # initial values the reference's initialisers fix exactly (flax_rbf.py:252-256; Dense bias)
    assert float(tree_ref["params"]["rbf_list"]["log_sigs"].abs().max()) == 0.0
    assert float(tree["params"]["rbf_list"]["log_sigs"].abs().max()) == 0.0
