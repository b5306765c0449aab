"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN SOURCE on its shipped checkpoints.

Runs ONLY in the build container (needs /root/reference); the GPU box uses the committed fixtures.
JAX / Flax / chex / optax / flax_rbf are not installable here, so the reference's files
are imported unmodified under the torch-CPU stand-ins of ``oracle/refshim`` (which implement library
semantics only -- element-wise jnp functions, vmap/scan loops, nn.Dense, autograd as jax.grad):

    model fixtures     <Module> = src/irbfn_mpc/model.py: WCRBFNet | DeeperWCRBFNet | ClusterWCRBFNet,
                       its RBF layer and basis function from the flax_rbf snapshot
                       (deprecated/f1tenth_gym/examples/flax_rbf/flax_rbf/flax_rbf.py);
                       y64   = <Module>(**config).apply({'params': checkpoint}, x)          float64
                       g/*   = jax.value_and_grad(lambda p, x: mean(apply(p, x)**2)) w.r.t. every leaf and x
    dynamics fixture   src/irbfn_mpc/dynamics.py functions and the nested train_step_* functions of
                       scripts/train_nmpc.py / scripts/train_nmpc_frenet.py (cut out with ``ast``, run as is)

Every value written is what the reference's code returned.  Before a fixture is written the NumPy
oracle (oracle/irbfn_oracle.py, oracle/dynamics_oracle.py) must agree with it to 1e-12 (outputs) /
1e-9 (gradients); tests/test_oracle_vs_ref.py repeats that check from the committed files, and
re-executes the reference when /root/reference is present.

Fixture contents (per checkpoint):
  cfg_json              module attributes (num_regions, bounds, delta, ranges, basis name ...), plus
                        "source": how the values were produced and sha256 of the reference files run
  p/<leaf path>         parameters cast to float32 (what the reference's x64-disabled inference
                        path evaluates)
  x                     (N,D) float32 queries: uniform in the config box, default_rng(0)
  y64                   (N,O) float64 outputs
  g/<leaf>, g/x         float64 gradients of L = mean(y**2) w.r.t. every leaf and x
  [logits64]            ClusterWCRBFNet logits

Usage:  python tests/golden/make_golden.py [fixture-name ...]
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import time

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import irbfn_oracle as orc  # noqa: E402
from oracle import dynamics_oracle as dyn  # noqa: E402
from oracle import refshim  # noqa: E402

REF = os.path.join(refshim.REFERENCE_ROOT, "scripts")
OUT = os.path.dirname(os.path.abspath(__file__))
N = 1024
SOURCE = "reference code executed under oracle/refshim"

# (fixture name, config, checkpoint, module kind, queries)
CASES = [
    ("c1_cartesian_1x1000", "dnmpc_1regions_newnewdata_1stepst_l2_newarch_ksint", "checkpoint_9999", "wcrbf", N),
    ("c1b_nmpc_4_regions", "nmpc_4_regions", "checkpoint_0", "wcrbf", N),
    ("c2_frenet_12x100", "dnmpc_12regions_frenet_l1_bigdata", "checkpoint_9500", "wcrbf", N),
    ("c2b_frenet_1x1000_o10", "dnmpc_1regions_frenet_l1_bigdata_5stepint_highk", "checkpoint_3900", "wcrbf", N),
    ("nmpc_1_region_iq", "nmpc_1_region", "checkpoint_0", "wcrbf", N),
    ("dnmpc_4regions_iq", "dnmpc_4regions_correctdata", "checkpoint_999", "wcrbf", N),
    ("dnmpc_128regions_narrow", "dnmpc_128regions", "checkpoint_600", "wcrbf", N),
    ("dnmpc_108regions_wide", "dnmpc_108regions_frenet_l1_bigdata", "checkpoint_6000", "wcrbf", N),
    ("deeper_1x100", "dnmpc_1regions_frenet_l1_bigdata_5stepint_deeper", "checkpoint_9999", "deeper", N),
    ("cluster_500x10", "dnmpc_500_clusters", "checkpoint_0", "cluster", 256),   # (n, 500) logits: keep it small
    # round 2: the remaining evaluable shipped shapes of SURVEY 8(c)
    ("dnmpc_9regions_multiv", "dnmpc_9regions_newnewdata_1stepst_l1_newarch_ksint_multiv", "checkpoint_9999", "wcrbf", N),
    ("warmstart_1x500_o10", "dnmpc_warmstart_constraint_centers_mode", "checkpoint_8400", "wcrbf", N),
    ("warmstart_27x500_o10", "dnmpc_27regions_warmstart_constraint_centers_mode", "checkpoint_1400", "wcrbf", 512),
]

CFG_KEYS = ("in_features", "out_features", "num_kernels", "num_regions", "basis_func",
            "lower_bounds", "upper_bounds", "dimension_ranges", "activation_idx", "delta")

REF_FILES = ("src/irbfn_mpc/model.py", "src/irbfn_mpc/dynamics.py", refshim.FLAX_RBF_SNAPSHOT,
             "scripts/train_nmpc.py", "scripts/train_nmpc_frenet.py")


def ref_digests():
    out = {}
    for rel in REF_FILES:
        with open(os.path.join(refshim.REFERENCE_ROOT, rel), "rb") as f:
            out[rel] = hashlib.sha256(f.read()).hexdigest()[:16]
    return out


def flatten(tree, prefix=""):
    out = {}
    for k, v in tree.items():
        if isinstance(v, dict):
            out.update(flatten(v, prefix + k + "/"))
        else:
            out[prefix + k] = np.asarray(v)
    return out


def unflatten(flat):
    tree = {}
    for k, v in flat.items():
        node = tree
        parts = k.split("/")
        for p in parts[:-1]:
            node = node.setdefault(p, {})
        node[parts[-1]] = v
    return tree


def build_reference_module(ref, cfg, kind):
    """Construct the reference's module the way scripts/train_nmpc_frenet.py:246-307 does (keyword
    arguments from the run's yaml; basis function resolved by name from flax_rbf)."""
    basis = ref.basis(cfg["basis_func"])
    if kind == "cluster":                                               # train_nmpc_frenet.py:262-270
        return ref.model.ClusterWCRBFNet(in_features=cfg["in_features"], out_features=cfg["out_features"],
                                         num_kernels=cfg["num_kernels"], basis_func=basis,
                                         num_regions=cfg["num_regions"])
    cls = ref.model.DeeperWCRBFNet if kind == "deeper" else ref.model.WCRBFNet
    return cls(in_features=cfg["in_features"], out_features=cfg["out_features"], num_kernels=cfg["num_kernels"],
               basis_func=basis, num_regions=cfg["num_regions"], lower_bounds=cfg["lower_bounds"],
               upper_bounds=cfg["upper_bounds"], dimension_ranges=cfg["dimension_ranges"],
               activation_idx=cfg["activation_idx"], delta=cfg["delta"])


def reference_eval(ref, module, kind, params, x):
    """-> (y, grads of mean(y^2) keyed like the fixture, logits | None), all float64 NumPy, all computed
    by the reference's own forward code (autodiff over it for the gradients)."""
    jnp = ref.jnp
    variables = {"params": params}
    out = module.apply(variables, jnp.array(x))
    y, logits = (out if kind == "cluster" else (out, None))

    def loss_fn(p_and_x):
        o = module.apply(p_and_x[0], p_and_x[1])
        o = o[0] if kind == "cluster" else o
        return (o ** 2).mean()

    _, (gp, gx) = ref.jax.value_and_grad(loss_fn)((variables, jnp.array(x)))
    grads = flatten(ref.jax.tree_util.tree_map(lambda t: t.numpy(), gp["params"]))
    grads["x"] = gx.numpy()
    return y.numpy(), grads, (None if logits is None else logits.numpy())


def oracle_eval(x, params, cfg, kind):
    """float64 oracle outputs and (where the oracle has an analytic VJP) gradients of mean(y^2)."""
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    rb = params["rbf_list"]
    grads = {}
    logits = None
    if kind == "cluster":
        y, logits = orc.cluster_forward(x, params, cfg)
        gate, _ = orc.softmax_gate(x, params["cluster"]["kernel"], params["cluster"]["bias"])
        dy = 2 * y / y.size
        g = orc.backward(x, rb["centers"], rb["log_sigs"], params["linear"]["kernel"],
                         params["linear"]["bias"], gate, dy, basis)
        dlogits = gate * (g["gamma"] - (g["gamma"] * gate).sum(-1, keepdims=True))
        grads["cluster/kernel"] = x.T @ dlogits
        grads["cluster/bias"] = dlogits.sum(0)
        grads["x"] = g["x"] + dlogits @ np.asarray(params["cluster"]["kernel"], np.float64).T
    elif kind == "deeper":
        return orc.deeper_forward(x, params, cfg), None, None
    else:
        y = orc.wcrbf_forward(x, params, cfg)
        dy = 2 * y / y.size
        g = orc.wcrbf_vjp(x, params, cfg, dy)
        grads["x"] = g["x"]
    grads["rbf_list/centers"] = g["centers"]
    grads["rbf_list/log_sigs"] = g["log_sigs"]
    grads["linear/kernel"] = g["kernel"]
    grads["linear/bias"] = g["bias"]
    return y, grads, logits


def load_case(conf, ckpt, kind):
    cfg_all = yaml.safe_load(open(os.path.join(REF, "configs", conf + ".yaml")))
    cfg = {k: cfg_all[k] for k in CFG_KEYS if k in cfg_all}
    cfg["kind"] = kind
    cfg["source_config"] = "scripts/configs/%s.yaml" % conf
    cfg["source_checkpoint"] = "scripts/ckpts/%s/%s" % (conf, ckpt)
    tree = orc.load_flax_checkpoint(os.path.join(REF, "ckpts", conf, ckpt))["params"]["params"]
    flat32 = {k: v.astype(np.float32) for k, v in flatten(tree).items()}
    return cfg, flat32


def make_model_fixture(ref, name, conf, ckpt, kind, n):
    t0 = time.time()
    cfg, flat32 = load_case(conf, ckpt, kind)
    cfg["source"] = SOURCE
    cfg["reference_sha256"] = ref_digests()
    if "warmstart" in conf:
        cfg["note"] = ("trained with WarmstartCenteredRBFLayer, whose source is absent from the reference "
                       "(SURVEY F2); evaluated with the snapshot's RBFLayer arithmetic on the same leaves")
    params = unflatten({k: v.astype(np.float64) for k, v in flat32.items()})
    x32 = orc.uniform_queries(cfg, 4096, seed=0)[:n].astype(np.float32)
    x = x32.astype(np.float64)

    module = build_reference_module(ref, cfg, kind)
    y, grads, logits = reference_eval(ref, module, kind, params, x)
    for k, g in grads.items():
        if not np.isfinite(g).all():
            raise AssertionError("%s: reference gradient %s not finite (x == centre?)" % (name, k))

    # the oracle must agree before anything is written
    oy, ograds, ologits = oracle_eval(x, params, cfg, kind)
    err_y = np.abs(oy - y).max() / np.abs(y).max()
    assert err_y < 1e-12, (name, err_y)
    worst = 0.0
    if ograds is not None:
        for k, g in ograds.items():
            e = np.linalg.norm(g - grads[k]) / max(np.linalg.norm(grads[k]), 1e-300)
            worst = max(worst, e)
            assert e < 1e-9, (name, k, e)
    if logits is not None:
        assert np.abs(ologits - logits).max() < 1e-10

    out = {"cfg_json": np.array(json.dumps(cfg)), "x": x32, "y64": y}
    for k, v in flat32.items():
        out["p/" + k] = v
    for k, v in grads.items():
        out["g/" + k] = v
    if logits is not None:
        out["logits64"] = logits
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **out)
    print("%-26s kind=%-7s n=%4d  oracle vs reference: y %.1e  grads %.1e   %.0f KB  %.0f s" % (
        name, kind, n, err_y, worst, os.path.getsize(path) / 1024, time.time() - t0), flush=True)


# --------------------------------------------------------------------------------------------------
# dynamics + training-step fixture
# --------------------------------------------------------------------------------------------------
class _State:
    """The two members of flax.training.train_state.TrainState the reference's steps touch."""

    def __init__(self, params):
        self.params, self.grads = params, None

    def apply_gradients(self, grads):
        self.grads = grads
        return self


class _Passthrough:
    """wcrbf stand-in whose 'parameters' ARE the predictions: the gradients the reference's step hands to
    apply_gradients are then d loss / d y_pred, produced by the reference's own loss code."""

    @staticmethod
    def apply(params, x):
        return params["y_pred"]


STEPS = {  # mode -> (script, nested step, D, O)
    dyn.MODE_FRENET_ONEINT: ("scripts/train_nmpc_frenet.py", "train_step_oneint", 8, 2),
    dyn.MODE_FRENET_FULLINT: ("scripts/train_nmpc_frenet.py", "train_step_fullint", 8, 10),
    dyn.MODE_ST_ONEINT: ("scripts/train_nmpc.py", "train_step_oneint", 7, 2),
    dyn.MODE_ST_FULLINT: ("scripts/train_nmpc.py", "train_step_fullint", 7, 10),
}


def step_namespace(ref, wcrbf, P):
    ns = {"jax": ref.jax, "jnp": ref.jnp, "optax": ref.optax, "wcrbf": wcrbf, "dyn_params": ref.jnp.array(P)}
    for fn in ("integrate_frenet_mult", "dynamic_frenet_onestep_aux", "integrate_st_mult", "dynamic_st_onestep_aux"):
        ns[fn] = getattr(ref.dynamics, fn)
    return ns


def f32(a):
    """Inputs are rounded to float32 so that a float32 device run sees bit-identical inputs."""
    return np.asarray(a, np.float32).astype(np.float64)


def dynamics_inputs(rng, B, D):
    lo = np.array([-0.2, -0.3, 1.0, -1.0, 3.0, -2.6, -1.0, -0.1])[:D]
    hi = np.array([2.0, 0.3, 7.0, 1.0, 7.0, 2.6, 1.0, 0.1])[:D]
    return f32(rng.uniform(lo, hi, (B, D)))


def make_dynamics_fixture(ref):
    t0 = time.time()
    rng = np.random.default_rng(11)
    P = dyn.default_dyn_params()
    B = 512
    out = {"dyn_params": P}
    meta = {"source": SOURCE, "reference_sha256": ref_digests(), "lines": {}}
    jnp = ref.jnp

    # the four integrators, on their own (dynamics.py:94-100,103-187,284-290,293-383)
    fr = f32(np.concatenate([dynamics_inputs(rng, B, 8)[:, [0, 0, 1, 2, 3, 5, 6, 7]], rng.normal(size=(B, 10)) * 2], 1))
    st = f32(np.concatenate([rng.uniform([-1, -1, -0.5, 0.5, -1, -1, -0.3], [1, 1, 0.5, 8.0, 1, 1, 0.3], (B, 7)),
                             rng.normal(size=(B, 10)) * 3], 1))
    out["frenet_mult_in"], out["st_mult_in"] = fr, st
    out["frenet_mult_out"] = ref.dynamics.integrate_frenet_mult(jnp.array(fr), jnp.array(P)).numpy()
    out["st_mult_out"] = ref.dynamics.integrate_st_mult(jnp.array(st), jnp.array(P)).numpy()
    for O in (2, 10):
        fa = f32(np.concatenate([dynamics_inputs(rng, B, 8)[:, [0, 1, 2, 3, 5, 6, 7]], rng.normal(size=(B, O)) * 2], 1))
        out["frenet_aux_in_o%d" % O] = fa
        out["frenet_aux_out_o%d" % O] = ref.dynamics.dynamic_frenet_onestep_aux(jnp.array(fa), jnp.array(P)).numpy()
    sa = st[:, :9]
    out["st_aux_in"] = sa
    out["st_aux_out"] = ref.dynamics.dynamic_st_onestep_aux(jnp.array(sa), jnp.array(P)).numpy()

    assert np.abs(dyn.integrate_frenet_mult(fr, P).numpy() - out["frenet_mult_out"]).max() < 1e-12
    assert np.abs(dyn.integrate_st_mult(st, P).numpy() - out["st_mult_out"]).max() < 1e-12
    assert np.abs(dyn.dynamic_st_onestep_aux(sa, P).numpy() - out["st_aux_out"]).max() < 1e-12
    for O in (2, 10):
        assert np.abs(dyn.dynamic_frenet_onestep_aux(out["frenet_aux_in_o%d" % O], P).numpy()
                      - out["frenet_aux_out_o%d" % O]).max() < 1e-12

    # the four training steps' losses and d loss / d y_pred, from the scripts' own nested functions
    for mode, (script, fn_name, D, O) in STEPS.items():
        step, lines = refshim.nested_function(script, "main", fn_name, step_namespace(ref, _Passthrough, P))
        meta["lines"]["mode%d" % mode] = "%s:%d-%d %s" % (script, lines[0], lines[1], fn_name)
        x = dynamics_inputs(rng, B, D)
        y = f32(rng.normal(size=(B, O)))
        yp = f32(y + rng.normal(size=(B, O)) * 0.7)
        state = _State({"y_pred": jnp.array(yp)})
        res = step(state, jnp.array(x), jnp.array(y))
        loss = float(res[1])
        g = state.grads["y_pred"].numpy()
        out["step%d_x" % mode], out["step%d_y" % mode], out["step%d_ypred" % mode] = x, y, yp
        out["step%d_loss" % mode] = np.array([float(r) for r in res[1:]])     # loss[, pred_loss, int_loss]
        out["step%d_dypred" % mode] = g
        ol, opl, oil, og = dyn.training_loss_and_grad(mode, x, yp, y, P)
        assert abs(ol - loss) <= 1e-12 * max(1, abs(loss)), (mode, ol, loss)
        assert np.linalg.norm(og - g) <= 1e-10 * np.linalg.norm(g), (mode,)
        if len(res) > 2:
            assert abs(opl - float(res[2])) < 1e-12 and abs(oil - float(res[3])) < 1e-10

    out["meta_json"] = np.array(json.dumps(meta))
    path = os.path.join(OUT, "dynamics_ref.npz")
    np.savez_compressed(path, **out)
    print("dynamics_ref: integrators + 4 training steps executed from the reference, oracle agrees  %.0f KB  %.0f s"
          % (os.path.getsize(path) / 1024, time.time() - t0), flush=True)


def make_train_step_fixture(ref):
    """One FULL reference training step per loss mode with the reference's real module inside: parameter
    gradients of (network -> dynamics loss), which is what irbfn_b200.train reproduces on the device."""
    t0 = time.time()
    P = dyn.default_dyn_params()
    rng = np.random.default_rng(23)
    B = 128
    pick = {dyn.MODE_FRENET_ONEINT: ("dnmpc_12regions_frenet_l1_bigdata", "checkpoint_9500"),
            dyn.MODE_FRENET_FULLINT: ("dnmpc_1regions_frenet_l1_bigdata_5stepint_highk", "checkpoint_3900"),
            dyn.MODE_ST_ONEINT: ("dnmpc_1regions_newnewdata_1stepst_l2_newarch_ksint", "checkpoint_9999")}
    out = {"dyn_params": P}
    meta = {"source": SOURCE, "reference_sha256": ref_digests(), "cases": {}}
    for mode, (conf, ckpt) in pick.items():
        script, fn_name, D, O = STEPS[mode]
        cfg, flat32 = load_case(conf, ckpt, "wcrbf")
        assert cfg["in_features"] == D and cfg["out_features"] == O, (conf, cfg["in_features"], cfg["out_features"])
        params = unflatten({k: v.astype(np.float64) for k, v in flat32.items()})
        module = build_reference_module(ref, cfg, "wcrbf")
        step, lines = refshim.nested_function(script, "main", fn_name, step_namespace(ref, module, P))
        x = orc.uniform_queries(cfg, B, seed=5).astype(np.float32).astype(np.float64)
        y = f32(rng.normal(size=(B, O)))
        state = _State({"params": params})
        res = step(state, ref.jnp.array(x), ref.jnp.array(y))
        grads = flatten(ref.jax.tree_util.tree_map(lambda t: t.numpy(), state.grads["params"]))
        tag = "mode%d" % mode
        meta["cases"][tag] = {"cfg": cfg, "step": "%s:%d-%d %s" % (script, lines[0], lines[1], fn_name)}
        out[tag + "/x"], out[tag + "/y"] = x.astype(np.float32), y
        out[tag + "/loss"] = np.array([float(r) for r in res[1:]])
        for k, v in flat32.items():
            out[tag + "/p/" + k] = v
        for k, v in grads.items():
            out[tag + "/g/" + k] = v
        # oracle: forward -> dynamics loss cotangent -> analytic VJP
        yp = orc.wcrbf_forward(x, params, cfg)
        ol, _, _, dyp = dyn.training_loss_and_grad(mode, x, yp, y, P)
        og = orc.wcrbf_vjp(x, params, cfg, dyp)
        assert abs(ol - float(res[1])) <= 1e-11 * max(1.0, abs(ol)), (mode, ol, float(res[1]))
        for leaf, key in (("centers", "rbf_list/centers"), ("log_sigs", "rbf_list/log_sigs"),
                          ("kernel", "linear/kernel"), ("bias", "linear/bias")):
            e = np.linalg.norm(og[leaf] - grads[key]) / np.linalg.norm(grads[key])
            assert e < 1e-9, (mode, leaf, e)
    out["meta_json"] = np.array(json.dumps(meta))
    path = os.path.join(OUT, "train_step_ref.npz")
    np.savez_compressed(path, **out)
    print("train_step_ref: %d full reference training steps, oracle agrees  %.0f KB  %.0f s"
          % (len(pick), os.path.getsize(path) / 1024, time.time() - t0), flush=True)


def make_checkpoint_fixture():
    """Reference-held test data for the product's checkpoint codec (irbfn_b200/checkpoint.py):
      * the raw bytes of one small shipped checkpoint (scripts/ckpts/nmpc_4_regions/checkpoint_0, 70 KB),
        so the reader decodes REAL flax.serialization output under test;
      * the key / shape / dtype layout of a shipped checkpoint written with the current optimiser chain
        (optax.chain(clip_by_global_norm, adam), scripts/train_nmpc.py:231-233), which
        TrainState.state_dict() must reproduce."""
    import shutil
    src = os.path.join(REF, "ckpts", "nmpc_4_regions", "checkpoint_0")
    shutil.copyfile(src, os.path.join(OUT, "ref_ckpt_nmpc_4_regions_checkpoint_0.msgpack"))
    os.chmod(os.path.join(OUT, "ref_ckpt_nmpc_4_regions_checkpoint_0.msgpack"), 0o644)

    def layout(node):
        if isinstance(node, dict):
            return {str(k): layout(v) for k, v in node.items()}
        a = np.asarray(node)
        return [list(a.shape), a.dtype.name]

    out = {}
    for conf, ckpt in (("dnmpc_12regions_frenet_l1_bigdata", "checkpoint_9500"), ("nmpc_4_regions", "checkpoint_0")):
        out["scripts/ckpts/%s/%s" % (conf, ckpt)] = layout(orc.load_flax_checkpoint(os.path.join(REF, "ckpts", conf, ckpt)))
    with open(os.path.join(OUT, "ckpt_layouts.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("checkpoint fixtures: raw nmpc_4_regions/checkpoint_0 (%d B) + layouts of %d shipped checkpoints"
          % (os.path.getsize(src), len(out)), flush=True)


def main():
    os.makedirs(OUT, exist_ok=True)
    only = set(sys.argv[1:])
    if not only or "checkpoints" in only:
        make_checkpoint_fixture()
        if only == {"checkpoints"}:
            return
    ref = refshim.load("float64")
    if not only or "dynamics_ref" in only:
        make_dynamics_fixture(ref)
    if not only or "train_step_ref" in only:
        make_train_step_fixture(ref)
    for name, conf, ckpt, kind, n in CASES:
        if only and name not in only:
            continue
        make_model_fixture(ref, name, conf, ckpt, kind, n)


if __name__ == "__main__":
    main()
