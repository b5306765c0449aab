"""Generate tests/golden/*.npz from the reference's shipped checkpoints + configs.

Runs ONLY in the build container (needs /root/reference); the GPU box uses the committed
fixtures.  JAX / Flax / flax_rbf are not installable here, so the vectors are
produced by the NumPy oracle (oracle/irbfn_oracle.py, float64) and cross-checked, before
anything is written, against a second, independent restatement: a literal torch-float64
transcription of the reference forward (sqrt / divide / tanh form, model.py:70-95,169-198 and
flax_rbf.py:275-283) differentiated by torch autograd.  A fixture is only written when both
agree to 1e-9 (outputs) / 1e-7 (gradients).

Fixture contents (per checkpoint):
  cfg_json              module attributes (num_regions, bounds, delta, ranges, basis name ...)
  p/<leaf path>         parameters cast to float32 (what the reference's x64-disabled inference
                        path evaluates)
  x                     (N,D) float32 queries: uniform in the config box, default_rng(0)
  y64                   (N,O) float64 outputs of the oracle on float32 params and x
  g/<leaf>, g/x         float64 gradients of L = mean(y**2) w.r.t. every leaf and x
  [logits64]            ClusterWCRBFNet logits

Usage:  python tests/golden/make_golden.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import irbfn_oracle as orc  # noqa: E402

REF = "/root/reference/scripts"
OUT = os.path.dirname(os.path.abspath(__file__))
N = 1024

# (fixture name, config, checkpoint, module kind)
CASES = [
    ("c1_cartesian_1x1000", "dnmpc_1regions_newnewdata_1stepst_l2_newarch_ksint", "checkpoint_9999", "wcrbf"),
    ("c1b_nmpc_4_regions", "nmpc_4_regions", "checkpoint_0", "wcrbf"),
    ("c2_frenet_12x100", "dnmpc_12regions_frenet_l1_bigdata", "checkpoint_9500", "wcrbf"),
    ("c2b_frenet_1x1000_o10", "dnmpc_1regions_frenet_l1_bigdata_5stepint_highk", "checkpoint_3900", "wcrbf"),
    ("nmpc_1_region_iq", "nmpc_1_region", "checkpoint_0", "wcrbf"),
    ("dnmpc_4regions_iq", "dnmpc_4regions_correctdata", "checkpoint_999", "wcrbf"),
    ("dnmpc_128regions_narrow", "dnmpc_128regions", "checkpoint_600", "wcrbf"),
    ("dnmpc_108regions_wide", "dnmpc_108regions_frenet_l1_bigdata", "checkpoint_6000", "wcrbf"),
    ("deeper_1x100", "dnmpc_1regions_frenet_l1_bigdata_5stepint_deeper", "checkpoint_9999", "deeper"),
    ("cluster_500x10", "dnmpc_500_clusters", "checkpoint_0", "cluster"),
]

CFG_KEYS = ("in_features", "out_features", "num_kernels", "num_regions", "basis_func",
            "lower_bounds", "upper_bounds", "dimension_ranges", "activation_idx", "delta")


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


# ---- literal torch restatement (independent of the oracle's analytic VJP) ----------------
def torch_forward(x, p, cfg, kind):
    """Transcribes the reference arithmetic op for op in torch float64."""
    basis = cfg["basis_func"]
    c, ls = p["rbf_list/centers"], p["rbf_list/log_sigs"]
    diff = x[:, None, None, :] - c[None]
    d = (diff ** 2).sum(-1) ** 0.5 / torch.exp(ls)[None]          # flax_rbf.py:280
    phi = torch.exp(-1 * d ** 2) if basis == "gaussian" else 1 / (1 + d ** 2)
    logits = None
    if kind == "cluster":
        logits = x @ p["cluster/kernel"] + p["cluster/bias"]        # model.py:402
        gamma = torch.softmax(logits, dim=-1)                       # model.py:403
    else:
        S = len(cfg["lower_bounds"])
        gs = []
        for j in range(S):                                          # model.py:74-86
            lb = torch.tensor(cfg["lower_bounds"][j], dtype=x.dtype)[None]
            ub = torch.tensor(cfg["upper_bounds"][j], dtype=x.dtype)[None]
            dl = float(cfg["delta"][j])
            gs.append(((torch.tanh(dl * (x[:, j:j + 1] - lb)) + 1) / 2)
                      * ((torch.tanh(dl * (ub - x[:, j:j + 1])) + 1) / 2))
        cols = []
        for i in range(cfg["num_regions"]):                         # model.py:88-93
            cur = gs[0][:, cfg["dimension_ranges"][i][0]]
            for j in range(1, S):
                cur = cur * gs[j][:, cfg["dimension_ranges"][i][j]]
            cols.append(cur)
        gamma = torch.stack(cols, dim=1)
    h = (gamma[:, :, None] * phi).sum(1)                            # model.py:187-193
    if kind == "deeper":                                            # model.py:285-287
        h = torch.relu(h @ p["linear_pre1/kernel"] + p["linear_pre1/bias"])
        h = torch.relu(h @ p["linear_pre2/kernel"] + p["linear_pre2/bias"])
    y = h @ p["linear/kernel"] + p["linear/bias"]                   # model.py:196
    return y, logits


def oracle_eval(x, params, cfg, kind):
    """float64 oracle outputs and gradients of L = mean(y^2) (analytic VJP)."""
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
        y = orc.deeper_forward(x, params, cfg)
        return y, None, None  # gradients come from the torch restatement only
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


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, conf, ckpt, kind in CASES:
        cfg_all = yaml.safe_load(open(os.path.join(REF, "configs", conf + ".yaml")))
        cfg = {k: cfg_all[k] for k in CFG_KEYS if k in cfg_all}
        cfg["kind"] = kind
        cfg["source_config"] = "scripts/configs/%s.yaml" % conf
        cfg["source_checkpoint"] = "scripts/ckpts/%s/%s" % (conf, ckpt)
        tree = orc.load_flax_checkpoint(os.path.join(REF, "ckpts", conf, ckpt))["params"]["params"]
        flat32 = {k: v.astype(np.float32) for k, v in flatten(tree).items()}
        params = unflatten({k: v.astype(np.float64) for k, v in flat32.items()})
        n = 256 if kind == "cluster" else N   # (n, 500) logits: keep the fixture small
        x32 = orc.uniform_queries(cfg, 4096, seed=0)[:n].astype(np.float32)
        x = x32.astype(np.float64)

        y, grads, logits = oracle_eval(x, params, cfg, kind)

        # independent check: literal torch restatement + autograd
        tp = {k: torch.tensor(v.astype(np.float64), requires_grad=True) for k, v in flat32.items()}
        tx = torch.tensor(x, requires_grad=True)
        ty, tlogits = torch_forward(tx, tp, cfg, kind)
        loss = (ty ** 2).mean()
        leaves = list(tp.keys())
        tg = torch.autograd.grad(loss, [tp[k] for k in leaves] + [tx])
        tgrads = {k: g.numpy() for k, g in zip(leaves + ["x"], tg)}
        scale = np.abs(y).max()
        err_y = np.abs(ty.detach().numpy() - y).max() / scale
        assert err_y < 1e-9, (name, err_y)
        if grads is None:
            grads = tgrads
        else:
            for k, g in grads.items():
                ref = tgrads[k]
                if not np.isfinite(ref).all():
                    raise AssertionError("%s: torch gradient %s not finite (x == centre?)" % (name, k))
                e = np.linalg.norm(g - ref) / max(np.linalg.norm(ref), 1e-300)
                assert e < 1e-7, (name, k, e)
        if logits is not None:
            assert np.abs(tlogits.detach().numpy() - logits).max() < 1e-9

        out = {"cfg_json": np.array(json.dumps(cfg)), "x": x32, "y64": y}
        for k, v in flat32.items():
            out["p/" + k] = v
        for k, v in grads.items():
            out["g/" + k] = v
        if logits is not None:
            out["logits64"] = logits
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **out)
        print("%-28s kind=%-7s y err vs torch %.1e  leaves %s  %.0f KB" % (
            name, kind, err_y, sorted(grads), os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
