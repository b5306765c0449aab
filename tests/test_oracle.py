"""CPU tests of the oracle (oracle/irbfn_oracle.py) against the committed golden fixtures.

The fixtures hold values produced by the reference's own source executed under oracle/refshim
(tests/golden/make_golden.py; tests/test_oracle_vs_ref.py is the pin itself).  Kept here: the survey's
independent spot value, the fp32-vs-fp64 margin, and a literal torch restatement as a third opinion.
"""
import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden, rel_l2, rel_max
from oracle import irbfn_oracle as orc


def _ids(paths):
    import os
    return [os.path.basename(p)[:-4] for p in paths]


@pytest.mark.parametrize("path", golden_paths(), ids=_ids(golden_paths()))
def test_oracle_reproduces_golden_outputs(path):
    cfg, params, arr = load_golden(path)
    x = arr["x"].astype(np.float64)
    if cfg["kind"] == "wcrbf":
        y = orc.wcrbf_forward(x, params, cfg)
    elif cfg["kind"] == "deeper":
        y = orc.deeper_forward(x, params, cfg)
    else:
        y, logits = orc.cluster_forward(x, params, cfg)
        assert np.abs(logits - arr["logits64"]).max() < 1e-10
    assert y.shape == arr["y64"].shape
    assert rel_max(y, arr["y64"]) < 1e-12


@pytest.mark.parametrize("path", [p for p in golden_paths() if "deeper" not in p and "cluster" not in p],
                         ids=_ids([p for p in golden_paths() if "deeper" not in p and "cluster" not in p]))
def test_oracle_reproduces_golden_gradients(path):
    cfg, params, arr = load_golden(path)
    x = arr["x"].astype(np.float64)
    y = orc.wcrbf_forward(x, params, cfg)
    g = orc.wcrbf_vjp(x, params, cfg, 2 * y / y.size)
    assert rel_l2(g["centers"], arr["g/rbf_list/centers"]) < 1e-10
    assert rel_l2(g["log_sigs"], arr["g/rbf_list/log_sigs"]) < 1e-10
    assert rel_l2(g["kernel"], arr["g/linear/kernel"]) < 1e-10
    assert rel_l2(g["bias"], arr["g/linear/bias"]) < 1e-10
    assert rel_l2(g["x"], arr["g/x"]) < 1e-10


def test_survey_spot_value():
    """nmpc_4_regions/checkpoint_0, first query of default_rng(0): computed in the
    survey session by a separate torch restatement on float64 params (the fixture holds the
    float32-cast params, hence 1e-6)."""
    import os
    import os
    from conftest import GOLDEN_DIR
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c1b_nmpc_4_regions.npz"))
    y = orc.wcrbf_forward(arr["x"][:1].astype(np.float64), params, cfg)
    spot = np.array([3.06386078, 16.57250513, 0.88298678, -30.269345])
    assert np.abs(y[0, :4] - spot).max() / np.abs(spot).max() < 1e-6


def test_fp32_oracle_within_tolerance_of_fp64():
    """F10: fp32 arithmetic itself is 2e-6..6e-6 from fp64 on shipped params."""
    for path in golden_paths():
        cfg, params, arr = load_golden(path)
        if cfg["kind"] != "wcrbf":
            continue
        y32 = orc.wcrbf_forward(arr["x"], params, cfg, dtype=np.float32)
        assert y32.dtype == np.float32
        assert rel_max(y32, arr["y64"]) < 2e-5


# ---- literal restatement of the reference forward (sqrt / divide / tanh / python loops over
# ---- regions), differentiated by autograd: independent of the oracle's analytic VJP ---------
def _literal_forward(x, c, ls, W, b, lbs, ubs, delta, ranges, basis):
    diff = x[:, None, None, :] - c[None]
    d = (diff ** 2).sum(-1) ** 0.5 / torch.exp(ls)[None]          # flax_rbf.py:280
    phi = torch.exp(-1 * d ** 2) if basis == "gaussian" else 1 / (1 + d ** 2)
    gs = []
    for j in range(len(lbs)):                                      # model.py:74-86
        lb = torch.tensor(lbs[j], dtype=x.dtype)[None]
        ub = torch.tensor(ubs[j], dtype=x.dtype)[None]
        gs.append(((torch.tanh(delta[j] * (x[:, j:j + 1] - lb)) + 1) / 2)
                  * ((torch.tanh(delta[j] * (ub - x[:, j:j + 1])) + 1) / 2))
    cols = []
    for i in range(len(ranges)):                                   # model.py:88-93
        cur = gs[0][:, ranges[i][0]]
        for j in range(1, len(lbs)):
            cur = cur * gs[j][:, ranges[i][j]]
        cols.append(cur)
    gamma = torch.stack(cols, 1)
    h = (gamma[:, :, None] * phi).sum(1)                           # model.py:187-193
    return h @ W + b                                               # model.py:196


def _random_case(seed, B=37, R=6, K=9, D=5, O=3):
    rng = np.random.default_rng(seed)
    lbs = [[-1.0, 0.0, 0.5], [-2.0, 0.1]]
    ubs = [[0.0, 0.6, 2.0], [0.2, 2.0]]
    ranges = [[i, j] for i in range(3) for j in range(2)]
    cfg = dict(num_regions=R, lower_bounds=lbs, upper_bounds=ubs, delta=[3.0, 7.0],
               dimension_ranges=ranges, in_features=D, out_features=O, num_kernels=K)
    p = {"rbf_list": {"centers": rng.normal(size=(R, K, D)), "log_sigs": rng.uniform(-0.5, 1, (R, K))},
         "linear": {"kernel": rng.normal(size=(K, O)), "bias": rng.normal(size=O)}}
    x = rng.uniform(-1.5, 1.5, (B, D))
    return cfg, p, x


@pytest.mark.parametrize("basis", ["gaussian", "inverse_quadratic"])
def test_oracle_matches_literal_restatement_and_autograd(basis):
    cfg, p, x = _random_case(3)
    cfg["basis_func"] = basis
    y = orc.wcrbf_forward(x, p, cfg)
    dy = np.random.default_rng(5).normal(size=y.shape)
    g = orc.wcrbf_vjp(x, p, cfg, dy)

    t = lambda a: torch.tensor(np.asarray(a, np.float64), requires_grad=True)
    tx, tc, tls = t(x), t(p["rbf_list"]["centers"]), t(p["rbf_list"]["log_sigs"])
    tW, tb = t(p["linear"]["kernel"]), t(p["linear"]["bias"])
    ty = _literal_forward(tx, tc, tls, tW, tb, cfg["lower_bounds"], cfg["upper_bounds"],
                          cfg["delta"], cfg["dimension_ranges"], basis)
    assert rel_max(y, ty.detach().numpy()) < 1e-12
    gx, gc, gls, gW, gb = torch.autograd.grad((ty * torch.tensor(dy)).sum(), [tx, tc, tls, tW, tb])
    assert rel_l2(g["x"], gx.numpy()) < 1e-10
    assert rel_l2(g["centers"], gc.numpy()) < 1e-10
    assert rel_l2(g["log_sigs"], gls.numpy()) < 1e-10
    assert rel_l2(g["kernel"], gW.numpy()) < 1e-10
    assert rel_l2(g["bias"], gb.numpy()) < 1e-10


def test_gate_is_not_normalised_and_overlaps_sum_above_one():
    """F4: overlapping regions (nmpc_4_regions) make the gates sum to more than 1."""
    import os
    import os
    from conftest import GOLDEN_DIR
    cfg, _, arr = load_golden(os.path.join(GOLDEN_DIR, "c1b_nmpc_4_regions.npz"))
    gam = orc.region_activation(arr["x"], cfg["num_regions"], cfg["lower_bounds"],
                                cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"])
    assert gam.min() >= 0 and gam.sum(1).max() > 1.5


def test_fewer_ranges_than_regions_leaves_zero_gate():
    cfg, p, x = _random_case(1)
    gam = orc.region_activation(x, 8, cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                cfg["dimension_ranges"])
    assert gam.shape == (x.shape[0], 8) and np.all(gam[:, 6:] == 0)


def test_empty_batch():
    cfg, p, x = _random_case(2)
    cfg["basis_func"] = "gaussian"
    y = orc.wcrbf_forward(x[:0], p, cfg)
    assert y.shape == (0, 3)


def test_query_on_centre_has_finite_analytic_gradient():
    """F8: the reference's sqrt formulation yields NaN at x == centre; the oracle's VJP is
    written on d^2 and returns the analytic limit."""
    cfg, p, x = _random_case(4)
    cfg["basis_func"] = "gaussian"
    x[0] = p["rbf_list"]["centers"][2, 3]
    y = orc.wcrbf_forward(x, p, cfg)
    g = orc.wcrbf_vjp(x, p, cfg, np.ones_like(y))
    for k in ("x", "centers", "log_sigs", "kernel", "bias"):
        assert np.isfinite(g[k]).all()
    # the literal formulation does produce NaN there
    t = lambda a: torch.tensor(np.asarray(a, np.float64), requires_grad=True)
    tx, tc = t(x), t(p["rbf_list"]["centers"])
    ty = _literal_forward(tx, tc, t(p["rbf_list"]["log_sigs"]), t(p["linear"]["kernel"]),
                          t(p["linear"]["bias"]), cfg["lower_bounds"], cfg["upper_bounds"],
                          cfg["delta"], cfg["dimension_ranges"], "gaussian")
    (gx,) = torch.autograd.grad(ty.sum(), [tx])
    assert not torch.isfinite(gx).all()


def test_softmax_gate_rows_sum_to_one():
    rng = np.random.default_rng(0)
    gate, logits = orc.softmax_gate(rng.normal(size=(11, 8)), rng.normal(size=(8, 50)), rng.normal(size=50))
    assert np.allclose(gate.sum(1), 1) and logits.shape == (11, 50)


def test_unsupported_basis_raises():
    cfg, p, x = _random_case(2)
    with pytest.raises(ValueError):
        orc.rbf_activations(x, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], 7)


def test_jacobian_oracle_matches_finite_differences():
    """wcrbf_jacobian (jacrev by one-hot VJPs) against central differences of the float64 forward."""
    import os
    from conftest import GOLDEN_DIR
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    x = arr["x"][:16].astype(np.float64)
    J = orc.wcrbf_jacobian(x, params, cfg)
    h = 1e-6
    for j in range(x.shape[1]):
        e = np.zeros_like(x)
        e[:, j] = h
        fd = (orc.wcrbf_forward(x + e, params, cfg) - orc.wcrbf_forward(x - e, params, cfg)) / (2 * h)
        assert np.abs(fd - J[:, :, j]).max() <= 1e-6 * max(1.0, np.abs(J).max())
