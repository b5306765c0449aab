"""GPU parity of the VJP (irbfn_backward through the C ABI) against the golden gradients and
the float64 oracle.

Tolerance (BASELINE.json north_star): per leaf ||g - g_ref||_2 / ||g_ref||_2 <= 1e-4.
"""
import os

import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden, rel_l2, rel_max
from oracle import irbfn_oracle as orc
from test_forward_gpu import SHAPES, _gd, _ops, _synthetic, _t

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-4
WCRBF = [p for p in golden_paths() if "deeper" not in p and "cluster" not in p]


@pytest.mark.parametrize("path", WCRBF, ids=[os.path.basename(p)[:-4] for p in WCRBF])
def test_backward_matches_golden(path, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(path)
    dev = cuda_device
    rb, lin = params["rbf_list"], params["linear"]
    y64 = arr["y64"]
    dy = _t(2 * y64 / y64.size, dev)
    g = ops.backward(_t(arr["x"], dev), _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev),
                     dy, gd=_gd(cfg, dev), basis=orc.BASIS_IDS[cfg["basis_func"]],
                     grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    assert rel_l2(g["centers"].cpu().numpy(), arr["g/rbf_list/centers"]) < GRAD_TOL
    assert rel_l2(g["log_sigs"].cpu().numpy(), arr["g/rbf_list/log_sigs"]) < GRAD_TOL
    assert rel_l2(g["kernel"].cpu().numpy(), arr["g/linear/kernel"]) < GRAD_TOL
    assert rel_l2(g["bias"].cpu().numpy(), arr["g/linear/bias"]) < GRAD_TOL
    assert rel_l2(g["x"].cpu().numpy(), arr["g/x"]) < GRAD_TOL


def test_cluster_backward_matches_golden(cuda_device):
    """ClusterWCRBFNet: the kernel returns dgate; the softmax + Dense(D->R) chain rule is host side."""
    ops = _ops()
    cfg, params, arr = load_golden([p for p in golden_paths() if "cluster" in p][0])
    dev = cuda_device
    rb, lin, cl = params["rbf_list"], params["linear"], params["cluster"]
    x = _t(arr["x"], dev)
    Wc, bc = _t(cl["kernel"], dev), _t(cl["bias"], dev)
    gate, _ = ops.softmax_gate(x, Wc, bc)
    y64 = arr["y64"]
    dy = _t(2 * y64 / y64.size, dev)
    g = ops.backward(x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev), dy, gate=gate,
                     basis=orc.BASIS_IDS[cfg["basis_func"]],
                     grad_mask=ops.GRAD_PARAMS | ops.GRAD_X | ops.GRAD_GATE)
    dlogits = gate * (g["gate"] - (g["gate"] * gate).sum(-1, keepdim=True))
    assert rel_l2((x.T @ dlogits).cpu().numpy(), arr["g/cluster/kernel"]) < GRAD_TOL
    assert rel_l2(dlogits.sum(0).cpu().numpy(), arr["g/cluster/bias"]) < GRAD_TOL
    assert rel_l2((g["x"] + dlogits @ Wc.T).cpu().numpy(), arr["g/x"]) < GRAD_TOL
    assert rel_l2(g["centers"].cpu().numpy(), arr["g/rbf_list/centers"]) < GRAD_TOL
    assert rel_l2(g["log_sigs"].cpu().numpy(), arr["g/rbf_list/log_sigs"]) < GRAD_TOL
    assert rel_l2(g["kernel"].cpu().numpy(), arr["g/linear/kernel"]) < GRAD_TOL


def test_deeper_backward_matches_golden(cuda_device):
    """DeeperWCRBFNet: hidden activations out, cotangent dhidden in; the MLP tail is differentiated
    by torch autograd on the device."""
    ops = _ops()
    cfg, params, arr = load_golden([p for p in golden_paths() if "deeper" in p][0])
    dev = cuda_device
    rb = params["rbf_list"]
    x = _t(arr["x"], dev)
    c, ls = _t(rb["centers"], dev), _t(rb["log_sigs"], dev)
    gd = _gd(cfg, dev)
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    h = ops.forward(x, c, ls, gd=gd, basis=basis, want_y=False, want_hidden=True).requires_grad_(True)
    p = {k: {kk: _t(vv, dev).requires_grad_(True) for kk, vv in v.items()} for k, v in params.items() if k != "rbf_list"}
    t = torch.relu(h @ p["linear_pre1"]["kernel"] + p["linear_pre1"]["bias"])
    t = torch.relu(t @ p["linear_pre2"]["kernel"] + p["linear_pre2"]["bias"])
    y = t @ p["linear"]["kernel"] + p["linear"]["bias"]
    (y ** 2).mean().backward()
    g = ops.backward(x, c, ls, None, None, dhidden=h.grad.contiguous(), gd=gd, basis=basis,
                     grad_mask=ops.GRAD_CENTERS | ops.GRAD_LOG_SIGS | ops.GRAD_X)
    assert rel_l2(g["centers"].cpu().numpy(), arr["g/rbf_list/centers"]) < GRAD_TOL
    assert rel_l2(g["log_sigs"].cpu().numpy(), arr["g/rbf_list/log_sigs"]) < GRAD_TOL
    assert rel_l2(g["x"].cpu().numpy(), arr["g/x"]) < GRAD_TOL
    assert rel_l2(p["linear_pre1"]["kernel"].grad.cpu().numpy(), arr["g/linear_pre1/kernel"]) < GRAD_TOL


@pytest.mark.parametrize("basis", [0, 1])
@pytest.mark.parametrize("shape", SHAPES, ids=["B%d_R%d_K%d_D%d_O%d" % s for s in SHAPES])
def test_backward_ragged_shapes_match_oracle(shape, basis, cuda_device):
    ops = _ops()
    B, R, K, D, O = shape
    cfg, p, x = _synthetic(13, B, R, K, D, O)
    cfg["basis_func"] = "gaussian" if basis == 0 else "inverse_quadratic"
    dev = cuda_device
    dy = np.random.default_rng(3).normal(size=(B, O)).astype(np.float32)
    params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
              "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    ref = orc.wcrbf_vjp(x.astype(np.float64), params, cfg, dy.astype(np.float64))
    g = ops.backward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev),
                     gd=_gd(cfg, dev), basis=basis, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    for name in ("centers", "log_sigs", "kernel", "bias", "x"):
        assert rel_l2(g[name].cpu().numpy(), ref[name]) < GRAD_TOL, name
    # external-gate form: dgate is returned, dx carries the RBF-layer term only
    gamma = orc.region_activation(x.astype(np.float64), R, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    ref2 = orc.backward(x.astype(np.float64), p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma,
                        dy.astype(np.float64), basis)
    g2 = ops.backward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev),
                      gate=_t(gamma, dev), basis=basis, grad_mask=ops.GRAD_X | ops.GRAD_GATE)
    assert rel_l2(g2["x"].cpu().numpy(), ref2["x"]) < GRAD_TOL
    assert rel_l2(g2["gate"].cpu().numpy(), ref2["gamma"]) < GRAD_TOL


def test_backward_parameters_larger_than_shared_memory(cuda_device):
    ops = _ops()
    cfg, p, x = _synthetic(5, 160, 64, 4096, 6, 16, S=3, nb=(4, 4, 4))
    cfg["basis_func"] = "gaussian"
    dev = cuda_device
    dy = np.random.default_rng(3).normal(size=(160, 16)).astype(np.float32)
    params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
              "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    ref = orc.wcrbf_vjp(x.astype(np.float64), params, cfg, dy.astype(np.float64), chunk=4)
    g = ops.backward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev),
                     gd=_gd(cfg, dev), basis=0, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    for name in ("centers", "log_sigs", "kernel", "bias", "x"):
        assert rel_l2(g[name].cpu().numpy(), ref[name]) < GRAD_TOL, name


WIDE_SHAPES = [
    # B, R, K, D, O : several regions under a wide readout -> irbfn_bwd_pw_kernel (lanes = regions).
    # group sizes 2..32, ragged last region group, odd K, O below / at the window, ragged query tiles
    (777, 40, 11, 6, 16), (4099, 17, 7, 8, 12), (333, 64, 20, 6, 16), (1500, 27, 50, 8, 10), (513, 9, 101, 7, 6),
    (2000, 5, 64, 5, 8), (90, 33, 3, 4, 5), (17, 2, 1, 8, 16),
]


@pytest.mark.parametrize("shape", WIDE_SHAPES, ids=["B%d_R%d_K%d_D%d_O%d" % s for s in WIDE_SHAPES])
def test_wide_readout_parameter_gradients(shape, cuda_device):
    ops = _ops()
    B, R, K, D, O = shape
    cfg, p, x = _synthetic(29, B, R, K, D, O, S=min(3, D), nb=(4, 3, 3))
    dev = cuda_device
    dy = np.random.default_rng(4).normal(size=(B, O)).astype(np.float32)
    params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
              "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    for basis in (0, 1):
        cfg["basis_func"] = "gaussian" if basis == 0 else "inverse_quadratic"
        ref = orc.wcrbf_vjp(x.astype(np.float64), params, cfg, dy.astype(np.float64), chunk=64)
        args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev))
        g = ops.backward(*args, gd=_gd(cfg, dev), basis=basis, grad_mask=ops.GRAD_PARAMS)
        for name in ("centers", "log_sigs", "kernel", "bias"):
            assert rel_l2(g[name].cpu().numpy(), ref[name]) < GRAD_TOL, (name, basis)
        g2 = ops.backward(*args, gd=_gd(cfg, dev), basis=basis, grad_mask=ops.GRAD_PARAMS)
        for name in ("centers", "log_sigs", "kernel", "bias"):
            assert torch.equal(g[name], g2[name]), "not deterministic: %s" % name


def test_query_on_centre_gives_finite_analytic_limit(cuda_device):
    """F8: the reference's sqrt form gives NaN here; the kernels work on d^2."""
    ops = _ops()
    cfg, p, x = _synthetic(4, 64, 3, 9, 5, 3)
    cfg["basis_func"] = "gaussian"
    x[0] = p["centers"][1, 4]
    x[7] = p["centers"][2, 0]
    dev = cuda_device
    dy = np.ones((64, 3), np.float32)
    params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
              "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    ref = orc.wcrbf_vjp(x.astype(np.float64), params, cfg, dy.astype(np.float64))
    g = ops.backward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev),
                     gd=_gd(cfg, dev), basis=0, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    for name in ("centers", "log_sigs", "kernel", "bias", "x"):
        assert torch.isfinite(g[name]).all()
        assert rel_l2(g[name].cpu().numpy(), ref[name]) < GRAD_TOL, name


def test_training_size_gradients_are_additive_and_deterministic(cuda_device):
    """Size-independent properties at the C3 size (262,144 samples, C2 shape): the gradient of
    the whole batch equals the sum of the gradients of its two halves; repeated launches are
    bit-identical; a subsample agrees with the oracle."""
    ops = _ops()
    cfg, params, arr = load_golden([p for p in golden_paths() if "c2_frenet_12x100" in p][0])
    dev = cuda_device
    rb, lin = params["rbf_list"], params["linear"]
    B = 262144
    x = _t(orc.uniform_queries(cfg, B, seed=2), dev)
    c, ls, W = _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev)
    dy = torch.randn((B, 2), generator=torch.Generator().manual_seed(7)).to(dev) / B
    gd = _gd(cfg, dev)
    full = ops.backward(x, c, ls, W, dy, gd=gd, grad_mask=ops.GRAD_PARAMS)
    again = ops.backward(x, c, ls, W, dy, gd=gd, grad_mask=ops.GRAD_PARAMS)
    h = B // 2
    lo = ops.backward(x[:h].contiguous(), c, ls, W, dy[:h].contiguous(), gd=gd, grad_mask=ops.GRAD_PARAMS)
    hi = ops.backward(x[h:].contiguous(), c, ls, W, dy[h:].contiguous(), gd=gd, grad_mask=ops.GRAD_PARAMS)
    for name in ("centers", "log_sigs", "kernel", "bias"):
        assert torch.equal(full[name], again[name]), name
        assert rel_l2((lo[name] + hi[name]).cpu().numpy(), full[name].cpu().numpy()) < GRAD_TOL, name
    n = 4096
    ref = orc.wcrbf_vjp(x[:n].cpu().numpy().astype(np.float64), params, cfg, dy[:n].cpu().numpy().astype(np.float64))
    sub = ops.backward(x[:n].contiguous(), c, ls, W, dy[:n].contiguous(), gd=gd, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    for name in ("centers", "log_sigs", "kernel", "bias", "x"):
        assert rel_l2(sub[name].cpu().numpy(), ref[name]) < GRAD_TOL, name


def test_backward_errors(cuda_device):
    ops = _ops()
    from irbfn_b200._lib import IrbfnError
    cfg, p, x = _synthetic(2, 8, 2, 3, 4, 2)
    dev = cuda_device
    args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev))
    dy = torch.ones((8, 2), device=dev)
    with pytest.raises(IrbfnError) as e:
        ops.backward(*args, dy, gd=_gd(cfg, dev), grad_mask=0)
    assert e.value.code == -9
    with pytest.raises(IrbfnError) as e:   # dgate only exists for an external gate
        ops.backward(*args, dy, gd=_gd(cfg, dev), grad_mask=ops.GRAD_GATE)
    assert e.value.code == -9


def test_graph_captured_train_step_equals_eager_steps(cuda_device):
    """GraphedTrainStep (forward, loss, VJP, clip + Adam in one CUDA graph) walks the same trajectory as
    eager train_step calls, and capturing it does not advance the parameters."""
    import os
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.train import GraphedTrainStep, TrainState, train_step
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    mk = lambda: WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                          cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                          cfg["delta"])
    x = torch.tensor(arr["x"][:1000], device=cuda_device)
    y = torch.tensor((arr["y64"][:1000] + 0.1).astype(np.float32), device=cuda_device)
    for loss in ("l2", "frenet_oneint"):
        eager = TrainState(mk(), device_params({"params": params}, cuda_device), lr=1e-3, device=cuda_device)
        graphed = TrainState(mk(), device_params({"params": params}, cuda_device), lr=1e-3, device=cuda_device)
        step = GraphedTrainStep(graphed, 1000, loss=loss).load(x, y)
        assert torch.equal(graphed.flat, eager.flat) and graphed.step == 0
        for _ in range(4):
            le = train_step(eager, x, y, loss=loss)
            lg = step()
            assert torch.allclose(le, lg, rtol=1e-6, atol=0)
        assert graphed.step == 4
        assert torch.allclose(graphed.flat, eager.flat, rtol=1e-5, atol=1e-7)
