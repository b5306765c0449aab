"""Gate-sparse, region-binned forward (irbfn_forward_sparse) against the DENSE float64 oracle.

The sparse path may drop a local network for a query only when its gate is below tol_abs / (R * max_o sum_k |W_ko|),
so every output stays within tol_abs of the dense result; with the tolerances used here that keeps the norm-wise
output error <= 1e-5 exactly like the dense kernel."""
import os

import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden, rel_max
from oracle import irbfn_oracle as orc
from test_forward_gpu import SHAPES, _gd, _ops, _synthetic, _t

pytestmark = pytest.mark.gpu
WCRBF = [p for p in golden_paths() if "deeper" not in p and "cluster" not in p]


@pytest.mark.parametrize("path", WCRBF, ids=[os.path.basename(p)[:-4] for p in WCRBF])
def test_sparse_forward_matches_dense_oracle_on_shipped_checkpoints(path, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(path)
    rb, lin = params["rbf_list"], params["linear"]
    dev = cuda_device
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    args = (_t(arr["x"], dev), _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev), _t(lin["bias"], dev))
    scale = np.abs(arr["y64"]).max()
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    y = ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=basis, tol_abs=1e-6 * scale, counters=cnt)
    assert rel_max(y.cpu().numpy(), arr["y64"]) < 1e-5
    skipped, done = (int(v) for v in cnt.cpu())
    assert done > 0 and skipped >= 0
    # tol_abs = 0 drops only gates that are exactly zero: the dense kernel's result to fp32 rounding
    y0 = ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=basis, tol_abs=0.0)
    yd = ops.forward(*args, gd=_gd(cfg, dev), basis=basis)
    assert rel_max(y0.cpu().numpy(), arr["y64"]) < 1e-5
    assert rel_max(y0.cpu().numpy(), yd.cpu().numpy()) < 1e-5      # (summation orders differ: split-K vs one pass)


@pytest.mark.parametrize("shape", [(1, 1, 1, 1, 1), (1023, 3, 7, 4, 5), (5000, 4, 33, 6, 9), (4097, 12, 100, 8, 2),
                                   (300, 70, 9, 5, 3), (2000, 130, 6, 7, 18)],
                         ids=lambda s: "B%d_R%d_K%d_D%d_O%d" % s)
def test_sparse_forward_ragged_shapes(shape, cuda_device):
    ops = _ops()
    B, R, K, D, O = shape
    cfg, p, x = _synthetic(41, B, R, K, D, O, S=min(3, D), nb=(4, 3, 2))
    cfg["delta"] = [40.0] * len(cfg["delta"])          # sharp gates: most regions are negligible for a query
    dev = cuda_device
    for basis, name in ((0, "gaussian"), (1, "inverse_quadratic")):
        cfg["basis_func"] = name
        params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
                  "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
        ref = orc.wcrbf_forward(x.astype(np.float64), params, cfg)
        args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(p["bias"], dev))
        tol = 1e-6 * np.abs(ref).max()
        y = ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=basis, tol_abs=tol)
        assert np.abs(y.cpu().numpy() - ref).max() <= 1e-5 * np.abs(ref).max()
        y2 = ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=basis, tol_abs=tol)
        assert torch.equal(y, y2)                      # the binning order cannot change a query's result


def test_sparse_forward_on_the_stress_shape_skips_most_regions(cuda_device):
    """C5-like: 64 non-overlapping regions (delta = 100), parameters larger than shared memory (chunked, compact
    staging of the active regions only); random queries need the binning to skip uniformly."""
    ops = _ops()
    dev = cuda_device
    R, K, D, O, B = 64, 512, 6, 16, 20000
    rng = np.random.default_rng(7)
    edges = np.linspace(-1.0, 1.0, 5)
    cfg = dict(num_regions=R, lower_bounds=[list(edges[:-1])] * 3, upper_bounds=[list(edges[1:])] * 3, delta=[100.0] * 3,
               dimension_ranges=[[i, j, k] for i in range(4) for j in range(4) for k in range(4)],
               in_features=D, out_features=O, num_kernels=K, basis_func="gaussian", activation_idx=[0, 1, 2])
    p = {"centers": rng.uniform(-1, 1, (R, K, D)).astype(np.float32), "log_sigs": rng.uniform(0, 2, (R, K)).astype(np.float32),
         "kernel": rng.standard_normal((K, O)).astype(np.float32), "bias": np.zeros(O, np.float32)}
    x = rng.uniform(-1, 1, (B, D)).astype(np.float32)
    params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]}, "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    ref = orc.wcrbf_forward(x.astype(np.float64), params, cfg, chunk=256)
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    y = ops.forward_sparse(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(p["bias"], dev),
                           gd=_gd(cfg, dev), basis=0, tol_abs=1e-6 * np.abs(ref).max(), counters=cnt)
    assert np.abs(y.cpu().numpy() - ref).max() <= 1e-5 * np.abs(ref).max()
    skipped, done = (int(v) for v in cnt.cpu())
    assert skipped > 4 * done, (skipped, done)         # > 80 % of the (warp, region) evaluations skipped


def test_sparse_forward_errors(cuda_device):
    from irbfn_b200._lib import IrbfnError
    ops = _ops()
    cfg, p, x = _synthetic(3, 64, 2, 5, 4, 2)
    dev = cuda_device
    args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(p["bias"], dev))
    with pytest.raises(IrbfnError):
        ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=0, tol_abs=-1.0)
    with pytest.raises(IrbfnError):
        ops.forward_sparse(*args, gd=_gd(cfg, dev), basis=7, tol_abs=0.0)


# ---------------------------------------------------------------------------------------------------
# Gate-sparse parameter gradients (irbfn_backward_sparse): the terms of (query, region) pairs with gamma <= gate_tol are
# left out.  Checked against the reference-executed gradients of the shipped checkpoints and against the DENSE float64
# oracle at the gradient tolerance every backward test uses (rel-L2 <= 1e-4 per leaf).
# ---------------------------------------------------------------------------------------------------
GRAD_TOL = 1e-4
GATE_TOL = 2.0 ** -24
LEAVES = (("centers", "g/rbf_list/centers"), ("log_sigs", "g/rbf_list/log_sigs"), ("kernel", "g/linear/kernel"),
          ("bias", "g/linear/bias"))


def _rel_l2(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


@pytest.mark.parametrize("path", WCRBF, ids=[os.path.basename(p)[:-4] for p in WCRBF])
def test_sparse_backward_matches_reference_gradients_on_shipped_checkpoints(path, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(path)
    rb, lin = params["rbf_list"], params["linear"]
    dev = cuda_device
    y64 = arr["y64"]
    args = (_t(arr["x"], dev), _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev),
            _t(2 * y64 / y64.size, dev))
    kw = dict(gd=_gd(cfg, dev), basis=orc.BASIS_IDS[cfg["basis_func"]], grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
    g = ops.backward(*args, gate_tol=GATE_TOL, **kw)
    for leaf, key in LEAVES:
        assert _rel_l2(g[leaf].cpu().numpy(), arr[key]) < GRAD_TOL, leaf
    assert _rel_l2(g["x"].cpu().numpy(), arr["g/x"]) < GRAD_TOL          # dx is not affected by gate_tol
    # deterministic: same fixed partition and order as the dense kernel
    g2 = ops.backward(*args, gate_tol=GATE_TOL, **kw)
    for leaf, _ in LEAVES:
        assert torch.equal(g[leaf], g2[leaf]), leaf
    # gate_tol = 0 leaves out only terms whose gate is exactly 0
    g0 = ops.backward(*args, gate_tol=0.0, **kw)
    for leaf, key in LEAVES:
        assert _rel_l2(g0[leaf].cpu().numpy(), arr[key]) < GRAD_TOL, leaf


@pytest.mark.parametrize("shape", [(1023, 3, 7, 4, 2, 3), (5000, 12, 100, 8, 2, 8), (3000, 27, 30, 8, 10, 3), (2049, 64, 40, 6, 16, 3),
                                   (777, 5, 33, 5, 4, 5), (4000, 130, 6, 7, 18, 3)],
                         ids=lambda s: "B%d_R%d_K%d_D%d_O%d_S%d" % s)
def test_sparse_backward_matches_dense_oracle_with_sharp_gates(shape, cuda_device):
    """Uniform queries, sharp gates (most (query, region) pairs are skipped), narrow and wide outputs, odd K; gates with
    up to 4 split dimensions (folded into dx inside the kernel) and more (gate_bwd_kernel); 130 regions (dense dx)."""
    ops = _ops()
    B, R, K, D, O, S = shape
    cfg, p, x = _synthetic(43, B, R, K, D, O, S=S, nb=(4, 3, 2))
    cfg["delta"] = [40.0] * len(cfg["delta"])
    dev = cuda_device
    rng = np.random.default_rng(5)
    dy = (rng.standard_normal((B, O)) / B).astype(np.float32)
    x64 = x.astype(np.float64)
    gamma = orc.region_activation(x64, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                  cfg["dimension_ranges"])
    assert (gamma <= GATE_TOL).mean() > 0.3            # the case really is sparse
    for basis in (0, 1):
        ref = orc.backward(x64, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, dy.astype(np.float64), basis)
        args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(dy, dev))
        g = ops.backward(*args, gd=_gd(cfg, dev), basis=basis, gate_tol=GATE_TOL, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
        for leaf in ("centers", "log_sigs", "kernel", "bias"):
            assert _rel_l2(g[leaf].cpu().numpy(), ref[leaf]) < GRAD_TOL, (basis, leaf)
        # dx: binned queries, regions skipped per warp; the full input gradient (RBF layers + gate path)
        cfg["basis_func"] = "gaussian" if basis == 0 else "inverse_quadratic"
        params = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
                  "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
        refx = orc.wcrbf_vjp(x64, params, cfg, dy.astype(np.float64))["x"]
        assert _rel_l2(g["x"].cpu().numpy(), refx) < GRAD_TOL, basis
        g2 = ops.backward(*args, gd=_gd(cfg, dev), basis=basis, gate_tol=GATE_TOL, grad_mask=ops.GRAD_X)
        assert torch.equal(g["x"], g2["x"])                # the binning order cannot change a query's gradient
        # an externally supplied gate (what a training step hands over from irbfn_forward_gate) takes the same path
        gate = ops.region_gate(_t(x, dev), _gd(cfg, dev), R)
        ge = ops.backward(*args, gate=gate, basis=basis, gate_tol=GATE_TOL)
        for leaf in ("centers", "log_sigs", "kernel", "bias"):
            assert torch.equal(g[leaf], ge[leaf]), (basis, leaf)


def test_sparse_backward_errors(cuda_device):
    from irbfn_b200._lib import IrbfnError
    ops = _ops()
    cfg, p, x = _synthetic(3, 64, 2, 5, 4, 2)
    dev = cuda_device
    args = (_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
            torch.zeros((64, 2), device=dev))
    with pytest.raises(IrbfnError):
        ops.backward(*args, gd=_gd(cfg, dev), basis=0, gate_tol=-1.0)
    with pytest.raises(IrbfnError):
        ops.backward(*args, gd=_gd(cfg, dev), basis=0, gate_tol=float("nan"))


def test_module_autograd_uses_the_sparse_vjp_when_gate_tol_is_set(cuda_device):
    """torch.autograd.grad through WCRBFNet.apply (the stand-in for jax.grad through the Flax module): with
    ``module.gate_tol`` set the custom VJP is irbfn_backward_sparse; gradients agree with the dense VJP."""
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    leaves = [dp["params"]["rbf_list"]["centers"], dp["params"]["rbf_list"]["log_sigs"], dp["params"]["linear"]["kernel"]]
    for t in leaves:
        t.requires_grad_(True)
    x = torch.tensor(np.tile(arr["x"], (8, 1)).astype(np.float32), device=cuda_device, requires_grad=True)

    def grads():
        y = net.apply(dp, x)
        return torch.autograd.grad((y * y).mean(), leaves + [x])

    dense = grads()
    net.gate_tol = GATE_TOL
    sparse = grads()
    net.gate_tol = None
    for a, b in zip(sparse, dense):
        assert _rel_l2(a.cpu().numpy(), b.cpu().numpy()) < 1e-5
    again = grads()
    for a, b in zip(again, dense):
        assert torch.equal(a, b)                      # switching it off restores the dense VJP


def test_gate_sparse_training_step_matches_the_dense_step(cuda_device):
    """TrainState(gate_tol=..., fwd_tol_abs=...): gate-sparse VJP (and forward) inside the data-parallel step; the
    flat gradient and the loss agree with the dense step, and an optimiser step moves the parameters alike."""
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.train import TrainState, grad_step, train_step
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    lo = np.array([min(b) for b in cfg["lower_bounds"]], np.float32)
    hi = np.array([max(b) for b in cfg["upper_bounds"]], np.float32)
    g = torch.Generator().manual_seed(3)
    n = 1 << 18                       # the gate-sparse forward is taken from 2^18 samples per shard on
    x = (torch.rand((n, 8), generator=g) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(cuda_device)
    y = torch.randn((n, 2), generator=g).to(cuda_device)
    dense = TrainState(net, dp, lr=1e-3, device=cuda_device)
    yscale = float(net.apply(dp, x).abs().max())
    for kw in (dict(gate_tol=GATE_TOL), dict(gate_tol=GATE_TOL, fwd_tol_abs=1e-6 * yscale)):
        sparse = TrainState(net, dp, lr=1e-3, device=cuda_device, **kw)
        l0 = grad_step(dense, x, y)
        l1 = grad_step(sparse, x, y)
        assert abs(float(l0) - float(l1)) <= 1e-5 * abs(float(l0))
        for k in dense.gviews:
            assert _rel_l2(sparse.gviews[k].cpu().numpy(), dense.gviews[k].cpu().numpy()) < 1e-5, (kw, k)
        train_step(sparse, x, y)
        d2 = TrainState(net, dp, lr=1e-3, device=cuda_device)
        train_step(d2, x, y)
        assert _rel_l2(sparse.flat.cpu().numpy(), d2.flat.cpu().numpy()) < 1e-6
