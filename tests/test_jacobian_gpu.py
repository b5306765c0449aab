"""One-pass Jacobian dy/dx (irbfn_jacobian) against jacrev of the oracle
(one float64 VJP per output column) on shipped parameters and ragged synthetic shapes.

Tolerance: the gradient tolerance of BASELINE.json's north_star, norm-wise over the whole tensor,
||J - J_ref||_2 / ||J_ref||_2 <= 1e-4; the forward value that rides along: 1e-5 (max-norm)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, load_golden, rel_l2, rel_max
from oracle import irbfn_oracle as orc
from test_forward_gpu import SHAPES, _gd, _ops, _synthetic, _t

pytestmark = pytest.mark.gpu

WCRBF = ["c1_cartesian_1x1000", "c1b_nmpc_4_regions", "c2_frenet_12x100", "c2b_frenet_1x1000_o10",
         "dnmpc_128regions_narrow", "dnmpc_4regions_iq"]


@pytest.mark.parametrize("name", WCRBF)
def test_jacobian_matches_oracle_on_shipped_parameters(name, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, name + ".npz"))
    n = 512
    x = arr["x"][:n]
    J_ref = orc.wcrbf_jacobian(x.astype(np.float64), params, cfg)
    rb, lin = params["rbf_list"], params["linear"]
    jac, y = ops.jacobian(_t(x, cuda_device), _t(rb["centers"], cuda_device), _t(rb["log_sigs"], cuda_device),
                          _t(lin["kernel"], cuda_device), _t(lin["bias"], cuda_device), gd=_gd(cfg, cuda_device),
                          basis=orc.BASIS_IDS[cfg["basis_func"]], want_y=True)
    assert rel_l2(jac.cpu().numpy(), J_ref) <= 1e-4
    assert rel_max(y.cpu().numpy(), arr["y64"][:n]) <= 1e-5


@pytest.mark.parametrize("basis", [0, 1])
@pytest.mark.parametrize("shape", SHAPES, ids=["B%d_R%d_K%d_D%d_O%d" % s for s in SHAPES])
def test_jacobian_ragged_shapes_match_oracle(shape, basis, cuda_device):
    ops = _ops()
    B, R, K, D, O = shape
    B = min(B, 700)
    cfg, p, x = _synthetic(23, B, R, K, D, O)
    cfg["basis_func"] = "gaussian" if basis == 0 else "inverse_quadratic"
    tree = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
            "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    J_ref = orc.wcrbf_jacobian(x.astype(np.float64), tree, cfg)
    jac = ops.jacobian(_t(x, cuda_device), _t(p["centers"], cuda_device), _t(p["log_sigs"], cuda_device),
                       _t(p["kernel"], cuda_device), gd=_gd(cfg, cuda_device), basis=basis)
    assert jac.shape == (B, O, D)
    assert rel_l2(jac.cpu().numpy(), J_ref) <= 1e-4


def test_jacobian_rows_equal_the_vjp_kernel(cuda_device):
    """Each Jacobian row is the input gradient irbfn_backward returns for a one-hot cotangent."""
    ops = _ops()
    cfg, p, x = _synthetic(5, 3000, 6, 40, 6, 3)
    d = cuda_device
    xs, c, ls, W = _t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d)
    gd = _gd(cfg, d)
    jac = ops.jacobian(xs, c, ls, W, gd=gd)
    for o in range(3):
        dy = torch.zeros((3000, 3), device=d)
        dy[:, o] = 1
        dx = ops.backward(xs, c, ls, W, dy, gd=gd, grad_mask=ops.GRAD_X)["x"]
        assert rel_l2(jac[:, o, :].cpu().numpy(), dx.cpu().numpy()) <= 2e-5


def test_jacobian_external_gate_and_region_readouts(cuda_device):
    """External gate (ClusterWCRBFNet): the Jacobian is the RBF-layer term, and region_y lets the
    caller add the gate's chain rule: J_full = J + sum_r region_y[:, r, :, None] * dgate[:, r, None, :]."""
    ops = _ops()
    cfg, p, x = _synthetic(9, 900, 5, 24, 5, 4)
    d = cuda_device
    x64 = x.astype(np.float64)
    gamma = orc.region_activation(x64, 5, cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"])
    xs, c, ls, W, b = (_t(a, d) for a in (x, p["centers"], p["log_sigs"], p["kernel"], p["bias"]))
    jac, y, ry = ops.jacobian(xs, c, ls, W, b, gate=_t(gamma, d), want_y=True, want_region_y=True)
    phi = orc.rbf_activations(x64, p["centers"], p["log_sigs"], 0)               # (B,R,K)
    ry_ref = phi @ p["kernel"].astype(np.float64)                                 # (B,R,O)
    assert rel_max(ry.cpu().numpy(), ry_ref) <= 1e-5
    # full Jacobian with the tanh gate's own derivative added on the host == fused-gate Jacobian
    full = ops.jacobian(xs, c, ls, W, gd=_gd(cfg, d))
    dgate = np.stack([orc.gate_vjp_x(x64, np.eye(5)[r][None].repeat(900, 0), 5, cfg["lower_bounds"], cfg["upper_bounds"],
                                     cfg["delta"], cfg["dimension_ranges"]) for r in range(5)], 1)   # (B,R,D)
    comp = jac.cpu().numpy() + np.einsum("bro,brj->boj", ry.cpu().numpy(), dgate)
    assert rel_l2(comp, full.cpu().numpy()) <= 2e-5


def test_jacobian_parameters_larger_than_shared_memory(cuda_device):
    """16 x 2048 kernels x 8 features do not fit one CTA's shared memory: one region per stage."""
    ops = _ops()
    cfg, p, x = _synthetic(31, 600, 16, 2048, 8, 2, S=2, nb=(4, 4))
    cfg["basis_func"] = "gaussian"
    tree = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
            "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    J_ref = orc.wcrbf_jacobian(x.astype(np.float64), tree, cfg, chunk=64)
    d = cuda_device
    jac = ops.jacobian(_t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d), gd=_gd(cfg, d))
    assert rel_l2(jac.cpu().numpy(), J_ref) <= 1e-4


def test_jacobian_finite_on_centre_and_errors(cuda_device):
    ops = _ops()
    from irbfn_b200 import _lib
    cfg, p, x = _synthetic(2, 64, 2, 8, 4, 2)
    x[:8] = p["centers"][0, :8]                      # queries exactly on centres (reference: NaN, F8)
    d = cuda_device
    jac = ops.jacobian(_t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d), gd=_gd(cfg, d))
    assert torch.isfinite(jac).all()
    with pytest.raises(_lib.IrbfnError):
        ops.jacobian(_t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d))     # no gate at all
    assert ops.jacobian(_t(x[:0], d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d),
                        gd=_gd(cfg, d)).shape == (0, 2, 4)


# ---------------------------------------------------------------------------------------------------
# Gate-sparse Jacobian (irbfn_jacobian_sparse): (query, region) pairs with an interval factor <= gate_tol are left out.
# ---------------------------------------------------------------------------------------------------
GATE_TOL = 2.0 ** -24


@pytest.mark.parametrize("name", WCRBF)
def test_sparse_jacobian_matches_oracle_on_shipped_parameters(name, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, name + ".npz"))
    n = 512
    x = arr["x"][:n]
    J_ref = orc.wcrbf_jacobian(x.astype(np.float64), params, cfg)
    rb, lin = params["rbf_list"], params["linear"]
    args = (_t(x, cuda_device), _t(rb["centers"], cuda_device), _t(rb["log_sigs"], cuda_device),
            _t(lin["kernel"], cuda_device), _t(lin["bias"], cuda_device))
    kw = dict(gd=_gd(cfg, cuda_device), basis=orc.BASIS_IDS[cfg["basis_func"]], want_y=True)
    jac, y = ops.jacobian(*args, gate_tol=GATE_TOL, **kw)
    assert rel_l2(jac.cpu().numpy(), J_ref) <= 1e-4
    assert rel_max(y.cpu().numpy(), arr["y64"][:n]) <= 1e-5
    jac2, y2 = ops.jacobian(*args, gate_tol=GATE_TOL, **kw)
    assert torch.equal(jac, jac2) and torch.equal(y, y2)   # the binning order cannot change a query's rows


@pytest.mark.parametrize("shape", [(5000, 12, 100, 8, 2, 3), (3001, 27, 30, 8, 10, 3), (2049, 64, 16, 6, 5, 3),
                                   (777, 5, 33, 5, 4, 2)], ids=lambda s: "B%d_R%d_K%d_D%d_O%d_S%d" % s)
@pytest.mark.parametrize("basis", [0, 1])
def test_sparse_jacobian_with_sharp_gates(shape, basis, cuda_device):
    """Uniform queries, sharp gates: most (query, region) pairs are skipped; against the dense float64 jacrev."""
    ops = _ops()
    B, R, K, D, O, S = shape
    cfg, p, x = _synthetic(47, B, R, K, D, O, S=S, nb=(4, 3, 2))
    cfg["delta"] = [40.0] * len(cfg["delta"])
    cfg["basis_func"] = "gaussian" if basis == 0 else "inverse_quadratic"
    tree = {"rbf_list": {"centers": p["centers"], "log_sigs": p["log_sigs"]},
            "linear": {"kernel": p["kernel"], "bias": p["bias"]}}
    idx = np.arange(0, B, 7)
    J_ref = orc.wcrbf_jacobian(x[idx].astype(np.float64), tree, cfg)
    d = cuda_device
    jac, y = ops.jacobian(_t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d), _t(p["bias"], d),
                          gd=_gd(cfg, d), basis=basis, want_y=True, gate_tol=GATE_TOL)
    assert rel_l2(jac[torch.tensor(idx, device=d)].cpu().numpy(), J_ref) <= 1e-4
    y_ref = orc.wcrbf_forward(x[idx].astype(np.float64), tree, cfg)
    assert rel_max(y[torch.tensor(idx, device=d)].cpu().numpy(), y_ref) <= 1e-5
    dense = ops.jacobian(_t(x, d), _t(p["centers"], d), _t(p["log_sigs"], d), _t(p["kernel"], d), gd=_gd(cfg, d), basis=basis)
    assert rel_l2(jac.cpu().numpy(), dense.cpu().numpy().astype(np.float64)) <= 1e-5
