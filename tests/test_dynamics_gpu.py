"""Loss-side dynamics on the GPU against the torch-float64 restatement of
the reference's dynamics.py and training losses (oracle/dynamics_oracle.py).

Tolerances: rollout states  max|s - s_ref| / max|s_ref| <= 1e-5 (fp32 Euler steps);
loss terms rel <= 1e-5; cotangent dy  ||g - g_ref||_2 / ||g_ref||_2 <= 1e-4."""
import numpy as np
import pytest
import torch

from irbfn_b200 import _lib, ops
from oracle import dynamics_oracle as dyn

pytestmark = pytest.mark.gpu
P = dyn.default_dyn_params()


def _frenet_x(rng, B):
    lo = np.array([-0.2, -0.3, 1.0, -1.0, 3.0, -2.6, -1.0, -0.1])
    hi = np.array([2.0, 0.3, 7.0, 1.0, 7.0, 2.6, 1.0, 0.1])
    return rng.uniform(lo, hi, (B, 8))


def _controls(rng, B, O):
    u = rng.normal(size=(B, O))
    if O == 10:
        u[:, :5] *= 6.0      # some accelerations beyond a_max = 9.51
        u[:, 5:] *= 2.5      # some steering velocities beyond sv_max = 3.2
    else:
        u *= np.array([6.0, 2.5])
    return u


def _dev(a, dev):
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device=dev)


@pytest.mark.parametrize("mode,D,O", [(0, 8, 2), (0, 8, 10), (1, 8, 10), (2, 7, 2), (2, 7, 10), (3, 7, 10)])
@pytest.mark.parametrize("B", [1, 1000, 70001])
def test_loss_and_cotangent_match_oracle(mode, D, O, B, cuda_device):
    rng = np.random.default_rng(100 * mode + O + B)
    x = _frenet_x(rng, B) if D == 8 else rng.uniform(-1.0, 8.0, (B, D))
    y = _controls(rng, B, O)
    yp = y + rng.normal(size=(B, O)) * 0.7
    x32, y32, yp32 = (a.astype(np.float32) for a in (x, y, yp))
    loss, pl, il, g = dyn.training_loss_and_grad(mode, x32.astype(np.float64), yp32.astype(np.float64),
                                                 y32.astype(np.float64), P)
    losses, dy = ops.dynamics_loss(mode, _dev(x32, cuda_device), _dev(yp32, cuda_device), _dev(y32, cuda_device), P)
    lv = losses.cpu().numpy()
    assert abs(lv[0] - pl) <= 1e-5 * abs(pl)
    # the integration term is a difference of two fp32 rollouts that share their initial state
    # ((v0 + a dt) - (v0 + a' dt), exactly what the reference computes in fp32): cancellation leaves
    # ~1e-4 relative accuracy, so this term gets the gradient tolerance
    assert abs(lv[1] - il) <= 2e-4 * abs(il) + 1e-12
    e = np.linalg.norm(dy.cpu().numpy() - g) / np.linalg.norm(g)
    assert e <= 1e-4, e


def test_shard_losses_and_cotangents_add_up(cuda_device):
    """global_batch > B: two shards evaluated separately give the one-device loss and cotangent."""
    rng = np.random.default_rng(5)
    B = 4096
    x, y = _frenet_x(rng, B), _controls(rng, B, 10)
    yp = y + rng.normal(size=(B, 10)) * 0.5
    xd, yd, ypd = _dev(x, cuda_device), _dev(y, cuda_device), _dev(yp, cuda_device)
    l_all, g_all = ops.dynamics_loss(1, xd, ypd, yd, P)
    h = 1500
    l0, g0 = ops.dynamics_loss(1, xd[:h].contiguous(), ypd[:h].contiguous(), yd[:h].contiguous(), P, global_batch=B)
    l1, g1 = ops.dynamics_loss(1, xd[h:].contiguous(), ypd[h:].contiguous(), yd[h:].contiguous(), P, global_batch=B)
    assert torch.allclose(l0 + l1, l_all, rtol=1e-6)
    assert torch.equal(torch.cat([g0, g1]), g_all)
    # deterministic
    l2, g2 = ops.dynamics_loss(1, xd, ypd, yd, P)
    assert torch.equal(l2, l_all) and torch.equal(g2, g_all)


def test_identical_prediction_has_zero_integration_loss(cuda_device):
    rng = np.random.default_rng(6)
    x, y = _frenet_x(rng, 2048), _controls(rng, 2048, 10)
    xd, yd = _dev(x, cuda_device), _dev(y, cuda_device)
    for mode in (0, 1):
        losses, dy = ops.dynamics_loss(mode, xd, yd.clone(), yd, P)
        assert losses.cpu().tolist() == [0.0, 0.0]
        assert float(dy.abs().max()) == 0.0


@pytest.mark.parametrize("model,W,fn", [
    (_lib.DYN_FRENET_MULT, 18, dyn.integrate_frenet_mult),
    (_lib.DYN_FRENET_ONESTEP_AUX, 9, dyn.dynamic_frenet_onestep_aux),
    (_lib.DYN_FRENET_ONESTEP_AUX, 17, dyn.dynamic_frenet_onestep_aux),
    (_lib.DYN_ST_MULT, 17, dyn.integrate_st_mult),
    (_lib.DYN_ST_ONESTEP_AUX, 9, dyn.dynamic_st_onestep_aux),
])
def test_rollouts_match_oracle(model, W, fn, cuda_device):
    rng = np.random.default_rng(model * 31 + W)
    B = 5000
    if model in (_lib.DYN_FRENET_MULT,):
        xu = np.concatenate([rng.uniform(-0.3, 0.3, (B, 1)), _frenet_x(rng, B)[:, [0, 1, 2, 3, 5, 6, 7]],
                             _controls(rng, B, 10)], 1)
    elif model == _lib.DYN_FRENET_ONESTEP_AUX:
        xu = np.concatenate([_frenet_x(rng, B)[:, [0, 1, 2, 3, 5, 6, 7]], _controls(rng, B, W - 7)], 1)
    else:
        # [X, Y, DELTA, V, PSI, PSI_DOT, BETA]: speeds on both sides of the V > 3 switch
        st = rng.uniform([-1, -1, -0.5, 0.5, -3, -1, -0.3], [1, 1, 0.5, 8.0, 3, 1, 0.3], (B, 7))
        xu = np.concatenate([st, _controls(rng, B, 10)[:, : W - 7]], 1)
    xu32 = xu.astype(np.float32)
    ref = fn(xu32.astype(np.float64), P).numpy()
    out = ops.integrate(model, _dev(xu32, cuda_device), P).cpu().numpy()
    assert out.shape == ref.shape
    assert np.abs(out - ref).max() / np.abs(ref).max() <= 1e-5


def test_dynamics_errors(cuda_device):
    x = torch.zeros((8, 8), device=cuda_device)
    y3 = torch.zeros((8, 3), device=cuda_device)
    y2 = torch.zeros((8, 2), device=cuda_device)
    with pytest.raises(_lib.IrbfnError):
        ops.dynamics_loss(0, x, y3, y3, P)               # O must be 2 or 10
    with pytest.raises(_lib.IrbfnError):
        ops.dynamics_loss(1, x, y2, y2, P)               # five-step rollout needs 10 controls
    with pytest.raises(_lib.IrbfnError):
        ops.dynamics_loss(7, x, y2, y2, P)
    with pytest.raises(_lib.IrbfnError):
        ops.integrate(_lib.DYN_FRENET_MULT, torch.zeros((4, 17), device=cuda_device), P)
    with pytest.raises(ValueError):
        ops.dynamics_loss(0, x.cpu(), y2, y2, P)         # no CPU path


def test_train_step_with_frenet_dynamics_loss_matches_oracle_chain(cuda_device):
    """grad_step(loss='frenet_oneint') == oracle VJP of the network fed with the oracle's float64
    loss cotangent (shipped C2 parameters, scripts/train_nmpc_frenet.py:339-392)."""
    import os
    from conftest import GOLDEN_DIR, load_golden, rel_l2
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.train import TrainState, grad_step
    from oracle import irbfn_oracle as orc

    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    x = arr["x"][:2048].astype(np.float32)
    rng = np.random.default_rng(3)
    y64 = orc.wcrbf_forward(x.astype(np.float64), params, cfg)
    t = (y64 + rng.normal(size=y64.shape) * 0.5).astype(np.float32)
    loss, pl, il, dy = dyn.training_loss_and_grad(0, x.astype(np.float64), y64, t.astype(np.float64), P)
    ref = orc.wcrbf_vjp(x.astype(np.float64), params, cfg, dy)
    state = TrainState(net, device_params({"params": params}, cuda_device), lr=1e-3, device=cuda_device)
    lval = grad_step(state, _dev(x, cuda_device), _dev(t, cuda_device), loss="frenet_oneint")
    assert abs(float(lval) - loss) <= 1e-4 * abs(loss)
    for name in ("centers", "log_sigs", "kernel", "bias"):
        assert rel_l2(state.gviews[name].cpu().numpy(), ref[name]) <= 2e-4, name


@pytest.mark.parametrize("kind", ["l2", "l1"])
@pytest.mark.parametrize("B,O,gb", [(1, 2, 1), (1000, 10, 4000), (262144, 2, 262144), (33333, 16, 33333), (5, 37, 5)])
def test_prediction_loss_matches_float64(kind, B, O, gb, cuda_device):
    """irbfn_prediction_loss against the float64 spelling of optax.l2_loss(...).mean() / jnp.abs(...).mean()
    (scripts/train_nmpc.py:287, scripts/train_nmpc_frenet.py:347): loss to 1e-6 relative, the cotangent exactly
    the float32 formula, and bit-identical run to run."""
    from irbfn_b200.train import loss_and_cotangent
    rng = np.random.default_rng(B + O)
    yp, y = rng.normal(size=(B, O)).astype(np.float32), rng.normal(size=(B, O)).astype(np.float32)
    if B > 3:
        yp[3] = y[3]                                    # exact ties: sign(0) = 0
    d = yp.astype(np.float64) - y.astype(np.float64)
    want = (0.5 * d * d if kind == "l2" else np.abs(d)).sum() / (gb * O)
    ypd, yd = torch.from_numpy(yp).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    loss, dy = ops.prediction_loss(ypd, yd, gb, kind)
    loss2, dy2 = ops.prediction_loss(ypd, yd, gb, kind)
    torch.cuda.synchronize()
    assert abs(float(loss[0]) - want) <= 1e-6 * abs(want) + 1e-12
    assert torch.equal(loss, loss2) and torch.equal(dy, dy2)
    inv = np.float32(1.0 / (float(gb) * O))
    dy_want = (yp - y) * inv if kind == "l2" else np.sign(yp - y) * inv
    assert np.array_equal(dy.cpu().numpy(), dy_want.astype(np.float32))
    lt, dyt = loss_and_cotangent(ypd, yd, gb, kind)     # the torch spelling kept for the CPU tests
    assert abs(float(lt) - want) <= 2e-5 * abs(want) + 1e-12
    assert torch.allclose(dyt, dy, rtol=1e-6, atol=0)
    with pytest.raises(ValueError):
        ops.prediction_loss(ypd, yd, gb, "huber")
    with pytest.raises(Exception):
        ops.prediction_loss(ypd, yd, B - 1 if B > 1 else 0, kind)   # global batch below the shard
