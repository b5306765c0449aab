"""CUDA kernels against values the REFERENCE'S OWN CODE produced (tests/golden/dynamics_ref.npz,
train_step_ref.npz: dynamics.py and the nested train_step_* functions of scripts/train_nmpc*.py executed
unmodified under oracle/refshim by tests/golden/make_golden.py).  The model fixtures (y64, g/*) used by
test_forward_gpu.py / test_backward_gpu.py are reference-executed too.

Tolerances: rollout states max|s - s_ref| / max|s_ref| <= 1e-5; loss terms rel <= 1e-5 (the integration
term is a difference of two fp32 rollouts from one initial state: 2e-4, see test_dynamics_gpu.py);
cotangents and parameter gradients ||g - g_ref||_2 / ||g_ref||_2 <= 1e-4."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, rel_l2
from irbfn_b200 import _lib, ops

pytestmark = pytest.mark.gpu


def _dev(a, dev):
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device=dev)


def test_rollouts_match_reference_executed_dynamics(cuda_device):
    z = np.load(os.path.join(GOLDEN_DIR, "dynamics_ref.npz"))
    P = z["dyn_params"]
    cases = [(_lib.DYN_FRENET_MULT, "frenet_mult_in", "frenet_mult_out"), (_lib.DYN_ST_MULT, "st_mult_in", "st_mult_out"),
             (_lib.DYN_ST_ONESTEP_AUX, "st_aux_in", "st_aux_out"),
             (_lib.DYN_FRENET_ONESTEP_AUX, "frenet_aux_in_o2", "frenet_aux_out_o2"),
             (_lib.DYN_FRENET_ONESTEP_AUX, "frenet_aux_in_o10", "frenet_aux_out_o10")]
    for model, kin, kout in cases:
        assert np.array_equal(z[kin].astype(np.float32).astype(np.float64), z[kin])      # inputs are fp32-exact
        out = ops.integrate(model, _dev(z[kin], cuda_device), P).cpu().numpy()
        ref = z[kout]
        assert out.shape == ref.shape, (kin, out.shape, ref.shape)
        assert np.abs(out - ref).max() / np.abs(ref).max() <= 1e-5, kin


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_fused_loss_matches_the_reference_train_step_functions(mode, cuda_device):
    """loss, its terms and d loss / d y_pred of train_step_oneint / train_step_fullint (both scripts)."""
    z = np.load(os.path.join(GOLDEN_DIR, "dynamics_ref.npz"))
    P = z["dyn_params"]
    x, y, yp = z["step%d_x" % mode], z["step%d_y" % mode], z["step%d_ypred" % mode]
    ref_loss, ref_g = z["step%d_loss" % mode], z["step%d_dypred" % mode]
    losses, dy = ops.dynamics_loss(mode, _dev(x, cuda_device), _dev(yp, cuda_device), _dev(y, cuda_device), P)
    lv = losses.cpu().numpy().astype(np.float64)
    assert abs(lv.sum() - ref_loss[0]) <= 2e-4 * abs(ref_loss[0])
    if len(ref_loss) == 3:                                   # the Frenet steps also return (pred_loss, int_loss)
        assert abs(lv[0] - ref_loss[1]) <= 1e-5 * abs(ref_loss[1])
        assert abs(lv[1] - ref_loss[2]) <= 2e-4 * abs(ref_loss[2])
    assert rel_l2(dy.cpu().numpy(), ref_g) <= 1e-4


def test_full_train_step_gradients_match_the_reference_step(cuda_device):
    """forward -> fused dynamics loss -> VJP (irbfn_b200.train.grad_step) against the parameter gradients
    jax.value_and_grad produced inside the reference's own train_step_* with its own WCRBFNet."""
    from irbfn_b200 import basis
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.train import TrainState, grad_step
    z = np.load(os.path.join(GOLDEN_DIR, "train_step_ref.npz"))
    meta = json.loads(str(z["meta_json"]))
    names = {0: "frenet_oneint", 1: "frenet_fullint", 2: "st_oneint", 3: "st_fullint"}
    for tag, case in meta["cases"].items():
        mode, cfg = int(tag[4:]), case["cfg"]
        net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], getattr(basis, cfg["basis_func"]),
                       cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"],
                       cfg["activation_idx"], cfg["delta"])
        tree = {"params": {"rbf_list": {"centers": z[tag + "/p/rbf_list/centers"], "log_sigs": z[tag + "/p/rbf_list/log_sigs"]},
                           "linear": {"kernel": z[tag + "/p/linear/kernel"], "bias": z[tag + "/p/linear/bias"]}}}
        st = TrainState(net, device_params(tree, cuda_device), lr=1e-3, device=cuda_device)
        st.dyn_params = z["dyn_params"]
        loss = grad_step(st, _dev(z[tag + "/x"], cuda_device), _dev(z[tag + "/y"], cuda_device), loss=names[mode])
        ref_loss = z[tag + "/loss"][0]
        assert abs(float(loss) - ref_loss) <= 2e-4 * abs(ref_loss), (tag, float(loss), ref_loss)
        for key, leaf in (("rbf_list/centers", "centers"), ("rbf_list/log_sigs", "log_sigs"),
                          ("linear/kernel", "kernel"), ("linear/bias", "bias")):
            e = rel_l2(st.gviews[leaf].cpu().numpy(), z[tag + "/g/" + key])
            assert e <= 1e-4, (tag, key, e)
