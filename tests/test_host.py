"""Host-side logic that needs no GPU: basis mapping, module attributes, checkpoint codec, launch planning."""
import ctypes as C
import os

import numpy as np
import pytest

from irbfn_b200 import _lib, basis, checkpoint
from irbfn_b200.model import ClusterWCRBFNet, DeeperWCRBFNet, WCRBFNet

B200_SMEM = 232448


def test_basis_maps_by_name_and_rejects_undefined_functions():
    assert basis.basis_id(basis.gaussian) == _lib.BASIS_GAUSSIAN
    assert basis.basis_id(basis.inverse_quadratic) == _lib.BASIS_INVERSE_QUADRATIC
    assert basis.basis_id("gaussian") == 0

    def gaussian_narrow(a):      # named in 2 reference configs, defined nowhere in the reference
        return a

    with pytest.raises(NotImplementedError):
        basis.basis_id(gaussian_narrow)
    a = np.linspace(0, 3, 7)
    # gaussian_wide / gaussian_wider (flax_rbf.py:40-47) run on the gaussian kernel with shifted log-widths
    assert basis.basis_id(basis.gaussian_wide) == _lib.BASIS_GAUSSIAN == basis.basis_id("gaussian_wider")
    for fn, c in ((basis.gaussian_wide, 0.1), (basis.gaussian_wider, 0.01)):
        np.testing.assert_allclose(fn(a), np.exp(-c * a ** 2))
        ls = 0.3
        d2 = a ** 2 * np.exp(2 * ls)                                   # alpha^2 = d^2 / sigma^2
        np.testing.assert_allclose(np.exp(-c * a ** 2), np.exp(-d2 * np.exp(-2 * (ls + basis.log_sig_shift(fn)))), rtol=1e-12)
    assert basis.log_sig_shift(basis.gaussian) == 0.0 == basis.log_sig_shift("inverse_quadratic")
    np.testing.assert_allclose(basis.gaussian(a), np.exp(-a ** 2))
    np.testing.assert_allclose(basis.inverse_quadratic(a), 1 / (1 + a ** 2))


def test_module_mirrors_reference_attributes():
    net = WCRBFNet(in_features=5, out_features=10, num_kernels=100, basis_func=basis.inverse_quadratic,
                   num_regions=4, lower_bounds=[[-1.0, 0.0], [-2.0, 0.5]], upper_bounds=[[0.2, 1.0], [0.6, 2.0]],
                   dimension_ranges=[[0, 0], [0, 1], [1, 0], [1, 1]], activation_idx=[0, 1], delta=[100.0, 100.0])
    # Flax freezes list attributes to tuples (what makes them static for jit, model.py:42)
    assert net.lower_bounds == ((-1.0, 0.0), (-2.0, 0.5)) and isinstance(net.dimension_ranges, tuple)
    assert net.num_split_dimensions == 2 and hash(net.lower_bounds) is not None
    assert not net.fixed_centers and not net.fixed_width and net.centers is None
    for cls in (DeeperWCRBFNet, ClusterWCRBFNet):
        m = cls(8, 10, 10, basis.gaussian, 3)
        assert (m.in_features, m.out_features, m.num_kernels, m.num_regions) == (8, 10, 10, 3)
    with pytest.raises(ValueError):
        WCRBFNet(9, 2, 10, basis.gaussian, 1)                      # D above the compiled maximum
    with pytest.raises(NotImplementedError):
        WCRBFNet(4, 2, 10, lambda a: a, 1)                          # unknown basis callable
    with pytest.raises(ValueError):
        WCRBFNet(4, 2, 10, basis.gaussian, 2, centers=np.zeros((3, 4)))   # centres of the wrong shape
    m = WCRBFNet(4, 2, 10, basis.gaussian, 2, centers=np.ones((10, 4)), fixed_centers=True)
    assert m.centers.shape == (2, 10, 4)                            # (K,D) broadcast over regions


def test_flax_msgpack_checkpoint_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    tree = {"step": np.int32(7),
            "params": {"params": {"rbf_list": {"centers": rng.normal(size=(2, 3, 4)), "log_sigs": np.zeros((2, 3))},
                                  "linear": {"kernel": rng.normal(size=(3, 2)).astype(np.float32),
                                             "bias": np.zeros(2, np.float32)}}},
            "opt_state": {"0": {"count": np.int32(7)}}}
    checkpoint.save_checkpoint(str(tmp_path), tree, step=5)
    path = checkpoint.save_checkpoint(str(tmp_path), tree, step=40)
    assert checkpoint.latest_checkpoint(str(tmp_path)) == path        # natural ordering: 40 > 5
    raw = checkpoint.load_checkpoint(path)
    assert raw["params"]["params"]["rbf_list"]["centers"].dtype == np.float64   # stored dtype survives
    p = checkpoint.restore_params(str(tmp_path))
    leaf = p["params"]["rbf_list"]["centers"]
    assert leaf.dtype == np.float32                                    # x64 is off on the inference path (F6)
    np.testing.assert_allclose(leaf, tree["params"]["params"]["rbf_list"]["centers"], rtol=1e-7)
    os.makedirs(tmp_path / "nothing")
    with pytest.raises(FileNotFoundError):
        checkpoint.restore_params(str(tmp_path / "nothing"))


def _plan(R, K, D, O, smem=B200_SMEM):
    lib = _lib.load()
    rc, kc, nt, sz = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
    code = lib.irbfn_forward_plan(R, K, D, O, smem, C.byref(rc), C.byref(kc), C.byref(nt), C.byref(sz))
    return code, rc.value, kc.value, nt.value, sz.value


@pytest.mark.parametrize("shape", [(1, 1000, 7, 2), (4, 100, 5, 10), (12, 100, 8, 2), (1, 1000, 8, 10),
                                   (27, 500, 8, 10), (128, 10, 7, 10), (500, 10, 8, 10)])
def test_shipped_shapes_are_staged_once(shape):
    R, K, D, O = shape
    code, rc, kc, nt, sz = _plan(R, K, D, O)
    assert code == 0 and sz <= B200_SMEM and nt in (256, 512)
    assert kc % 4 == 0 and 1 <= rc <= R
    if R <= 12:
        assert rc == R and kc >= K                          # whole parameter set resident
    else:
        assert rc < R or kc < K or sz > 100000              # the (R, tile) gate table forces region chunks


def test_stress_shape_is_chunked_within_shared_memory():
    code, rc, kc, nt, sz = _plan(64, 4096, 6, 16)           # BASELINE config 5: 7.6 MB of parameters
    assert code == 0 and sz <= B200_SMEM
    assert rc < 64 or kc < 4096
    assert kc % 4 == 0 and kc >= 64
    assert _plan(3520, 100, 4, 5)[0] == 0                   # the IROS-2023 many-region shape
    assert _plan(12, 100, 8, 2, smem=2048)[0] != 0          # impossible budget -> error code, no crash
    assert _plan(12, 100, 9, 2)[0] == -2                    # D > IRBFN_MAX_D


# ---- reference-held checkpoint data (tests/golden/make_golden.py: make_checkpoint_fixture) -----------------
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_reader_decodes_real_flax_checkpoint_bytes():
    """irbfn_b200/checkpoint.py on the raw bytes of a checkpoint the REFERENCE wrote
    (scripts/ckpts/nmpc_4_regions/checkpoint_0), leaf by leaf against the fixture parameters that
    tests/golden/make_golden.py decoded with the oracle's separate reader."""
    raw = checkpoint.load_checkpoint(os.path.join(GOLDEN, "ref_ckpt_nmpc_4_regions_checkpoint_0.msgpack"))
    assert set(raw) == {"step", "params", "opt_state"} and int(raw["step"]) == 5184000
    p = raw["params"]["params"]
    assert p["rbf_list"]["centers"].dtype == np.float64 and p["linear"]["kernel"].dtype == np.float32   # F6
    z = np.load(os.path.join(GOLDEN, "c1b_nmpc_4_regions.npz"))
    restored = checkpoint.restore_params(os.path.join(GOLDEN, "ref_ckpt_nmpc_4_regions_checkpoint_0.msgpack"))["params"]
    for path in ("rbf_list/centers", "rbf_list/log_sigs", "linear/kernel", "linear/bias"):
        a, b = path.split("/")
        assert restored[a][b].dtype == np.float32
        assert np.array_equal(restored[a][b], z["p/" + path]), path
        assert np.array_equal(p[a][b].astype(np.float32), z["p/" + path]), path


def _layout(node):
    if isinstance(node, dict):
        return {str(k): _layout(v) for k, v in node.items()}
    a = np.asarray(node)
    return [list(a.shape), a.dtype.name]


def test_train_state_writes_and_resumes_the_reference_checkpoint_layout(tmp_path):
    """TrainState.state_dict() has the key / shape structure of a checkpoint the reference wrote with
    optax.chain(clip_by_global_norm, adam) (layout recorded from scripts/ckpts/dnmpc_12regions_frenet_l1_bigdata/
    checkpoint_9500), survives the msgpack codec, and load_state_dict() resumes params, moments and count."""
    import json
    import torch
    from irbfn_b200.train import TrainState
    with open(os.path.join(GOLDEN, "ckpt_layouts.json")) as f:
        layouts = json.load(f)
    want = layouts["scripts/ckpts/dnmpc_12regions_frenet_l1_bigdata/checkpoint_9500"]
    z = np.load(os.path.join(GOLDEN, "c2_frenet_12x100.npz"))
    cfg = json.loads(str(z["cfg_json"]))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], basis.gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
    tree = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                       "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
    st = TrainState(net, tree, lr=1e-3, device="cpu")
    st.mu.copy_(torch.arange(st.mu.numel()) * 1e-3)
    st.nu.copy_(torch.arange(st.nu.numel()) * 2e-3)
    st.step = 41
    sd = st.state_dict()

    def shapes_only(n):
        return {k: shapes_only(v) for k, v in n.items()} if isinstance(n, dict) else n[0]
    assert shapes_only(_layout(sd)) == shapes_only(want)                 # same keys, nesting and shapes
    assert _layout(sd)["opt_state"]["1"]["0"]["count"] == want["opt_state"]["1"]["0"]["count"]   # int32 scalar
    path = checkpoint.save_checkpoint(str(tmp_path), sd, step=41)
    back = checkpoint.load_checkpoint(path)
    assert shapes_only(_layout(back)) == shapes_only(want)
    st2 = TrainState(net, tree, lr=1e-3, device="cpu")
    st2.flat.zero_()
    st2.load_state_dict(back)
    assert st2.step == 41 and float(st2.step_t) == 41.0
    assert torch.equal(st2.flat, st.flat) and torch.equal(st2.mu, st.mu) and torch.equal(st2.nu, st.nu)
    # the oldest shipped checkpoints hold a bare Adam state (opt_state['0'] = {count, mu, nu}): resumable too
    old = checkpoint.load_checkpoint(os.path.join(GOLDEN, "ref_ckpt_nmpc_4_regions_checkpoint_0.msgpack"))
    z4 = np.load(os.path.join(GOLDEN, "c1b_nmpc_4_regions.npz"))
    c4 = json.loads(str(z4["cfg_json"]))
    net4 = WCRBFNet(c4["in_features"], c4["out_features"], c4["num_kernels"], basis.inverse_quadratic, c4["num_regions"],
                    c4["lower_bounds"], c4["upper_bounds"], c4["dimension_ranges"], c4["activation_idx"], c4["delta"])
    st4 = TrainState(net4, old["params"], lr=1e-3, device="cpu").load_state_dict(old)
    assert st4.step == int(old["opt_state"]["0"]["count"]) and np.array_equal(st4.views["kernel"].numpy(), z4["p/linear/kernel"])
