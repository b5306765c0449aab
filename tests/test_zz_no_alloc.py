"""irbfn_forward_noalloc (include/irbfn_b200.h): the forward for callers that cannot allocate.  It has no counterpart
in the reference; its scratch is a static __device__ array inside the
library, so the check is against the existing entry point irbfn_forward on the same inputs, through the C ABI, on two
calls in a row (the array keeps the previous call's contents), plus the size limit's error path.

Collected last on purpose (file name).
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden, rel_max
from oracle import irbfn_oracle as orc

pytestmark = pytest.mark.gpu

OUT_TOL = 1e-5          # the forward tests' tolerance (tests/test_forward_gpu.py)
SCRATCH_BYTES = 8 << 20  # IRBFN_NOALLOC_SCRATCH_BYTES


def _fixture(name):
    return load_golden([p for p in golden_paths() if os.path.basename(p)[:-4] == name][0])


def _t(a, dev):
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device=dev)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _call_noalloc(lib, x, c, ls, W, b, gd, y, basis, stream):
    B, D = x.shape
    R, K = ls.shape
    return lib.irbfn_forward_noalloc(_ptr(x), _ptr(c), _ptr(ls), _ptr(W), _ptr(b), None, gd.ref(), _ptr(y), None,
                                     B, R, K, D, W.shape[1], basis, C.c_void_p(stream))


def _call_existing(lib, x, c, ls, W, b, gd, y, basis, stream):
    B, D = x.shape
    R, K = ls.shape
    O = W.shape[1]
    nbytes = lib.irbfn_forward_workspace_bytes(B, R, K, D, O, 1)
    ws = torch.empty(max(nbytes, 4), dtype=torch.uint8, device=x.device)
    rc = lib.irbfn_forward(_ptr(x), _ptr(c), _ptr(ls), _ptr(W), _ptr(b), None, gd.ref(), _ptr(y), None,
                           B, R, K, D, O, basis, _ptr(ws), nbytes, C.c_void_p(stream))
    return rc, nbytes


@pytest.mark.parametrize("name", ["c1_cartesian_1x1000", "c2_frenet_12x100"])
def test_noalloc_forward_matches_existing_entry_point_twice(name, cuda_device):
    from irbfn_b200 import _lib, ops
    lib = _lib.load()
    dev = cuda_device
    cfg, params, arr = _fixture(name)
    rb = params["rbf_list"]
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    c, ls = _t(rb["centers"], dev), _t(rb["log_sigs"], dev)
    W, b = _t(params["linear"]["kernel"], dev), _t(params["linear"]["bias"], dev)
    gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    x0 = arr["x"].astype(np.float32)
    lo, hi = x0.min(0), x0.max(0)
    g = torch.Generator(device="cpu").manual_seed(5)
    used_scratch = False
    # two calls in a row with DIFFERENT batches (sizes that take the split path, so the scratch array is used and
    # the second call finds the first one's partial sums in it), then the first batch again
    for B in (16384, 9001, 16384):
        assert ops.forward_path(B, ls.shape[0], ls.shape[1], x0.shape[1], W.shape[1]) == "split"
        x = (torch.tensor(lo) + torch.rand((B, x0.shape[1]), generator=g) * torch.tensor(hi - lo)).to(dev).contiguous()
        y_new = torch.full((B, W.shape[1]), float("nan"), device=dev)
        y_old = torch.full((B, W.shape[1]), float("nan"), device=dev)
        rc_old, nbytes = _call_existing(lib, x, c, ls, W, b, gd, y_old, basis, stream)
        assert rc_old == 0
        used_scratch |= nbytes > 0
        torch.cuda.synchronize(dev)
        free_before, _ = torch.cuda.mem_get_info(dev)
        rc = _call_noalloc(lib, x, c, ls, W, b, gd, y_new, basis, stream)
        assert rc == 0, lib.irbfn_last_error_detail()
        torch.cuda.synchronize(dev)
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert free_after >= free_before            # the call took no device memory
        assert torch.isfinite(y_new).all()
        assert rel_max(y_new.cpu().numpy(), y_old.cpu().numpy().astype(np.float64)) < OUT_TOL
        # and against the float64 oracle, like every forward test
        xs = x.cpu().numpy().astype(np.float64)
        gamma = orc.region_activation(xs, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                      cfg["dimension_ranges"])
        y_ref = orc.forward(xs, rb["centers"], rb["log_sigs"], params["linear"]["kernel"], params["linear"]["bias"],
                            gamma, basis)
        assert rel_max(y_new.cpu().numpy(), y_ref) < OUT_TOL
    assert used_scratch, "the test shapes no longer take the split path: pick sizes that use the scratch array"


def test_noalloc_forward_rejects_shapes_beyond_the_builtin_scratch(cuda_device):
    from irbfn_b200 import _lib
    lib = _lib.load()
    dev = cuda_device
    R, K, D, O = 1, 4096, 6, 16
    big = None
    for B in range(32768, 600000, 8192):
        if lib.irbfn_forward_workspace_bytes(B, R, K, D, O, 0) > SCRATCH_BYTES:
            big = B
            break
    assert big is not None, "no probe shape exceeds the built-in scratch; widen the search"
    x = torch.zeros((big, D), device=dev)
    gate = torch.ones((big, R), device=dev)
    c, ls = torch.zeros((R, K, D), device=dev), torch.zeros((R, K), device=dev)
    W, b = torch.zeros((K, O), device=dev), torch.zeros((O,), device=dev)
    y = torch.full((big, O), 7.0, device=dev)
    rc = lib.irbfn_forward_noalloc(_ptr(x), _ptr(c), _ptr(ls), _ptr(W), _ptr(b), _ptr(gate), None, _ptr(y), None,
                                   big, R, K, D, O, 0, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == -6                                  # IRBFN_ERR_WORKSPACE
    detail = lib.irbfn_last_error_detail().decode()
    assert "irbfn_forward_noalloc" in detail and str(SCRATCH_BYTES) in detail
    torch.cuda.synchronize(dev)
    assert bool((y == 7.0).all())                    # nothing was launched
