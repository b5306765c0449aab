"""The C (float32, OpenMP) oracle against the NumPy float64 oracle and the golden fixtures."""
import os

import numpy as np
import pytest

from conftest import golden_paths, load_golden, rel_l2, rel_max
from oracle import c_oracle
from oracle import irbfn_oracle as orc

WCRBF = [p for p in golden_paths() if "deeper" not in p and "cluster" not in p]


@pytest.mark.parametrize("path", WCRBF, ids=[os.path.basename(p)[:-4] for p in WCRBF])
def test_c_oracle_forward_matches_golden(path):
    cfg, params, arr = load_golden(path)
    rb, lin = params["rbf_list"], params["linear"]
    y = c_oracle.forward(arr["x"], rb["centers"], rb["log_sigs"], lin["kernel"], lin["bias"],
                         orc.BASIS_IDS[cfg["basis_func"]], cfg=cfg)
    # float32 arithmetic in the reference's sqrt/divide/tanh form: 2e-5 norm-wise (F10)
    assert rel_max(y, arr["y64"]) < 2e-5


@pytest.mark.parametrize("path", WCRBF[:4], ids=[os.path.basename(p)[:-4] for p in WCRBF[:4]])
def test_c_oracle_backward_matches_golden(path):
    cfg, params, arr = load_golden(path)
    rb, lin = params["rbf_list"], params["linear"]
    y64 = arr["y64"]
    dy = (2 * y64 / y64.size).astype(np.float32)
    g = c_oracle.backward(arr["x"], rb["centers"], rb["log_sigs"], lin["kernel"], dy,
                          orc.BASIS_IDS[cfg["basis_func"]], cfg=cfg)
    assert rel_l2(g["centers"], arr["g/rbf_list/centers"]) < 1e-4
    assert rel_l2(g["log_sigs"], arr["g/rbf_list/log_sigs"]) < 1e-4
    assert rel_l2(g["kernel"], arr["g/linear/kernel"]) < 1e-4
    assert rel_l2(g["bias"], arr["g/linear/bias"]) < 1e-4
    assert rel_l2(g["x"], arr["g/x"]) < 1e-4


def test_c_oracle_external_gate_and_hidden():
    cfg, params, arr = load_golden([p for p in golden_paths() if "cluster" in p][0])
    rb, lin, cl = params["rbf_list"], params["linear"], params["cluster"]
    gate, _ = orc.softmax_gate(arr["x"].astype(np.float64), cl["kernel"], cl["bias"])
    y, hid = c_oracle.forward(arr["x"], rb["centers"], rb["log_sigs"], lin["kernel"], lin["bias"],
                              orc.BASIS_IDS[cfg["basis_func"]], gate=gate, want_hidden=True)
    assert rel_max(y, arr["y64"]) < 2e-5
    assert rel_max(hid @ lin["kernel"].astype(np.float64) + lin["bias"], arr["y64"]) < 2e-5
