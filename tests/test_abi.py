"""The C-ABI library loads on a machine without a GPU, exports every symbol the header
declares, and validates arguments before it touches a device (no compute calls here)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from conftest import ROOT
from irbfn_b200 import _lib

HEADER = os.path.join(ROOT, "include", "irbfn_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(irbfn_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_and_loads():
    lib = _lib.load()
    assert lib.irbfn_abi_version() == _lib.ABI_VERSION


def test_every_declared_symbol_is_exported():
    declared = _declared_functions()
    assert set(declared) == set(_lib.EXPORTED_SYMBOLS)
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r" T (irbfn_[a-z0-9_]+)$", out, flags=re.M))
    assert set(declared) <= exported
    lib = _lib.load()
    for name in declared:
        assert getattr(lib, name) is not None


def test_every_declared_symbol_has_a_row_in_integration_md():
    """INTEGRATION.md section 0 names the reference interface each entry point replaces; a symbol added to the header
    without a row there fails here (suffix forms like "(+ `_workspace_bytes`)" count for their base symbol)."""
    src = open(HEADER).read()
    doc = open(os.path.join(os.path.dirname(HEADER), "..", "INTEGRATION.md")).read()
    syms = sorted(set(re.findall(r"\b(irbfn_[a-z0-9_]+)\s*\(", src)))
    suffixes = ("_workspace_bytes", "_available", "_max_n", "_buffer_bytes")
    missing = [x for x in syms if x not in doc
               and not any(x.endswith(suf) and x[: -len(suf)] in doc and suf in doc for suf in suffixes)]
    assert not missing, missing


def test_header_constants_match_binding():
    src = open(HEADER).read()
    val = lambda name: int(re.search(r"#define\s+%s\s+\(?(-?\w+)\)?" % name, src).group(1).rstrip("u"), 0)
    assert val("IRBFN_ABI_VERSION") == _lib.ABI_VERSION
    assert val("IRBFN_BASIS_GAUSSIAN") == _lib.BASIS_GAUSSIAN
    assert val("IRBFN_BASIS_INVERSE_QUADRATIC") == _lib.BASIS_INVERSE_QUADRATIC
    assert val("IRBFN_MAX_D") == _lib.MAX_D
    assert val("IRBFN_GRAD_CENTERS") == _lib.GRAD_CENTERS
    assert val("IRBFN_GRAD_LOG_SIGS") == _lib.GRAD_LOG_SIGS
    assert val("IRBFN_GRAD_KERNEL") == _lib.GRAD_KERNEL
    assert val("IRBFN_GRAD_X") == _lib.GRAD_X
    assert val("IRBFN_GRAD_GATE") == _lib.GRAD_GATE


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "irbfn_b200.h"\nint main(void){IrbfnGateDesc g; (void)g; return IRBFN_OK;}\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    "-c", str(src), "-o", str(tmp_path / "t.o")], check=True)


def test_argument_validation_happens_before_any_device_work():
    lib = _lib.load()
    one = C.c_void_p(16)  # never dereferenced: validation fails first
    fwd = lambda **k: lib.irbfn_forward(one, one, one, one, one, one, None, one, None,
                                        k.get("B", 4), k.get("R", 2), k.get("K", 3), k.get("D", 4),
                                        k.get("O", 2), k.get("basis", 0), None, 0, None)
    assert fwd(R=0) == -1                      # IRBFN_ERR_BAD_DIMS
    assert fwd(B=-1) == -1
    assert fwd(D=_lib.MAX_D + 1) == -2         # IRBFN_ERR_UNSUPPORTED_D
    assert fwd(basis=7) == -3                  # IRBFN_ERR_BAD_BASIS
    # neither / both gate inputs
    gd = _lib.GateDesc(1, 1, 1, 16, 16, 16, 16, 16)
    assert lib.irbfn_forward(one, one, one, one, one, None, None, one, None, 4, 2, 3, 4, 2, 0, None, 0, None) == -5
    assert lib.irbfn_forward(one, one, one, one, one, one, C.byref(gd), one, None, 4, 2, 3, 4, 2, 0, None, 0, None) == -5
    # no outputs requested
    assert lib.irbfn_forward(one, one, one, one, one, one, None, None, None, 4, 2, 3, 4, 2, 0, None, 0, None) == -4
    # empty batch is a no-op success even without a device
    assert fwd(B=0) == 0
    for code in range(0, -10, -1):
        assert lib.irbfn_error_string(code)
    assert b"unknown" in lib.irbfn_error_string(-99)


def test_forward_path_is_a_host_side_decision():
    """irbfn_forward_path (no device needed): the planner's batch sizes take the latency kernel, shard-sized batches
    the split kernel, the lattice the tiled one; asking for more than y, wide outputs or a large parameter set keeps
    the throughput kernels; bad dimensions are rejected."""
    lib = _lib.load()
    path = lib.irbfn_forward_path
    C2 = (12, 100, 8, 2)
    assert [path(B, *C2, 1) for B in (1, 400, 4096, 4369)] == [_lib.PATH_SMALL] * 4
    assert path(4370, *C2, 1) == _lib.PATH_SPLIT and path(32768, *C2, 1) == _lib.PATH_SPLIT
    assert path(1 << 20, *C2, 1) == _lib.PATH_TILED
    assert path(1, *C2, 0) == _lib.PATH_SPLIT                      # hidden activations / gate output wanted
    assert path(1, 12, 100, 8, 18, 1) == _lib.PATH_SPLIT           # more than one output window
    assert path(1, 64, 4096, 6, 16, 1) == _lib.PATH_SPLIT          # C5: 7.6 MB of parameters per CTA would not do
    assert path(1, 1, 2, 4, 1, 1) == _lib.PATH_SMALL               # a single kernel tile cannot be split anyway
    assert path(0, *C2, 1) == -1 and path(4, 12, 100, 9, 2, 1) == -1
    # the small path needs no workspace, whatever irbfn_forward_workspace_bytes offers for the split one
    assert lib.irbfn_forward_workspace_bytes(400, *C2, 1) > 0


def test_backward_argument_validation():
    lib = _lib.load()
    one = C.c_void_p(16)
    bwd = lambda mask, **k: lib.irbfn_backward(
        one, one, one, one, one, None, one, None, k.get("dC", one), one, one, one, k.get("dx", one), None,
        k.get("B", 4), k.get("R", 2), 3, k.get("D", 4), 2, 0, mask, None, 0, None)
    assert bwd(0) == -9                                    # IRBFN_ERR_BAD_MASK: empty
    assert bwd(_lib.GRAD_CENTERS, dC=None) == -9           # asks for an output with a NULL pointer
    assert bwd(_lib.GRAD_X, dx=None) == -9
    assert bwd(_lib.GRAD_PARAMS, R=0) == -1
    assert bwd(_lib.GRAD_PARAMS, D=9) == -2
    assert bwd(_lib.GRAD_PARAMS, B=0) == 0
    assert lib.irbfn_backward_workspace_bytes(1 << 18, 12, 100, 8, 2, _lib.GRAD_PARAMS) > 0
    # dx alone needs the (B,R) scratch for d gamma of the tanh gate plus the binning scratch of the gate-sparse pass
    # (irbfn_backward_sparse: 8 bytes per query, a histogram and the interval edges -- under 7 MB here)
    n = lib.irbfn_backward_workspace_bytes(1 << 18, 12, 100, 8, 2, _lib.GRAD_X)
    assert (1 << 18) * (12 * 4 + 8) <= n <= (1 << 18) * (12 * 4 + 8) + (5 << 20)


def test_no_cpu_fallback_when_library_missing(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()
