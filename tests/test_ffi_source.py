"""The XLA FFI layer (irbfn_b200/ffi/irbfn_ffi.cc + irbfn_flax.py) cannot be compiled or imported here (no JAX on
either machine).  What CAN be checked without JAX: every C entry point the handlers call is
declared in include/irbfn_b200.h with exactly the number of arguments passed; every handler the Python side
registers is defined in the C++ source; the binding is complete (forward, hidden, external gate, all backward
variants, Jacobian, losses) and is not imported by the product."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "irbfn_b200.h")
FFI_CC = os.path.join(ROOT, "irbfn_b200", "ffi", "irbfn_ffi.cc")
FFI_PY = os.path.join(ROOT, "irbfn_b200", "ffi", "irbfn_flax.py")


def _strip_comments(src):
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def _split_args(argtext):
    depth, cur, out = 0, "", []
    for ch in argtext:
        if ch in "([":
            depth += 1
        elif ch in ")]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def _declared():
    src = _strip_comments(open(HEADER).read())
    decl = {}
    for m in re.finditer(r"\b(irbfn_[a-z_0-9]+)\s*\(([^;{]*)\)\s*;", src):
        args = _split_args(m.group(2))
        decl[m.group(1)] = 0 if args == ["void"] else len(args)
    return decl


def _calls(src):
    """(name, number of arguments) of every irbfn_*(...) call in the C++ source."""
    out = []
    for m in re.finditer(r"\b(irbfn_[a-z_0-9]+)\s*\(", src):
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(src[i], 0)
            i += 1
        out.append((m.group(1), len(_split_args(src[m.end():i - 1]))))
    return out


def test_every_c_call_matches_the_header():
    decl = _declared()
    calls = _calls(_strip_comments(open(FFI_CC).read()))
    assert len(calls) >= 14
    for name, n in calls:
        assert name in decl, "%s is not declared in include/irbfn_b200.h" % name
        assert decl[name] == n, "%s: header takes %d arguments, irbfn_ffi.cc passes %d" % (name, decl[name], n)
    used = {n for n, _ in calls}
    for needed in ("irbfn_forward", "irbfn_forward_gate", "irbfn_forward_sparse", "irbfn_backward", "irbfn_softmax_gate",
                   "irbfn_jacobian", "irbfn_dynamics_loss", "irbfn_prediction_loss", "irbfn_integrate"):
        assert needed in used, needed


def test_registered_handlers_are_all_defined_and_none_is_elided():
    cc = open(FFI_CC).read()
    defined = set(re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(\s*(\w+)\s*,", _strip_comments(cc)))
    py = open(FFI_PY).read()
    block = re.search(r"for _name in \((.*?)\):", py, flags=re.S).group(1)
    registered = set(re.findall(r'"(\w+)"', block))
    assert registered == defined, (registered ^ defined)
    called = set(re.findall(r'ffi_call\(\s*"(\w+)"', py))
    assert called <= defined and {"IrbfnFwd", "IrbfnBwd", "IrbfnFwdHidden", "IrbfnBwdHidden", "IrbfnFwdExtGate",
                                  "IrbfnBwdExtGate", "IrbfnJacobian", "IrbfnDynamicsLoss"} <= called
    # number of Args + Rets bound == number of buffer parameters of the implementation, per handler
    for name, impl in re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(\s*(\w+)\s*,\s*(\w+)\s*,", _strip_comments(cc)):
        bind = re.search(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(\s*%s\s*,.*?\)\s*;" % name, _strip_comments(cc), flags=re.S).group(0)
        n_bound = len(re.findall(r"\.Arg<", bind)) + len(re.findall(r"\.Ret<", bind))
        n_attr = len(re.findall(r"\.Attr<", bind))
        sig = re.search(r"static ffi::Error %s\((.*?)\)\s*{" % impl, _strip_comments(cc), flags=re.S).group(1)
        params = _split_args(sig)
        n_buf = sum(1 for p in params if re.match(r"(F32|S32|RF32|RU8)\b", p))
        assert n_bound == n_buf, (name, n_bound, n_buf)
        assert len(params) == 1 + n_buf + n_attr, (name, params)          # stream + buffers + attributes


def test_flax_modules_reproduce_the_reference_scopes_and_are_not_imported_by_the_product():
    py = open(FFI_PY).read()
    for cls in ("class WCRBFNet(nn.Module)", "class DeeperWCRBFNet(nn.Module)", "class ClusterWCRBFNet(nn.Module)"):
        assert cls in py
    # the scope names the reference's setup() assigns (model.py:161-167, 254-256, 393)
    for scope in ("self.rbf_list = ", "self.linear = ", "self.linear_pre1 = ", "self.linear_pre2 = ", "self.cluster = "):
        assert scope in py, scope
    for leaf in ('"centers"', '"log_sigs"', '"kernel"', '"bias"'):
        assert leaf in py, leaf
    pkg = os.path.join(ROOT, "irbfn_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "irbfn_flax" not in open(os.path.join(pkg, fn)).read() or fn == "build.py", fn
