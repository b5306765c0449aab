"""In-tree build of libirbfn_b200.so (sm_100a only).

``python -m irbfn_b200.build`` compiles every translation unit under ``csrc/`` with nvcc
(in parallel, one process per TU) and links them into ``irbfn_b200/libirbfn_b200.so``.
nvcc cross-compiles without a GPU, so this runs on the CPU build box; the resulting .so
is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libirbfn_b200.so")
# measurement tooling (FP32 / XU peak probe): its own library, NOT part of the product ABI
BENCH_CSRC = os.path.join(HERE, "bench_csrc")
BENCH_LIB = os.path.join(HERE, "libirbfn_bench.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]

# (source file, extra defines, object name)
FEATURE_COUNTS = (4, 5, 6, 7, 8)


def _units():
    units = [("irbfn_api.cu", [], "irbfn_api.o"), ("irbfn_dynamics.cu", [], "irbfn_dynamics.o"),
             ("irbfn_table.cu", [], "irbfn_table.o"), ("irbfn_optim.cu", [], "irbfn_optim.o")]
    for name in ("irbfn_fwd_inst.cu", "irbfn_fwdm_inst.cu", "irbfn_bwd_x_inst.cu", "irbfn_bwd_xs_inst.cu", "irbfn_bwd_p_inst.cu", "irbfn_bwd_ps_inst.cu", "irbfn_bwd_pw_inst.cu",
                 "irbfn_jac_inst.cu"):
        if not os.path.exists(os.path.join(CSRC, name)):
            continue
        for d in FEATURE_COUNTS:
            units.append((name, ["-DIRBFN_D=%d" % d], "%s_d%d.o" % (name[:-3], d)))
    for name in ("irbfn_bwd_api.cu", "irbfn_jac_api.cu", "irbfn_sparse.cu", "irbfn_fwd_small.cu"):
        if os.path.exists(os.path.join(CSRC, name)):
            units.append((name, [], name[:-3] + ".o"))
    return units


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libirbfn_b200.so cannot be built")


def _source_digest():
    h = hashlib.sha256()
    for root in (CSRC, BENCH_CSRC, os.path.join(HERE, "..", "include")):
        for fn in sorted(os.listdir(root)):
            if fn.endswith((".cu", ".cuh", ".h")):
                with open(os.path.join(root, fn), "rb") as f:
                    h.update(fn.encode())
                    h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Build the shared library if sources changed. Returns its path."""
    stamp = os.path.join(OBJ, "digest.txt")
    digest = _source_digest()
    if not force and os.path.exists(LIB) and os.path.exists(BENCH_LIB) and os.path.exists(stamp):
        with open(stamp) as f:
            if f.read().strip() == digest:
                return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    units = _units()

    def compile_one(u):
        src, defs, obj = u
        cmd = [nvcc, *NVCC_FLAGS, *defs, "-c", os.path.join(CSRC, src), "-o", os.path.join(OBJ, obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        with open(os.path.join(OBJ, obj + ".log"), "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s %s:\n%s" % (src, defs, r.stderr[-4000:]))
        return obj

    with ThreadPoolExecutor(max_workers=min(len(units), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(compile_one, units))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB,
           *[os.path.join(OBJ, o) for o in objs]]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr[-4000:])
    cmd = [nvcc, *NVCC_FLAGS, "-shared", os.path.join(BENCH_CSRC, "fp32_peak.cu"), "-o", BENCH_LIB]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for the peak probe:\n" + r.stderr[-4000:])
    build_ffi(verbose)
    with open(stamp, "w") as f:
        f.write(digest)
    if verbose:
        print("built", LIB)
    return LIB


def build_ffi(verbose: bool = False):
    """The XLA FFI shim (ffi/irbfn_ffi.cc -> ffi/libirbfn_ffi.so) for the reference's JAX side.  It needs the XLA
    FFI headers that ship with jaxlib; without JAX (this container, the GPU box) it is skipped -- the product and
    its tests do not use it (irbfn_flax.py is the only consumer)."""
    try:
        import jax.ffi as jffi       # noqa: F401
        inc = jffi.include_dir()
    except Exception:
        if verbose:
            print("jax not importable: ffi/irbfn_ffi.cc not built (unverified source, see INTEGRATION.md)")
        return None
    out = os.path.join(HERE, "ffi", "libirbfn_ffi.so")
    cmd = [_nvcc(), "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-I", inc, "-I", os.path.join(HERE, "..", "include"),
           os.path.join(HERE, "ffi", "irbfn_ffi.cc"), "-L", HERE, "-lirbfn_b200", "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("irbfn_ffi.cc failed to build:\n" + r.stderr[-4000:])
    return out


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
