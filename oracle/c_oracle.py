"""ctypes wrapper + build recipe for oracle/irbfn_oracle.c (float32, OpenMP).

TEST INFRASTRUCTURE ONLY: the checker and the timed CPU baseline.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "irbfn_oracle.c")
LIB = os.path.join(HERE, "_build", "libirbfn_oracle.so")
# AVX2+FMA is the common denominator of the build container and the GPU box hosts
CFLAGS = ["-O3", "-mavx2", "-mfma", "-fopenmp", "-shared", "-fPIC"]


def build(force: bool = False) -> str:
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.run(["gcc", *CFLAGS, SRC, "-o", LIB, "-lm"], check=True)
    return LIB


class _Gate(C.Structure):
    _fields_ = [("S", C.c_int), ("num_ranges", C.c_int), ("num_bounds", C.c_int),
                ("lower", C.c_void_p), ("upper", C.c_void_p), ("offset", C.c_void_p),
                ("delta", C.c_void_p), ("ranges", C.c_void_p)]


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(build())
        p, i, i64 = C.c_void_p, C.c_int, C.c_int64
        lib.irbfn_oracle_num_threads.restype = i
        lib.irbfn_oracle_forward_f32.restype = i
        lib.irbfn_oracle_forward_f32.argtypes = [p, p, p, p, p, p, C.POINTER(_Gate), p, p, i64, i, i, i, i, i, i]
        lib.irbfn_oracle_backward_f32.restype = i
        lib.irbfn_oracle_backward_f32.argtypes = [p, p, p, p, p, C.POINTER(_Gate), p, p, p, p, p, p,
                                                  i64, i, i, i, i, i, i]
        _lib = lib
    return _lib


def num_threads() -> int:
    return int(_load().irbfn_oracle_num_threads())


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class GateArrays:
    """Flat packing of (lower_bounds, upper_bounds, delta, dimension_ranges) for the C oracle."""

    def __init__(self, cfg):
        lbs, ubs = cfg["lower_bounds"], cfg["upper_bounds"]
        S = len(lbs)
        offs = np.zeros(S + 1, np.int32)
        for d in range(S):
            offs[d + 1] = offs[d] + len(lbs[d])
        self.lower = _f32(np.concatenate([np.asarray(b, np.float64) for b in lbs]))
        self.upper = _f32(np.concatenate([np.asarray(b, np.float64) for b in ubs]))
        self.offset = offs
        self.delta = _f32(list(cfg["delta"])[:S])
        rng = np.asarray(cfg["dimension_ranges"], np.int32).reshape(-1, S)
        self.ranges = np.ascontiguousarray(rng)
        self.desc = _Gate(S, rng.shape[0], int(offs[-1]), _ptr(self.lower), _ptr(self.upper),
                          _ptr(self.offset), _ptr(self.delta), _ptr(self.ranges))


def forward(x, centers, log_sigs, kernel, bias, basis, cfg=None, gate=None, want_hidden=False,
            nthreads=0):
    lib = _load()
    x, centers, log_sigs = _f32(x), _f32(centers), _f32(log_sigs)
    B, D = x.shape
    R, K, _ = centers.shape
    y = hidden = None
    O = 0
    if kernel is not None:
        kernel, bias = _f32(kernel), _f32(bias)
        O = kernel.shape[1]
        y = np.empty((B, O), np.float32)
    if want_hidden:
        hidden = np.empty((B, K), np.float32)
    ga = GateArrays(cfg) if gate is None else None
    gate = _f32(gate) if gate is not None else None
    rc = lib.irbfn_oracle_forward_f32(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(bias),
                                      _ptr(gate), C.byref(ga.desc) if ga else None, _ptr(y),
                                      _ptr(hidden), B, R, K, D, O, basis, nthreads)
    if rc != 0:
        raise ValueError("irbfn_oracle_forward_f32 rejected its arguments")
    return (y, hidden) if want_hidden and y is not None else (hidden if want_hidden else y)


def backward(x, centers, log_sigs, kernel, dy, basis, cfg=None, gate=None, want_dx=True, nthreads=0):
    lib = _load()
    x, centers, log_sigs, kernel, dy = _f32(x), _f32(centers), _f32(log_sigs), _f32(kernel), _f32(dy)
    B, D = x.shape
    R, K, _ = centers.shape
    O = kernel.shape[1]
    out = {"centers": np.empty((R, K, D), np.float32), "log_sigs": np.empty((R, K), np.float32),
           "kernel": np.empty((K, O), np.float32), "bias": np.empty((O,), np.float32)}
    if want_dx:
        out["x"] = np.empty((B, D), np.float32)
    ga = GateArrays(cfg) if gate is None else None
    gate = _f32(gate) if gate is not None else None
    rc = lib.irbfn_oracle_backward_f32(_ptr(x), _ptr(centers), _ptr(log_sigs), _ptr(kernel), _ptr(gate),
                                       C.byref(ga.desc) if ga else None, _ptr(dy), _ptr(out["centers"]),
                                       _ptr(out["log_sigs"]), _ptr(out["kernel"]), _ptr(out["bias"]),
                                       _ptr(out.get("x")), B, R, K, D, O, basis, nthreads)
    if rc != 0:
        raise ValueError("irbfn_oracle_backward_f32 rejected its arguments")
    return out
