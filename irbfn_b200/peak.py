"""FP32-pipe / XU issue-rate micro-benchmark (bench_csrc/fp32_peak.cu -> libirbfn_bench.so).

MEASUREMENT TOOLING, not part of the product ABI: ``include/irbfn_b200.h`` and ``libirbfn_b200.so`` do
not know about it.  ``bench.py`` uses it for the roofline denominator (MEASURED_PEAKS.json has no FP32
number); ``scripts/dev_peak.py`` prints all probe modes."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libirbfn_bench.so")
_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -m irbfn_b200.build`" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        i, p = C.c_int, C.c_void_p
        lib.irbfn_fp32_microbench.restype = i
        lib.irbfn_fp32_microbench.argtypes = [i, i, i, p, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                              C.POINTER(C.c_float), p]
        lib.irbfn_pipe_microbench.restype = i
        lib.irbfn_pipe_microbench.argtypes = [i, i, i, p, C.POINTER(C.c_double), C.POINTER(C.c_float), p]
        _lib = lib
    return _lib


def _scratch(blocks, device):
    if blocks is None:
        blocks = torch.cuda.get_device_properties(device).multi_processor_count * 8
    return blocks, torch.empty(blocks * 512, device=device, dtype=torch.float32)


def fp32_microbench(mode: int, iters: int = 20000, blocks: Optional[int] = None, device=None):
    """-> (fp32 lane-ops/s, MUFU lane-ops/s, elapsed ms).  mode 0 FFMA, 1 FFMA2, 2 FADD2+FFMA2, 3 MUFU.EX2."""
    lib = _load()
    device = torch.device(device if device is not None else "cuda")
    with torch.cuda.device(device):
        blocks, scratch = _scratch(blocks, device)
        f, m, ms = C.c_double(), C.c_double(), C.c_float()
        rc = lib.irbfn_fp32_microbench(mode, iters, blocks, C.c_void_p(scratch.data_ptr()), C.byref(f), C.byref(m),
                                       C.byref(ms), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError("irbfn_fp32_microbench(mode=%d) failed: %d" % (mode, rc))
    return f.value, m.value, ms.value


def pipe_microbench(mode: int, iters: int = 20000, blocks: Optional[int] = None, device=None):
    """All-pipe variant: -> (fp32, MUFU, fp64 lane-ops/s, elapsed ms)."""
    lib = _load()
    device = torch.device(device if device is not None else "cuda")
    with torch.cuda.device(device):
        blocks, scratch = _scratch(blocks, device)
        rates, ms = (C.c_double * 3)(), C.c_float()
        rc = lib.irbfn_pipe_microbench(mode, iters, blocks, C.c_void_p(scratch.data_ptr()), rates, C.byref(ms),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError("irbfn_pipe_microbench(mode=%d) failed: %d" % (mode, rc))
    return rates[0], rates[1], rates[2], ms.value
