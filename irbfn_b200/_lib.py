"""ctypes binding of libirbfn_b200.so -- the C ABI declared in include/irbfn_b200.h.

There is no CPU fallback: if the shared library is missing the import of any compute entry
point raises, and every non-zero return code becomes an ``IrbfnError``.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libirbfn_b200.so")

BASIS_GAUSSIAN = 0
BASIS_INVERSE_QUADRATIC = 1
MAX_D = 8
MAX_S = 8
PATH_TILED, PATH_SPLIT, PATH_SMALL = 0, 1, 2      # irbfn_forward_path

GRAD_CENTERS = 0x01
GRAD_LOG_SIGS = 0x02
GRAD_KERNEL = 0x04
GRAD_X = 0x08
GRAD_GATE = 0x10
GRAD_PARAMS = GRAD_CENTERS | GRAD_LOG_SIGS | GRAD_KERNEL

ABI_VERSION = 1

LOSS_FRENET_ONEINT, LOSS_FRENET_FULLINT, LOSS_ST_ONEINT, LOSS_ST_FULLINT = 0, 1, 2, 3
DYN_FRENET_MULT, DYN_FRENET_ONESTEP_AUX, DYN_ST_MULT, DYN_ST_ONESTEP_AUX = 0, 1, 2, 3

# every symbol include/irbfn_b200.h declares (tests check the .so exports exactly these)
EXPORTED_SYMBOLS = (
    "irbfn_abi_version", "irbfn_error_string", "irbfn_last_error_detail", "irbfn_forward_workspace_bytes", "irbfn_forward",
    "irbfn_forward_gate", "irbfn_forward_streamed", "irbfn_upload_streamed", "irbfn_upload_streamed_available", "irbfn_forward_noalloc", "irbfn_forward_sparse_workspace_bytes", "irbfn_forward_sparse", "irbfn_forward_sparse_gate",
    "irbfn_backward_workspace_bytes", "irbfn_backward", "irbfn_backward_sparse", "irbfn_region_gate",
    "irbfn_softmax_gate", "irbfn_forward_plan",
    "irbfn_dynamics_loss_workspace_bytes", "irbfn_dynamics_loss", "irbfn_integrate", "irbfn_jacobian", "irbfn_jacobian_sparse_workspace_bytes", "irbfn_jacobian_sparse",
    "irbfn_gather_rows", "irbfn_lattice_queries", "irbfn_grid_index", "irbfn_grid_index_f64", "irbfn_forward_path", "irbfn_nearest_row_workspace_bytes", "irbfn_nearest_row",
    "irbfn_prediction_loss_workspace_bytes", "irbfn_prediction_loss",
    "irbfn_clip_adam_workspace_bytes", "irbfn_clip_adam",
    "irbfn_fused_update_max_n", "irbfn_fused_update_buffer_bytes", "irbfn_fused_update",
)


class IrbfnError(RuntimeError):
    def __init__(self, code: int, where: str, text: str):
        super().__init__("%s failed: %s (code %d)" % (where, text, code))
        self.code = code


class GateDesc(C.Structure):
    """struct IrbfnGateDesc (include/irbfn_b200.h)."""
    _fields_ = [
        ("num_split", C.c_int32), ("num_ranges", C.c_int32), ("num_bounds", C.c_int32),
        ("lower", C.c_void_p), ("upper", C.c_void_p), ("offset", C.c_void_p),
        ("delta", C.c_void_p), ("ranges", C.c_void_p),
    ]


_lib = None


def load():
    """dlopen the library (once) and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libirbfn_b200.so is not built (%s). Run `python -m irbfn_b200.build`; "
            "there is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    p, i, i64, u, sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint, C.c_size_t
    lib.irbfn_abi_version.restype = i
    lib.irbfn_abi_version.argtypes = []
    lib.irbfn_error_string.restype = C.c_char_p
    lib.irbfn_error_string.argtypes = [i]
    lib.irbfn_last_error_detail.restype = C.c_char_p
    lib.irbfn_last_error_detail.argtypes = []
    lib.irbfn_forward_workspace_bytes.restype = sz
    lib.irbfn_forward_workspace_bytes.argtypes = [i64, i, i, i, i, i]
    lib.irbfn_forward.restype = i
    lib.irbfn_forward.argtypes = [p, p, p, p, p, p, C.POINTER(GateDesc), p, p, i64, i, i, i, i, i,
                                  p, sz, p]
    lib.irbfn_forward_streamed.restype = i
    lib.irbfn_forward_streamed.argtypes = [p, p, p, p, p, p, C.POINTER(GateDesc), p, i64, i, i, i, i, i, p, i64, p, p]
    lib.irbfn_upload_streamed_available.restype = i
    lib.irbfn_upload_streamed_available.argtypes = []
    lib.irbfn_upload_streamed.restype = i
    lib.irbfn_upload_streamed.argtypes = [p, p, i64, i, p, i64, C.POINTER(C.c_void_p), i]
    lib.irbfn_forward_noalloc.restype = i
    lib.irbfn_forward_noalloc.argtypes = [p, p, p, p, p, p, C.POINTER(GateDesc), p, p, i64, i, i, i, i, i, p]
    lib.irbfn_forward_gate.restype = i
    lib.irbfn_forward_gate.argtypes = [p, p, p, p, p, C.POINTER(GateDesc), p, p, i64, i, i, i, i, i, p, sz, p]
    lib.irbfn_forward_sparse_workspace_bytes.restype = sz
    lib.irbfn_forward_sparse_workspace_bytes.argtypes = [i64, i]
    lib.irbfn_forward_sparse.restype = i
    lib.irbfn_forward_sparse.argtypes = [p, p, p, p, p, C.POINTER(GateDesc), p, i64, i, i, i, i, i, C.c_float, p, p, sz, p]
    lib.irbfn_forward_sparse_gate.restype = i
    lib.irbfn_forward_sparse_gate.argtypes = lib.irbfn_forward_sparse.argtypes
    lib.irbfn_backward_workspace_bytes.restype = sz
    lib.irbfn_backward_workspace_bytes.argtypes = [i64, i, i, i, i, u]
    lib.irbfn_backward.restype = i
    lib.irbfn_backward.argtypes = [p, p, p, p, p, C.POINTER(GateDesc), p, p, p, p, p, p, p, p,
                                   i64, i, i, i, i, i, u, p, sz, p]
    lib.irbfn_backward_sparse.restype = i
    lib.irbfn_backward_sparse.argtypes = [p, p, p, p, p, C.POINTER(GateDesc), p, p, p, p, p, p, p, p,
                                          i64, i, i, i, i, i, u, C.c_float, p, sz, p]
    lib.irbfn_region_gate.restype = i
    lib.irbfn_region_gate.argtypes = [p, C.POINTER(GateDesc), p, i64, i, i, p]
    lib.irbfn_softmax_gate.restype = i
    lib.irbfn_softmax_gate.argtypes = [p, p, p, p, p, i64, i, i, p]
    lib.irbfn_forward_plan.restype = i
    lib.irbfn_forward_plan.argtypes = [i, i, i, i, i, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(sz)]
    lib.irbfn_dynamics_loss_workspace_bytes.restype = sz
    lib.irbfn_dynamics_loss_workspace_bytes.argtypes = [i64]
    lib.irbfn_prediction_loss_workspace_bytes.restype = sz
    lib.irbfn_prediction_loss_workspace_bytes.argtypes = [i64, i]
    lib.irbfn_prediction_loss.restype = i
    lib.irbfn_prediction_loss.argtypes = [i, p, p, p, p, i64, i, i64, p, sz, p]
    lib.irbfn_clip_adam_workspace_bytes.restype = sz
    lib.irbfn_clip_adam_workspace_bytes.argtypes = [i64]
    lib.irbfn_clip_adam.restype = i
    lib.irbfn_clip_adam.argtypes = [p, p, p, p, i64, p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                    p, p, sz, p]
    lib.irbfn_fused_update_max_n.restype = i64
    lib.irbfn_fused_update_max_n.argtypes = []
    lib.irbfn_fused_update_buffer_bytes.restype = sz
    lib.irbfn_fused_update_buffer_bytes.argtypes = [i64, i]
    lib.irbfn_fused_update.restype = i
    lib.irbfn_fused_update.argtypes = [p, p, p, p, i64, i, p, i, i, i, p, C.c_double, C.c_double, C.c_double,
                                       C.c_double, C.c_double, p, p, p]
    lib.irbfn_dynamics_loss.restype = i
    lib.irbfn_dynamics_loss.argtypes = [i, p, p, p, C.POINTER(C.c_float), p, p, i64, i, i, i64, p, sz, p]
    lib.irbfn_integrate.restype = i
    lib.irbfn_integrate.argtypes = [i, p, C.POINTER(C.c_float), p, i64, i, p]
    lib.irbfn_jacobian_sparse_workspace_bytes.restype = sz
    lib.irbfn_jacobian_sparse_workspace_bytes.argtypes = [i64, i]
    lib.irbfn_jacobian_sparse.restype = i
    lib.irbfn_jacobian_sparse.argtypes = [p, p, p, p, p, C.POINTER(GateDesc), p, p, i64, i, i, i, i, i, C.c_float, p, sz, p]
    lib.irbfn_jacobian.restype = i
    lib.irbfn_jacobian.argtypes = [p, p, p, p, p, p, C.POINTER(GateDesc), p, p, p, i64, i, i, i, i, i, p]
    lib.irbfn_gather_rows.restype = i
    lib.irbfn_gather_rows.argtypes = [p, p, p, p, p, i64, i64, i, i, p, p]
    lib.irbfn_lattice_queries.restype = i
    lib.irbfn_lattice_queries.argtypes = [p, p, p, i, p, p, i64, i, p]
    lib.irbfn_grid_index.restype = i
    lib.irbfn_grid_index.argtypes = [p, p, p, p, i64, i, p]
    lib.irbfn_forward_path.restype = i
    lib.irbfn_forward_path.argtypes = [i64, i, i, i, i, i]
    lib.irbfn_grid_index_f64.restype = i
    lib.irbfn_grid_index_f64.argtypes = [p, p, p, p, i64, i, p]
    lib.irbfn_nearest_row_workspace_bytes.restype = sz
    lib.irbfn_nearest_row_workspace_bytes.argtypes = [i64, i64]
    lib.irbfn_nearest_row.restype = i
    lib.irbfn_nearest_row.argtypes = [p, p, p, p, i64, i64, i, p, sz, p]
    if lib.irbfn_abi_version() != ABI_VERSION:
        raise ImportError("libirbfn_b200.so ABI %d != binding ABI %d"
                          % (lib.irbfn_abi_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(code: int, where: str) -> None:
    if code != 0:
        text = load().irbfn_error_string(code).decode()
        if code == -7:
            text += " [" + load().irbfn_last_error_detail().decode() + "]"
        raise IrbfnError(code, where, text)
