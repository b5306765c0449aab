"""Radial basis functions, as callables with the names the reference's configs ``eval``
(``basis_func: gaussian`` ... at scripts/train_nmpc.py:207, irbfn_planner.py:72).

The module classes map a callable to a kernel id by ``__name__`` (SURVEY.md section 5,
constraint 1).  The kernels implement the two functions the shipped runs use (flax_rbf.py:34-37,
49-52); ``gaussian_wide`` / ``gaussian_wider`` (flax_rbf.py:40-47: exp(-0.1 a^2), exp(-0.01 a^2))
are the gaussian kernel with the width shifted -- exp(-c a^2) with a^2 = d^2 e^{-2 log_sig} equals
exp(-d^2 e^{-2 (log_sig - ln(c)/2)}) -- so they reuse it with ``log_sig_shift`` added to the
log-widths at the call boundary (gradients w.r.t. log_sig are unchanged by a constant shift).
Any other name raises, in particular ``gaussian_narrow`` / ``gaussian_narrower`` whose
definitions exist nowhere in the reference: they are not guessed.
The callables themselves evaluate the formula on any array type that supports ``**`` and
``exp`` via its namespace (torch or numpy); they are used for documentation and host checks,
never by the CUDA path.
"""
from __future__ import annotations

from ._lib import BASIS_GAUSSIAN, BASIS_INVERSE_QUADRATIC


def _exp(a):
    import numpy as np
    try:
        import torch
        if isinstance(a, torch.Tensor):
            return torch.exp(a)
    except ImportError:  # pragma: no cover
        pass
    return np.exp(a)


def gaussian(alpha):
    """phi = exp(-alpha^2)   (flax_rbf.py:34-37)"""
    return _exp(-1 * alpha ** 2)


def gaussian_wide(alpha):
    """phi = exp(-0.1 alpha^2)   (flax_rbf.py:40-42)"""
    return _exp(-0.1 * alpha ** 2)


def gaussian_wider(alpha):
    """phi = exp(-0.01 alpha^2)   (flax_rbf.py:45-47)"""
    return _exp(-0.01 * alpha ** 2)


def inverse_quadratic(alpha):
    """phi = 1 / (1 + alpha^2)   (flax_rbf.py:49-52)"""
    return 1 / (1 + alpha ** 2)


_IDS = {"gaussian": BASIS_GAUSSIAN, "inverse_quadratic": BASIS_INVERSE_QUADRATIC,
        "gaussian_wide": BASIS_GAUSSIAN, "gaussian_wider": BASIS_GAUSSIAN}
# exp(-c a^2): log_sig' = log_sig - ln(c) / 2
_LOG_SIG_SHIFT = {"gaussian_wide": 0.5 * 2.302585092994046, "gaussian_wider": 0.5 * 4.605170185988092}


def log_sig_shift(basis_func) -> float:
    """Constant added to log_sigs so that the kernel's basis evaluates ``basis_func`` (0 for the native two)."""
    name = basis_func if isinstance(basis_func, str) else getattr(basis_func, "__name__", None)
    return _LOG_SIG_SHIFT.get(name, 0.0)


def basis_id(basis_func) -> int:
    """Kernel id of a basis callable (by ``__name__``) or name string."""
    name = basis_func if isinstance(basis_func, str) else getattr(basis_func, "__name__", None)
    if name not in _IDS:
        raise NotImplementedError(
            "basis function %r has no sm_100a kernel; supported: %s" % (name, sorted(_IDS)))
    return _IDS[name]
