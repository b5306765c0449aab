"""Radial basis functions, as callables with the names the reference's configs ``eval``
(``basis_func: gaussian`` ... at scripts/train_nmpc.py:207, irbfn_planner.py:72).

The module classes map a callable to a kernel id by ``__name__`` (SURVEY.md section 5,
constraint 1).  Only the two functions the shipped runs use and the reference defines are
supported by the kernels (flax_rbf.py:34-37, 49-52); any other name raises, in particular
``gaussian_narrow`` / ``gaussian_narrower`` whose definitions exist nowhere in the reference.
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


def inverse_quadratic(alpha):
    """phi = 1 / (1 + alpha^2)   (flax_rbf.py:49-52)"""
    return 1 / (1 + alpha ** 2)


_IDS = {"gaussian": BASIS_GAUSSIAN, "inverse_quadratic": BASIS_INVERSE_QUADRATIC}


def basis_id(basis_func) -> int:
    """Kernel id of a basis callable (by ``__name__``) or name string."""
    name = basis_func if isinstance(basis_func, str) else getattr(basis_func, "__name__", None)
    if name not in _IDS:
        raise NotImplementedError(
            "basis function %r has no sm_100a kernel; supported: %s" % (name, sorted(_IDS)))
    return _IDS[name]
