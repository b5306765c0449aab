"""pytest configuration: the ``gpu`` marker and shared fixture helpers.

``-m "not gpu"`` runs on the CPU build box (oracle vs. golden vectors, host logic, ABI
symbols, gloo world_size-2 sharding); ``-m gpu`` runs the parity tests proper on a B200
through the C ABI.  Nothing here reads /root/reference.
"""
import glob
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_paths():
    """Model fixtures (one per shipped checkpoint).  ``*_ref.npz`` are the dynamics / training-step
    fixtures, which have their own layout."""
    return sorted(p for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")) if not p.endswith("_ref.npz"))


def load_golden(path):
    """-> (cfg dict, params tree (float32 leaves), arrays dict)."""
    z = np.load(path)
    cfg = json.loads(str(z["cfg_json"]))
    params = {}
    for k in z.files:
        if k.startswith("p/"):
            node = params
            parts = k[2:].split("/")
            for p in parts[:-1]:
                node = node.setdefault(p, {})
            node[parts[-1]] = z[k]
    arrays = {k: z[k] for k in z.files if not k.startswith("p/") and k != "cfg_json"}
    return cfg, params, arrays


def rel_max(a, ref):
    """norm-wise output error: max|a - ref| / max|ref|  (SURVEY.md 8c)."""
    a = np.asarray(a, np.float64)
    ref = np.asarray(ref, np.float64)
    return float(np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-300))


def rel_l2(a, ref):
    """per-leaf gradient error: ||a - ref||_2 / ||ref||_2."""
    a = np.asarray(a, np.float64).ravel()
    ref = np.asarray(ref, np.float64).ravel()
    return float(np.linalg.norm(a - ref) / max(np.linalg.norm(ref), 1e-300))


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
