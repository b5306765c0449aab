"""Flax legacy msgpack checkpoints (what the reference saves with
``flax.training.checkpoints.save_checkpoint`` at scripts/train_nmpc.py:497-502 and restores at
irbfn_planner.py:88-90), read and written with ``msgpack`` alone.

Format: a msgpack map ``{step, params: {params: {...}}, opt_state}`` whose array leaves are
ext type 1 = msgpack((shape, dtype_name, raw_bytes)) and whose NumPy scalars are ext type 3.
"""
from __future__ import annotations

import os
import re

import numpy as np


def _ext_hook(code, data):
    import msgpack
    if code == 1:
        shape, dtype, buf = msgpack.unpackb(data, raw=False)
        return np.frombuffer(buf, dtype=np.dtype(dtype)).reshape(shape).copy()
    if code == 3:
        dtype, buf = msgpack.unpackb(data, raw=False)
        return np.frombuffer(buf, dtype=np.dtype(dtype))[0]
    return msgpack.ExtType(code, data)


def load_checkpoint(path: str) -> dict:
    """Decode one checkpoint file -> the raw tree (NumPy leaves)."""
    import msgpack
    with open(path, "rb") as f:
        return msgpack.unpackb(f.read(), ext_hook=_ext_hook, raw=False, strict_map_key=False)


def latest_checkpoint(ckpt_dir: str, prefix: str = "checkpoint_"):
    """Highest-step ``checkpoint_<n>`` in a directory (Flax's natural ordering), or None."""
    best, best_n = None, -1
    for fn in os.listdir(ckpt_dir):
        m = re.fullmatch(re.escape(prefix) + r"(\d+)", fn)
        if m and int(m.group(1)) > best_n:
            best, best_n = os.path.join(ckpt_dir, fn), int(m.group(1))
    return best


def restore_params(ckpt: str) -> dict:
    """``{'params': {...}}`` of a checkpoint file or directory, leaves as float32 NumPy arrays
    (the reference's inference path runs with x64 disabled, so JAX canonicalises the float64
    leaves of the shipped checkpoints to float32)."""
    path = latest_checkpoint(ckpt) if os.path.isdir(ckpt) else ckpt
    if path is None:
        raise FileNotFoundError("no checkpoint_<n> file in %s" % ckpt)
    tree = load_checkpoint(path)
    params = tree["params"] if "params" in tree else tree

    def cast(node):
        if isinstance(node, dict):
            return {k: cast(v) for k, v in node.items()}
        return np.asarray(node, dtype=np.float32)

    return cast(params)


def _default(obj):
    import msgpack
    if isinstance(obj, np.ndarray):
        return msgpack.ExtType(1, msgpack.packb((list(obj.shape), obj.dtype.name, obj.tobytes()),
                                                use_bin_type=True))
    if isinstance(obj, np.generic):
        return msgpack.ExtType(3, msgpack.packb((obj.dtype.name, obj.tobytes()), use_bin_type=True))
    raise TypeError("cannot serialise %r" % type(obj))


def save_checkpoint(ckpt_dir: str, target: dict, step: int, prefix: str = "checkpoint_") -> str:
    """Write ``target`` (a tree of NumPy arrays / scalars / dicts) as ``checkpoint_<step>``."""
    import msgpack

    def to_np(node):
        if isinstance(node, dict):
            return {k: to_np(v) for k, v in node.items()}
        if hasattr(node, "detach"):
            return node.detach().cpu().numpy()
        return node

    os.makedirs(ckpt_dir, exist_ok=True)
    path = os.path.join(ckpt_dir, "%s%d" % (prefix, step))
    with open(path, "wb") as f:
        f.write(msgpack.packb(to_np(target), default=_default, use_bin_type=True))
    return path
