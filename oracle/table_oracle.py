"""TEST INFRASTRUCTURE -- NumPy restatement of the reference's table-side operations (SURVEY.md 8f
ranks 3 and 4).  Only tests/ and bench.py's cpu_baseline leg may import this.

  gather            batch_x = train_x[perm, :]                  scripts/train_nmpc_frenet.py:499-501
  remove_infeasible valid_ind = np.unique(np.where(outputs != -999)[0])   scripts/train_nmpc_frenet.py:49-52
  epoch_batches     perms[: steps * bs].reshape(steps, bs)      scripts/train_nmpc_frenet.py:491-497
  nearest_row       jnp.argmin(jnp.linalg.norm(all - q, axis=1))  src/irbfn_mpc/explicit_planner.py:21-25
  grid_index        min(shape - 1, np.searchsorted(keys, v, side="right"))   explicit_planner.py:165-175

Parity: gather / grid_index / remove_infeasible are exact index arithmetic.  nearest_row is evaluated in
float64 here; the device accumulates in float32, so indices are compared where the float64 runner-up is
not within float32 rounding of the winner (tests state the rule).
"""
import numpy as np


def gather(train_x, train_y, perm):
    return train_x[perm, :], train_y[perm, :]


def remove_infeasible(inputs, outputs, marker=-999):
    valid_ind = np.unique(np.where(outputs != marker)[0])
    return inputs[valid_ind], outputs[valid_ind]


def epoch_batches(perm, num_train, bs):
    num_steps = num_train // bs
    return np.asarray(perm)[: num_steps * bs].reshape((num_steps, bs))


def nearest_row(q, table, chunk=256):
    """(index, distance, runner-up distance) per query, float64, first minimum wins."""
    q = np.asarray(q, np.float64)
    t = np.asarray(table, np.float64)
    idx = np.empty(q.shape[0], np.int64)
    d1 = np.empty(q.shape[0])
    d2 = np.empty(q.shape[0])
    for s in range(0, q.shape[0], chunk):
        diff = t[None, :, :] - q[s:s + chunk, None, :]
        n = np.linalg.norm(diff, axis=2)
        i = np.argmin(n, axis=1)
        idx[s:s + chunk] = i
        d1[s:s + chunk] = n[np.arange(len(i)), i]
        if t.shape[0] > 1:
            n[np.arange(len(i)), i] = np.inf
            d2[s:s + chunk] = n.min(axis=1)
        else:
            d2[s:s + chunk] = np.inf
    return idx, d1, d2


def grid_index(q, keys):
    q = np.asarray(q)
    out = np.empty(q.shape, np.int32)
    for d, k in enumerate(keys):
        k = np.asarray(k, q.dtype)
        out[:, d] = np.minimum(len(k) - 1, np.searchsorted(k, q[:, d], side="right"))
    return out


def lattice(axes, fixed):
    """Cell 8 of scripts/irbfn_vs_explicit_frenet.ipynb: meshgrid(indexing="ij") over the chosen columns,
    flattened, the other columns constant."""
    cols = list(axes.keys())
    mesh = np.meshgrid(*[np.asarray(axes[c], np.float64) for c in cols], indexing="ij")
    n = mesh[0].size
    x = np.empty((n, len(fixed)), np.float32)
    x[:] = np.asarray(fixed, np.float32)
    for c, g in zip(cols, mesh):
        x[:, c] = g.flatten()
    return x


def explicit_lookup(queries, inputs, outputs, out_features):
    """Cells 4, 11, 12: inds = argmin_row ||inputs - q||; outputs[inds].reshape((-1, O), order="F")."""
    inds, d1, d2 = nearest_row(queries, inputs)
    return inds, np.asarray(outputs)[inds].reshape((-1, out_features), order="F"), d1, d2
