"""CPU oracle for the IRBFN hot path (NumPy).  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the arithmetic of the reference's batched IRBFN
evaluation.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.
The product path (``irbfn_b200``) never does.

PARITY PINNED BY EXECUTING THE REFERENCE: the reference ships no golden vectors or known-answer
tests for this path and JAX/Flax/flax_rbf are not installable here, so the
reference's own source files (src/irbfn_mpc/model.py, the flax_rbf snapshot) are imported UNMODIFIED
under the torch-CPU library stand-ins of ``oracle/refshim`` and run on the shipped checkpoints
(tests/golden/make_golden.py).  Every committed fixture value is what that code returned; this oracle
reproduces all of them to 1e-12 (outputs) / 1e-9 (gradients) -- tests/test_oracle_vs_ref.py, which also
re-executes the reference live wherever /root/reference exists.  Gradients are autodiff of the
reference's executed forward (the reference has no gradient code of its own, F7).

Reference lines followed (paths relative to the reference checkout):
  gate      : src/irbfn_mpc/model.py:70-95   (_region_activation)
  RBF layer : deprecated/f1tenth_gym/examples/flax_rbf/flax_rbf/flax_rbf.py:275-283
  basis     : flax_rbf.py:34-37 (gaussian), :49-52 (inverse_quadratic)
  blend     : src/irbfn_mpc/model.py:187-193
  readout   : src/irbfn_mpc/model.py:196  (nn.Dense: y = h @ kernel + bias)
  soft gate : src/irbfn_mpc/model.py:402-409 (ClusterWCRBFNet softmax gate)
  deep tail : src/irbfn_mpc/model.py:285-287 (DeeperWCRBFNet)

The arithmetic type is chosen by the caller (``dtype=np.float64`` for the
high-precision reference, ``np.float32`` for a like-for-like comparison).
"""
from __future__ import annotations

import numpy as np

GAUSSIAN = 0
INVERSE_QUADRATIC = 1
BASIS_IDS = {"gaussian": GAUSSIAN, "inverse_quadratic": INVERSE_QUADRATIC}


# --------------------------------------------------------------------------
# gate  (model.py:70-95)
# --------------------------------------------------------------------------
def region_gate_factors(x, lower_bounds, upper_bounds, delta, dtype=np.float64):
    """Per split dimension d: g_d[b, i] (model.py:74-86).

    Returns a list of S arrays of shape (B, n_d).
    """
    x = np.asarray(x, dtype=dtype)
    out = []
    for d in range(len(lower_bounds)):
        lb = np.asarray(lower_bounds[d], dtype=dtype)[None, :]
        ub = np.asarray(upper_bounds[d], dtype=dtype)[None, :]
        dl = dtype(delta[d])
        lower_diffs = x[:, d : d + 1] - lb            # model.py:75-77
        upper_diffs = ub - x[:, d : d + 1]            # model.py:78-81
        g = ((np.tanh(dl * lower_diffs) + 1) / 2) * ((np.tanh(dl * upper_diffs) + 1) / 2)
        out.append(g.astype(dtype))
    return out


def region_activation(x, num_regions, lower_bounds, upper_bounds, delta,
                      dimension_ranges, dtype=np.float64):
    """gamma[b, r] = prod_j g_j[b, ranges[r][j]]  (model.py:88-93).  Not normalised."""
    x = np.asarray(x, dtype=dtype)
    S = len(lower_bounds)
    if S > x.shape[1]:
        raise IndexError("more split dimensions than input features")
    gs = region_gate_factors(x, lower_bounds, upper_bounds, delta, dtype)
    gamma = np.zeros((x.shape[0], num_regions), dtype=dtype)
    for i in range(min(len(dimension_ranges), num_regions)):
        cur = gs[0][:, int(dimension_ranges[i][0])]
        for j in range(1, S):
            cur = cur * gs[j][:, int(dimension_ranges[i][j])]
        gamma[:, i] = cur
    return gamma


def softmax_gate(x, cluster_kernel, cluster_bias, dtype=np.float64):
    """ClusterWCRBFNet gate (model.py:402-406): softmax(x @ Wc + bc). Returns (gate, logits)."""
    x = np.asarray(x, dtype=dtype)
    logits = x @ np.asarray(cluster_kernel, dtype=dtype) + np.asarray(cluster_bias, dtype=dtype)
    z = logits - logits.max(axis=-1, keepdims=True)
    e = np.exp(z)
    return (e / e.sum(axis=-1, keepdims=True)).astype(dtype), logits.astype(dtype)


# --------------------------------------------------------------------------
# RBF layer + blend + readout
# --------------------------------------------------------------------------
def _basis(alpha, basis):
    if basis == GAUSSIAN:
        return np.exp(-1 * alpha ** 2)                # flax_rbf.py:36
    if basis == INVERSE_QUADRATIC:
        return np.ones_like(alpha) / (np.ones_like(alpha) + alpha ** 2)  # flax_rbf.py:51
    raise ValueError("unsupported basis id %r" % (basis,))


def rbf_activations(x, centers, log_sigs, basis, dtype=np.float64):
    """phi[b, r, k] following flax_rbf.py:275-283 literally (sqrt, divide, basis)."""
    x = np.asarray(x, dtype=dtype)
    c = np.asarray(centers, dtype=dtype)
    ls = np.asarray(log_sigs, dtype=dtype)
    diff = x[:, None, None, :] - c[None]              # (B,R,K,D)
    d = (np.sum(diff ** 2, axis=-1) ** 0.5) / np.exp(ls)[None]
    return _basis(d, basis).astype(dtype)


def blend(gamma, phi):
    """h[b,k] = sum_r gamma[b,r] phi[b,r,k]  (model.py:187-193)."""
    return np.sum(gamma[:, :, None] * phi, axis=1)


def forward(x, centers, log_sigs, kernel, bias, gamma, basis, dtype=np.float64, chunk=1024,
            return_hidden=False):
    """y = (sum_r gamma_r phi_r) @ kernel + bias, in query chunks to bound memory."""
    x = np.asarray(x, dtype=dtype)
    W = np.asarray(kernel, dtype=dtype)
    b = np.asarray(bias, dtype=dtype)
    B = x.shape[0]
    y = np.empty((B, W.shape[1]), dtype=dtype)
    hs = np.empty((B, W.shape[0]), dtype=dtype) if return_hidden else None
    for s in range(0, B, chunk):
        e = min(B, s + chunk)
        phi = rbf_activations(x[s:e], centers, log_sigs, basis, dtype)
        h = blend(np.asarray(gamma[s:e], dtype=dtype), phi)
        if return_hidden:
            hs[s:e] = h
        y[s:e] = h @ W + b                           # model.py:196
    return (y, hs) if return_hidden else y


def wcrbf_forward(x, params, cfg, dtype=np.float64, chunk=1024):
    """WCRBFNet.__call__ (model.py:169-198).  ``params`` = the 'params' sub-tree,
    ``cfg`` = dict with the module attributes of model.py:113-125."""
    gamma = region_activation(x, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"],
                              cfg["delta"], cfg["dimension_ranges"], dtype)
    return forward(x, params["rbf_list"]["centers"], params["rbf_list"]["log_sigs"],
                   params["linear"]["kernel"], params["linear"]["bias"], gamma,
                   BASIS_IDS[cfg["basis_func"]], dtype, chunk)


def deeper_forward(x, params, cfg, dtype=np.float64, chunk=1024):
    """DeeperWCRBFNet.__call__ (model.py:259-289)."""
    gamma = region_activation(x, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"],
                              cfg["delta"], cfg["dimension_ranges"], dtype)
    K = np.asarray(params["rbf_list"]["centers"]).shape[1]
    eye = np.eye(K, dtype=dtype)
    h = forward(x, params["rbf_list"]["centers"], params["rbf_list"]["log_sigs"], eye,
                np.zeros(K, dtype), gamma, BASIS_IDS[cfg["basis_func"]], dtype, chunk)
    p1 = h @ np.asarray(params["linear_pre1"]["kernel"], dtype) + np.asarray(params["linear_pre1"]["bias"], dtype)
    p2 = np.maximum(p1, 0) @ np.asarray(params["linear_pre2"]["kernel"], dtype) + np.asarray(params["linear_pre2"]["bias"], dtype)
    return np.maximum(p2, 0) @ np.asarray(params["linear"]["kernel"], dtype) + np.asarray(params["linear"]["bias"], dtype)


def cluster_forward(x, params, cfg, dtype=np.float64, chunk=1024):
    """ClusterWCRBFNet.__call__ (model.py:395-414). Returns (out, logits)."""
    gate, logits = softmax_gate(x, params["cluster"]["kernel"], params["cluster"]["bias"], dtype)
    y = forward(x, params["rbf_list"]["centers"], params["rbf_list"]["log_sigs"],
                params["linear"]["kernel"], params["linear"]["bias"], gate,
                BASIS_IDS[cfg["basis_func"]], dtype, chunk)
    return y, logits


# --------------------------------------------------------------------------
# analytic VJP (what jax.value_and_grad of the forward definition returns, written on
# d^2 so that x == centre gives the finite limit)
# --------------------------------------------------------------------------
def gate_vjp_x(x, dgamma, num_regions, lower_bounds, upper_bounds, delta, dimension_ranges,
               dtype=np.float64):
    """dx[b, j] = sum_r dgamma[b,r] * d gamma_r / d x_j  for the tanh gate."""
    x = np.asarray(x, dtype=dtype)
    B, D = x.shape
    S = len(lower_bounds)
    gs, dgs = [], []
    for d in range(S):
        lb = np.asarray(lower_bounds[d], dtype=dtype)[None, :]
        ub = np.asarray(upper_bounds[d], dtype=dtype)[None, :]
        dl = dtype(delta[d])
        tl = np.tanh(dl * (x[:, d : d + 1] - lb))
        tu = np.tanh(dl * (ub - x[:, d : d + 1]))
        sl, su = (tl + 1) / 2, (tu + 1) / 2
        gs.append(sl * su)
        dgs.append(dl * (1 - tl ** 2) / 2 * su - sl * dl * (1 - tu ** 2) / 2)
    dx = np.zeros((B, D), dtype=dtype)
    for i in range(min(len(dimension_ranges), num_regions)):
        idx = [int(v) for v in dimension_ranges[i]]
        for j in range(S):
            prod = dgs[j][:, idx[j]]
            for jj in range(S):
                if jj != j:
                    prod = prod * gs[jj][:, idx[jj]]
            dx[:, j] += dgamma[:, i] * prod
    return dx


def backward(x, centers, log_sigs, kernel, bias, gamma, dy, basis, dtype=np.float64, chunk=512):
    """VJP of ``forward`` for cotangent dy (B,O).

    Returns dict: centers (R,K,D), log_sigs (R,K), kernel (K,O), bias (O,),
    x (B,D) [through the RBF layers only: the gate term is separate, see
    gate_vjp_x], gamma (B,R).
    """
    x = np.asarray(x, dtype=dtype)
    c = np.asarray(centers, dtype=dtype)
    ls = np.asarray(log_sigs, dtype=dtype)
    W = np.asarray(kernel, dtype=dtype)
    gamma = np.asarray(gamma, dtype=dtype)
    dy = np.asarray(dy, dtype=dtype)
    B, D = x.shape
    R, K, _ = c.shape
    e = np.exp(-2 * ls)                               # 1/sigma^2
    dC = np.zeros_like(c)
    dLS = np.zeros_like(ls)
    dW = np.zeros_like(W)
    dx = np.zeros_like(x)
    dgamma = np.zeros_like(gamma)
    for s in range(0, B, chunk):
        t = min(B, s + chunk)
        diff = x[s:t, None, None, :] - c[None]        # (b,R,K,D)
        d2 = np.sum(diff ** 2, axis=-1)               # (b,R,K)
        a2 = d2 * e[None]
        if basis == GAUSSIAN:
            phi = np.exp(-a2)
            dphi_da2 = -phi
        elif basis == INVERSE_QUADRATIC:
            phi = 1 / (1 + a2)
            dphi_da2 = -phi * phi
        else:
            raise ValueError("unsupported basis id %r" % (basis,))
        h = np.sum(gamma[s:t, :, None] * phi, axis=1)
        dW += h.T @ dy[s:t]
        dh = dy[s:t] @ W.T                            # (b,K)
        dphi = dh[:, None, :] * gamma[s:t, :, None]   # dL/dphi
        dgamma[s:t] = np.sum(dh[:, None, :] * phi, axis=-1)
        da2 = dphi * dphi_da2                         # dL/d(alpha^2)
        dd2 = da2 * e[None]
        dC += np.sum(-2 * dd2[..., None] * diff, axis=0)
        dLS += np.sum(da2 * (-2) * a2, axis=0)
        dx[s:t] = np.sum(2 * dd2[..., None] * diff, axis=(1, 2))
    return {"centers": dC, "log_sigs": dLS, "kernel": dW, "bias": dy.sum(axis=0),
            "x": dx, "gamma": dgamma}


def wcrbf_vjp(x, params, cfg, dy, dtype=np.float64, chunk=512):
    """Full VJP of WCRBFNet.apply w.r.t. params and x (including the gate's x-dependence)."""
    gamma = region_activation(x, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"],
                              cfg["delta"], cfg["dimension_ranges"], dtype)
    g = backward(x, params["rbf_list"]["centers"], params["rbf_list"]["log_sigs"],
                 params["linear"]["kernel"], params["linear"]["bias"], gamma, dy,
                 BASIS_IDS[cfg["basis_func"]], dtype, chunk)
    g["x"] = g["x"] + gate_vjp_x(x, g["gamma"], cfg["num_regions"], cfg["lower_bounds"],
                                 cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dtype)
    return g


# --------------------------------------------------------------------------
# Flax msgpack checkpoint reader (flax.serialization format; ext type 1 = ndarray)
# --------------------------------------------------------------------------
def wcrbf_jacobian(x, params, cfg, dtype=np.float64, chunk=512):
    """J[b,o,j] = d y[b,o] / d x[b,j]: jax.jacrev of WCRBFNet.__call__ (model.py:169-198), built
    the way jacrev does it -- one VJP with a one-hot cotangent per output column."""
    x = np.asarray(x, dtype=dtype)
    O = np.asarray(params["linear"]["kernel"]).shape[1]
    J = np.empty((x.shape[0], O, x.shape[1]), dtype=dtype)
    for o in range(O):
        dy = np.zeros((x.shape[0], O), dtype=dtype)
        dy[:, o] = 1
        J[:, o, :] = wcrbf_vjp(x, params, cfg, dy, dtype, chunk)["x"]
    return J


def load_flax_checkpoint(path):
    """Decode a Flax legacy msgpack checkpoint (saved at scripts/train_nmpc.py:497-502)
    with ``msgpack`` alone.  Returns the raw tree {step, params:{params:{...}}, opt_state}."""
    import msgpack

    def ext_hook(code, data):
        if code == 1:
            shape, dtype, buf = msgpack.unpackb(data, raw=False)
            return np.frombuffer(buf, dtype=np.dtype(dtype)).reshape(shape).copy()
        if code == 3:
            dtype, buf = msgpack.unpackb(data, raw=False)
            return np.frombuffer(buf, dtype=np.dtype(dtype))[0]
        return msgpack.ExtType(code, data)

    with open(path, "rb") as f:
        return msgpack.unpackb(f.read(), ext_hook=ext_hook, raw=False, strict_map_key=False)


def query_box(cfg):
    """Per-dimension [min lower_bounds, max upper_bounds] box of a config."""
    lo = np.array([min(b) for b in cfg["lower_bounds"]], dtype=np.float64)
    hi = np.array([max(b) for b in cfg["upper_bounds"]], dtype=np.float64)
    return lo, hi


def uniform_queries(cfg, n, seed=0):
    """n queries uniform in the config box (NumPy default_rng(seed)); columns beyond the
    split dimensions are drawn from [-1, 1]."""
    lo, hi = query_box(cfg)
    D = int(cfg["in_features"])
    rng = np.random.default_rng(seed)
    u = rng.random((n, D))
    lo_f = np.full(D, -1.0)
    hi_f = np.full(D, 1.0)
    lo_f[: len(lo)] = lo[:D]
    hi_f[: len(hi)] = hi[:D]
    return lo_f + u * (hi_f - lo_f)
