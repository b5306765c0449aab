"""TEST INFRASTRUCTURE -- NumPy float64 restatement of the reference's optimiser update.  Only tests/ may
import this.

The reference builds ``optax.chain(optax.clip_by_global_norm(max_grad_norm), optax.adam(lr))``
(scripts/train_nmpc.py:231-233; scripts/train_nmpc_frenet.py:306-309) and applies it with
``state.apply_gradients(grads=grads)`` (scripts/train_nmpc.py:299).  optax is a dependency of the reference
(pyproject.toml, no pinned version), absent here; the functions below restate its published definitions:

  global_norm            sqrt(sum over leaves of sum(x^2))
  clip_by_global_norm    trigger = g_norm < max_norm; g = select(trigger, g, (g / g_norm) * max_norm)
  scale_by_adam          mu = (1-b1) g + b1 mu; nu = (1-b2) g^2 + b2 nu; count += 1;
                         mu_hat = mu / (1 - b1^count); nu_hat = nu / (1 - b2^count);
                         u = mu_hat / (sqrt(nu_hat + eps_root) + eps), eps_root = 0
  scale(-lr)             u *= -lr;  apply_updates: p += u

Parity unpinned (no optax here to run against); pinned by the definitions above and by a hand-computed
two-step case in tests/test_optim.py.
"""
import numpy as np


def clip_by_global_norm(g, max_norm):
    g = np.asarray(g, np.float64)
    g_norm = np.sqrt(np.sum(g * g))
    if max_norm <= 0 or g_norm < max_norm:
        return g, g_norm
    return (g / g_norm) * max_norm, g_norm


def adam_step(p, g, mu, nu, count, lr, b1=0.9, b2=0.999, eps=1e-8, max_norm=0.0):
    """One update; returns (p, g_clipped, mu, nu, count, g_norm), all float64."""
    p, mu, nu = (np.asarray(a, np.float64) for a in (p, mu, nu))
    g, g_norm = clip_by_global_norm(g, max_norm)
    mu = (1 - b1) * g + b1 * mu
    nu = (1 - b2) * (g * g) + b2 * nu
    count = count + 1
    mu_hat = mu / (1 - b1 ** count)
    nu_hat = nu / (1 - b2 ** count)
    p = p + (-lr) * (mu_hat / (np.sqrt(nu_hat) + eps))
    return p, g, mu, nu, count, g_norm
