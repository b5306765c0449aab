"""The dynamics / loss restatement (oracle/dynamics_oracle.py) is self-consistent on the CPU:
autograd cotangents agree with central finite differences, the one-step functions agree with the
first step of the rollouts, and documented reference quirks are kept."""
import numpy as np
import torch

from oracle import dynamics_oracle as dyn

P = dyn.default_dyn_params()


def _x(rng, B, D):
    lo = np.array([-0.2, -0.3, 1.0, -1.0, 3.0, -2.6, -1.0, -0.1])[:D]
    hi = np.array([2.0, 0.3, 7.0, 1.0, 7.0, 2.6, 1.0, 0.1])[:D]
    return rng.uniform(lo, hi, (B, D))


def test_cotangent_matches_finite_differences():
    rng = np.random.default_rng(0)
    for mode, D, O in [(0, 8, 2), (0, 8, 10), (1, 8, 10), (2, 7, 2), (3, 7, 10)]:
        x, y = _x(rng, 6, D), rng.normal(size=(6, O))
        yp = y + rng.normal(size=(6, O))
        _, _, _, g = dyn.training_loss_and_grad(mode, x, yp, y, P)
        f = lambda v: float(dyn.training_loss(mode, x, v, y, P)[0])
        h = 1e-6
        for b, o in [(0, 0), (3, O - 1), (5, O // 2)]:
            e = np.zeros_like(yp)
            e[b, o] = h
            fd = (f(yp + e) - f(yp - e)) / (2 * h)
            assert abs(fd - g[b, o]) <= 1e-6 + 1e-5 * abs(fd), (mode, b, o, fd, g[b, o])


def test_onestep_aux_is_first_step_of_rollout():
    rng = np.random.default_rng(1)
    x = _x(rng, 32, 8)
    u = rng.normal(size=(32, 10))
    init8 = x[:, [0, 0, 1, 2, 3, 5, 6, 7]]
    roll = dyn.integrate_frenet_mult(np.concatenate([init8, u], 1), P).numpy()
    # aux reads a = x_u[8] = u[1] and deltv = x_u[9] = u[2] (dynamics.py:326-327): feed them in those slots
    u_aux = np.zeros((32, 10))
    u_aux[:, 1], u_aux[:, 2] = u[:, 0], u[:, 5]
    aux = dyn.dynamic_frenet_onestep_aux(np.concatenate([x[:, [0, 1, 2, 3, 5, 6, 7]], u_aux], 1), P).numpy()
    assert np.allclose(aux, roll[:, 0, 1:7], rtol=1e-12, atol=1e-12)


def test_two_output_network_reads_clamped_control_index():
    rng = np.random.default_rng(2)
    x7 = _x(rng, 8, 8)[:, [0, 1, 2, 3, 5, 6, 7]]
    u = rng.normal(size=(8, 2))
    out = dyn.dynamic_frenet_onestep_aux(np.concatenate([x7, u], 1), P).numpy()
    # both a and deltv come from the last column (JAX clamps x_u[9] to x_u[8])
    assert np.allclose(out[:, 2], x7[:, 2] + np.clip(u[:, 1], -P[10], P[10]) * P[8])
    assert np.allclose(out[:, 1], x7[:, 1] + np.clip(u[:, 1], -P[9], P[9]) * P[8])


def test_single_track_switches_model_at_3_mps():
    st = np.array([[0, 0, 0.1, 2.0, 0.2, 0.3, 0.05], [0, 0, 0.1, 5.0, 0.2, 0.3, 0.05]])
    xu = np.concatenate([st, np.ones((2, 10))], 1)
    out = dyn.integrate_st_mult(xu, P).numpy()
    assert out[0, 0, 5] == st[0, 5] and out[0, 0, 6] == st[0, 6]      # kinematic: PSI_DOT, BETA frozen
    assert out[1, 0, 5] != st[1, 5] and out[1, 0, 6] != st[1, 6]      # dynamic
    aux = dyn.dynamic_st_onestep_aux(xu[:, :9], P).numpy()
    assert np.allclose(aux[0], out[0, 0])
