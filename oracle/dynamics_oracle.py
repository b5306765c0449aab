"""TEST INFRASTRUCTURE -- CPU restatement of the reference's loss-side vehicle dynamics and training
losses.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
may import this; the product (irbfn_b200/) never does.

Parity pinned by executing the reference: src/irbfn_mpc/dynamics.py and the nested train_step_*
functions of scripts/train_nmpc.py / scripts/train_nmpc_frenet.py are run UNMODIFIED under oracle/refshim
(tests/golden/make_golden.py -> tests/golden/dynamics_ref.npz, train_step_ref.npz); this restatement
(torch CPU, float64 by default, gradients from torch autograd) reproduces those values to 1e-12
(tests/test_oracle_vs_ref.py) and is the independent check of the hand-derived adjoints in
irbfn_b200/csrc/irbfn_dynamics.cu.

Restated (paths relative to the reference checkout):
  dynamic_st_onestep           src/irbfn_mpc/dynamics.py:9-91
  integrate_st_mult            src/irbfn_mpc/dynamics.py:94-100
  dynamic_st_onestep_aux       src/irbfn_mpc/dynamics.py:103-187
  dynamic_frenet_onestep       src/irbfn_mpc/dynamics.py:190-281
  integrate_frenet_mult        src/irbfn_mpc/dynamics.py:284-290
  dynamic_frenet_onestep_aux   src/irbfn_mpc/dynamics.py:293-383
  loss of train_step_oneint    scripts/train_nmpc_frenet.py:339-392   (mode 0)
  loss of train_step_fullint   scripts/train_nmpc_frenet.py:395-420   (mode 1)
  loss of train_step_oneint    scripts/train_nmpc.py:258-300          (mode 2)
  loss of train_step_fullint   scripts/train_nmpc.py:303-421          (mode 3)

JAX semantics kept on purpose: jnp.clip is minimum(maximum(.)) (gradient 1 inside, 0 outside);
static out-of-range indices clamp (x_u[9] on a 9-vector reads x_u[8], which
dynamic_frenet_onestep_aux hits when the network has 2 outputs); d|z|/dz = sign(z), 0 at 0.
"""
from __future__ import annotations

import numpy as np
import torch

G = 9.81
MODE_FRENET_ONEINT, MODE_FRENET_FULLINT, MODE_ST_ONEINT, MODE_ST_FULLINT = 0, 1, 2, 3


def default_dyn_params(mu=1.0489, cs=4.718):
    """dyn_params of scripts/train_nmpc_frenet.py:322-338 (args.mu, args.cs are CLI values)."""
    return np.array([mu, 1.0489, 0.04712, 0.15875, 0.17145, cs, cs, 0.074, 0.1, 3.2, 9.51, 0.4189, 7.0])


def _t(a, dtype):
    return a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a), dtype=dtype)


def _clip(v, lo, hi):
    return torch.clamp(v, min=lo, max=hi)


def _col(a, i):
    """a[:, i] with JAX's clamped static index."""
    return a[:, min(max(i, 0), a.shape[1] - 1)]


# ---------------------------------------------------------------------------- Frenet
def frenet_step_ls(st, a_raw, dv_raw, P):
    """One step of dynamic_frenet_onestep (dynamics.py:190-281): x + deriv_x_ls * dt.
    st: tuple (s, ey, delta, vx, vy, wz, epsi, cur) of (B,) tensors."""
    LF, LR, dt, sv_max, a_max, s_max = P[3], P[4], P[8], P[9], P[10], P[11]
    s, ey, dl, vx, vy, wz, ep, cu = st
    dc = _clip(dl, -s_max, s_max)
    a = _clip(a_raw, -a_max, a_max)
    dv = _clip(dv_raw, -sv_max, sv_max)
    sd = (vx * torch.cos(ep)) / (1 - ey * cu)
    return (s + sd * dt, ey + vx * torch.sin(ep) * dt, dl + dv * dt, vx + a * dt, vy, wz,
            ep + ((vx * torch.tan(dc)) / (LR + LF) - cu * ((vx * torch.cos(ep)) / (1 - cu * ey))) * dt, cu)


def integrate_frenet_mult(x_and_u, P, dtype=torch.float64):
    """(B,18) -> (B,5,8)   dynamics.py:284-290; seq = u.reshape(5,2,order='F')."""
    xu = _t(x_and_u, dtype)
    st = tuple(xu[:, i] for i in range(8))
    out = []
    for i in range(5):
        st = frenet_step_ls(st, xu[:, 8 + i], xu[:, 13 + i], P)
        out.append(torch.stack(st, -1))
    return torch.stack(out, 1)


def dynamic_frenet_onestep_aux(x_u, P, dtype=torch.float64):
    """(B,7+O) -> (B,6)   dynamics.py:293-383."""
    xu = _t(x_u, dtype)
    LF, LR, dt, sv_max, a_max, s_max = P[3], P[4], P[8], P[9], P[10], P[11]
    ey, dl, vx, vy, wz, ep, cu = (xu[:, i] for i in range(7))
    dc = _clip(dl, -s_max, s_max)
    a = _clip(_col(xu, 8), -a_max, a_max)
    dv = _clip(_col(xu, 9), -sv_max, sv_max)
    epd = (vx * torch.tan(dc)) / (LR + LF) - cu * ((vx * torch.cos(ep)) / (1 - cu * ey))
    z = torch.zeros_like(ey)
    return torch.stack([ey + vx * torch.sin(ep) * dt, dl + dv * dt, vx + a * dt, vy + z, wz + z, ep + epd * dt], -1)


# ---------------------------------------------------------------------------- single track
def _st_derivs(X, P):
    """f (dynamic) and f_ks (kinematic) of dynamics.py:47-88 for state columns X (B,7) + controls."""
    mu, m, I, lf, lr, C_Sf, C_Sr, h = (P[i] for i in range(8))
    sv_max, a_max, s_max, v_max = P[9], P[10], P[11], P[12]
    x7, a_raw, sv_raw = X
    DELTA = _clip(x7[:, 2], -s_max, s_max)
    V = _clip(x7[:, 3], -v_max, v_max)
    PSI, PSI_DOT, BETA = x7[:, 4], x7[:, 5], x7[:, 6]
    ACCL = _clip(a_raw, -a_max, a_max)
    SV = _clip(sv_raw, -sv_max, sv_max)
    rf = G * lr - ACCL * h
    rr = G * lf + ACCL * h
    f = torch.stack([
        V * torch.cos(PSI + BETA), V * torch.sin(PSI + BETA), SV, ACCL, PSI_DOT,
        ((mu * m) / (I * (lf + lr))) * (lf * C_Sf * rf * DELTA + (lr * C_Sr * rr - lf * C_Sf * rf) * BETA
                                        - (lf * lf * C_Sf * rf + lr * lr * C_Sr * rr) * (PSI_DOT / V)),
        (mu / (V * (lr + lf))) * (C_Sf * rf * DELTA - (C_Sr * rr + C_Sf * rf) * BETA
                                  + (C_Sr * rr * lr - C_Sf * rf * lf) * (PSI_DOT / V)) - PSI_DOT], -1)
    z = torch.zeros_like(V)
    f_ks = torch.stack([V * torch.cos(PSI), V * torch.sin(PSI), SV, ACCL, (V / (lr + lf)) * torch.tan(DELTA), z, z], -1)
    return f, f_ks, V


def integrate_st_mult(x_and_u, P, dtype=torch.float64):
    """(B,17) -> (B,5,7)   dynamics.py:94-100: x + select(V > 3, f, f_ks) * dt, five steps."""
    xu = _t(x_and_u, dtype)
    x7 = xu[:, :7]
    out = []
    for i in range(5):
        f, f_ks, V = _st_derivs((x7, xu[:, 7 + i], xu[:, 12 + i]), P)
        x7 = x7 + torch.where((V > 3.0)[:, None], f, f_ks) * P[8]
        out.append(x7)
    return torch.stack(out, 1)


def dynamic_st_onestep_aux(x_u, P, dtype=torch.float64):
    """(B,9+) -> (B,7)   dynamics.py:103-187: x_u[:7] + f_ks * dt."""
    xu = _t(x_u, dtype)
    _, f_ks, _ = _st_derivs((xu[:, :7], xu[:, 7], xu[:, 8]), P)
    return xu[:, :7] + f_ks * P[8]


# ---------------------------------------------------------------------------- training losses
def training_loss(mode, x, y_pred, y, P, dtype=torch.float64):
    """(loss, pred_loss, int_loss) as torch scalars for the four reference training steps.
    int_loss is the weighted term that enters the loss (mode 0 returns 100 * int_loss like the
    reference's aux output, scripts/train_nmpc_frenet.py:384)."""
    x, yp, y = _t(x, dtype), _t(y_pred, dtype), _t(y, dtype)
    if mode == MODE_FRENET_ONEINT:          # scripts/train_nmpc_frenet.py:343-384
        init = x[:, [0, 1, 2, 3, 5, 6, 7]]
        pred_loss = (yp - y).abs().mean()
        act = dynamic_frenet_onestep_aux(torch.cat([init, y], 1), P, dtype)
        pre = dynamic_frenet_onestep_aux(torch.cat([init, yp], 1), P, dtype)
        int_loss = 100.0 * (pre - act).abs().mean()
    elif mode == MODE_FRENET_FULLINT:       # scripts/train_nmpc_frenet.py:399-413
        init = x[:, [0, 0, 1, 2, 3, 5, 6, 7]]
        pred_loss = (yp - y).abs().mean()
        act = integrate_frenet_mult(torch.cat([init, y], 1), P, dtype)
        pre = integrate_frenet_mult(torch.cat([init, yp], 1), P, dtype)
        int_loss = (pre - act).abs().mean()
    elif mode == MODE_ST_ONEINT:            # scripts/train_nmpc.py:260-293
        init = torch.zeros((x.shape[0], 7), dtype=dtype)
        init[:, 3], init[:, 6], init[:, 5] = x[:, 0], x[:, 5], x[:, 6]
        pred_loss = (0.5 * (yp - y) ** 2).mean()
        act = dynamic_st_onestep_aux(torch.cat([init, y], 1), P, dtype)
        pre = dynamic_st_onestep_aux(torch.cat([init, yp], 1), P, dtype)
        sel = [0, 1, 3, 4]
        int_loss = (0.5 * (pre[:, sel] - act[:, sel]) ** 2).mean()
    elif mode == MODE_ST_FULLINT:           # scripts/train_nmpc.py:306-413
        DT, WB, MAXV, MINV, MAXS = 0.1, 0.33, 7.0, 0.0, 0.4189

        def roll(u):
            X = torch.zeros(x.shape[0], dtype=dtype)
            Y, dl, yaw = X.clone(), X.clone(), X.clone()
            v = _clip(x[:, 0], MINV, MAXV)
            for i in range(5):
                a, dv = u[:, i], u[:, i + 5]
                X = X + v * torch.cos(yaw) * DT
                Y = Y + v * torch.sin(yaw) * DT
                dl = _clip(dl + dv * DT, -MAXS, MAXS)
                v = _clip(v + a * DT, MINV, MAXV)
                yaw = yaw + (v / WB) * torch.tan(dl) * DT
            return torch.stack([X, Y, dl, v, yaw], -1)

        pred_loss = (yp[:, [0, 5]] - y[:, [0, 5]]).abs().mean()
        # the reference's middle term is |first_states_pred - first_states_pred| == 0 (train_nmpc.py:404)
        int_loss = (roll(yp) - roll(y)).abs().mean()
    else:
        raise ValueError("mode")
    return pred_loss + int_loss, pred_loss, int_loss


def training_loss_and_grad(mode, x, y_pred, y, P):
    """float64 (loss, pred_loss, int_loss, dloss/dy_pred (B,O)) by autograd."""
    yp = torch.as_tensor(np.asarray(y_pred), dtype=torch.float64).clone().requires_grad_(True)
    loss, pl, il = training_loss(mode, x, yp, y, P)
    (g,) = torch.autograd.grad(loss, yp)
    return float(loss.detach()), float(pl.detach()), float(il.detach()), g.numpy()
