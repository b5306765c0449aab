"""The optimiser update: oracle/optim_oracle.py against a hand-computed case and the
torch spelling on the CPU; irbfn_clip_adam (two CUDA launches) against both on the GPU.

Tolerance: parameters and moments within 2e-6 relative (norm-wise) of the float64 oracle after 25 steps --
float32 arithmetic of the update itself, nothing accumulates beyond that."""
import numpy as np
import pytest
import torch

from oracle import optim_oracle as oo


def test_oracle_matches_a_hand_computed_step():
    # one parameter, g = 3, max_norm = 1 -> clipped to 1; mu = 0.1, nu = 0.001; bias-corrected both to 1
    p, g, mu, nu, c, gn = oo.adam_step([2.0], [3.0], [0.0], [0.0], 0, lr=0.5, max_norm=1.0)
    assert gn == 3.0 and np.allclose(g, [1.0]) and np.allclose(mu, [0.1]) and np.allclose(nu, [0.001]) and c == 1
    assert np.allclose(p, [2.0 - 0.5 * 1.0 / (1.0 + 1e-8)])
    # second step, g = 0.5 (< max_norm: untouched)
    p2, g2, mu2, nu2, c2, gn2 = oo.adam_step(p, [0.5], mu, nu, c, lr=0.5, max_norm=1.0)
    assert gn2 == 0.5 and g2[0] == 0.5 and c2 == 2
    assert np.allclose(mu2, [0.9 * 0.1 + 0.1 * 0.5]) and np.allclose(nu2, [0.999 * 0.001 + 0.001 * 0.25])
    mh, nh = mu2[0] / (1 - 0.81), nu2[0] / (1 - 0.999 ** 2)
    assert np.allclose(p2, [p[0] - 0.5 * mh / (np.sqrt(nh) + 1e-8)])
    # the boundary: g_norm == max_norm is NOT below it, optax rescales (a no-op up to rounding)
    _, g3, *_ = oo.adam_step([0.0, 0.0], [0.6, 0.8], [0, 0], [0, 0], 0, lr=0.1, max_norm=1.0)
    assert np.allclose(g3, [0.6, 0.8])


def test_torch_spelling_matches_oracle_on_cpu():
    from irbfn_b200.train import clip_and_adam_torch_
    rng = np.random.default_rng(0)
    n = 1000
    p = rng.normal(size=n); mu = np.zeros(n); nu = np.zeros(n); c = 0
    pt, mt, nt = (torch.tensor(a, dtype=torch.float64) for a in (p, mu, nu))
    for step in range(1, 8):
        g = rng.normal(size=n) * (10.0 if step % 2 else 0.001)      # alternately clipped and not
        p, _, mu, nu, c, gn = oo.adam_step(p, g, mu, nu, c, lr=1e-2, max_norm=1.0)
        gt = torch.tensor(g, dtype=torch.float64)
        gnt = clip_and_adam_torch_(pt, gt, mt, nt, step, lr=1e-2, max_norm=1.0)
        assert abs(float(gnt) - gn) < 1e-12 * gn
    assert np.allclose(pt.numpy(), p, rtol=1e-12, atol=1e-14)
    assert np.allclose(mt.numpy(), mu, rtol=1e-12) and np.allclose(nt.numpy(), nu, rtol=1e-12)


def _rel(a, ref):
    return float(np.abs(np.asarray(a, np.float64) - ref).max() / max(np.abs(ref).max(), 1e-300))


@pytest.mark.gpu
@pytest.mark.parametrize("n,max_norm", [(1, 1.0), (11002, 1.0), (11002, 0.0), (1900000, 1.0), (4097, 1e9)])
def test_clip_adam_kernel_matches_oracle(n, max_norm, cuda_device):
    from irbfn_b200 import ops
    rng = np.random.default_rng(n)
    p = rng.normal(size=n).astype(np.float32)
    pd = torch.from_numpy(p.copy()).to(cuda_device)
    mud, nud = torch.zeros(n, device=cuda_device), torch.zeros(n, device=cuda_device)
    cnt = torch.zeros(1, device=cuda_device, dtype=torch.float64)
    p = p.astype(np.float64); mu = np.zeros(n); nu = np.zeros(n); c = 0
    for step in range(25):
        g = (rng.normal(size=n) * (3.0 if step % 3 == 0 else 1e-3 / np.sqrt(n))).astype(np.float32)
        gd = torch.from_numpy(g.copy()).to(cuda_device)
        gn_d = ops.clip_adam_(pd, gd, mud, nud, cnt, 1e-3, 0.9, 0.999, 1e-8, max_norm)
        p, gc, mu, nu, c, gn = oo.adam_step(p, g, mu, nu, c, lr=1e-3, max_norm=max_norm)
        assert abs(float(gn_d[0]) - gn) <= 1e-6 * gn
        assert _rel(gd.cpu().numpy(), gc) < 1e-6                    # the gradient is left clipped
    assert float(cnt[0]) == 25.0
    assert _rel(pd.cpu().numpy(), p) < 2e-6
    assert _rel(mud.cpu().numpy(), mu) < 2e-6 and _rel(nud.cpu().numpy(), nu) < 2e-6


@pytest.mark.gpu
def test_clip_adam_is_deterministic_and_checks_arguments(cuda_device):
    from irbfn_b200 import ops
    n = 50000
    g0 = torch.randn(n, device=cuda_device)
    outs = []
    for _ in range(2):
        p, g = torch.ones(n, device=cuda_device), g0.clone()
        mu, nu = torch.zeros(n, device=cuda_device), torch.zeros(n, device=cuda_device)
        cnt = torch.zeros(1, device=cuda_device, dtype=torch.float64)
        gn = ops.clip_adam_(p, g, mu, nu, cnt, 1e-3, max_norm=1.0)
        outs.append((p.clone(), gn.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    with pytest.raises(Exception):
        ops.clip_adam_(p, g[:-1].contiguous(), mu, nu, cnt, 1e-3)                       # length mismatch
    with pytest.raises(Exception):
        ops.clip_adam_(p, g, mu, nu, torch.zeros(1, device=cuda_device), 1e-3)          # count must be float64
    with pytest.raises(Exception):
        ops.clip_adam_(p, g, mu, nu, cnt, 1e-3, b1=1.0)                                 # b1 outside [0, 1)


@pytest.mark.gpu
@pytest.mark.parametrize("n,max_norm", [(1, 1.0), (11002, 1.0), (11002, 0.0), (131071, 1.0), (4097, 1e9)])
def test_fused_single_launch_update_matches_oracle(n, max_norm, cuda_device):
    """irbfn_fused_update without peers (one process): global norm + clip + Adam in ONE launch, against the
    optax restatement and against the two-launch irbfn_clip_adam."""
    from irbfn_b200 import ops
    rng = np.random.default_rng(n)
    p0 = rng.normal(size=n).astype(np.float32)
    dev = cuda_device
    p, mu, nu = torch.tensor(p0, device=dev), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    p2, mu2, nu2 = p.clone(), mu.clone(), nu.clone()
    cnt, cnt2 = torch.zeros(1, device=dev, dtype=torch.float64), torch.zeros(1, device=dev, dtype=torch.float64)
    gn = torch.zeros(1, device=dev)
    rp, rmu, rnu, rc = p0.astype(np.float64), np.zeros(n), np.zeros(n), 0
    for step in range(4):
        g0 = (rng.normal(size=n) * (3.0 if step % 2 == 0 else 1e-3)).astype(np.float32)   # clipped / not clipped
        g = torch.tensor(np.concatenate([g0, [7.5]]).astype(np.float32), device=dev)      # [grad | loss]
        g2 = torch.tensor(g0, device=dev)
        ops.fused_update_(p, mu, nu, g, n, 1, cnt, 1e-3, max_norm=max_norm, grad_norm=gn)
        ops.clip_adam_(p2, g2, mu2, nu2, cnt2, 1e-3, max_norm=max_norm)
        rp, rg, rmu, rnu, rc, rnorm = oo.adam_step(rp, g0.astype(np.float64), rmu, rnu, rc, lr=1e-3, max_norm=max_norm)
        assert float(cnt) == step + 1 and float(g[n]) == 7.5             # the trailing float is left alone
        assert abs(float(gn) - rnorm) <= 1e-5 * rnorm
        assert _rel(g[:n].cpu().numpy(), rg) <= 2e-6                     # the gradient is clipped in place
        assert _rel(p.cpu().numpy(), rp) <= 2e-6 and _rel(mu.cpu().numpy(), rmu) <= 2e-6
        assert torch.allclose(p, p2, rtol=1e-6, atol=1e-7) and torch.allclose(nu, nu2, rtol=1e-5, atol=1e-12)
    from irbfn_b200._lib import IrbfnError
    with pytest.raises(IrbfnError):                                      # beyond the single-CTA limit
        ops.fused_update_(None, None, None, torch.zeros((1 << 17) + 8, device=dev), (1 << 17) + 4, 1, None, 1e-3,
                          apply_update=False)
