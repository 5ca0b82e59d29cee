"""Pins for the mixture-ELBO oracle: closed forms and central finite differences (float64)."""
import math

import numpy as np
import torch

from oracle import elbo as oelbo
from oracle import models as omodels


def _fd_grad(f, x, e=1e-6):
    g = np.zeros_like(x)
    for k in range(x.size):
        xp, xm = x.copy(), x.copy()
        xp[k] += e
        xm[k] -= e
        g[k] = (f(xp) - f(xm)) / (2 * e)
    return g


def test_layout_sorted_names_uci():
    rows, L, d = omodels.layout(omodels.uci_bnn(13), "normal")
    assert L == 3302 and d == 6604
    by = {r["name"]: r for r in rows}
    # sorted leaf names: b1_loc, b1_scale, b2_.., b3_.., prec_.., w1_.., w2_.., w3_..
    assert by["b1"]["loc"] == 0 and by["b1"]["rho"] == 50 and by["b2"]["loc"] == 100
    assert by["prec"]["lat"] == 0 and by["w1"]["lat"] == 1
    assert omodels.layout(omodels.mnist_mlp(), "normal")[1:] == (89610, 179220)      # config C4
    assert omodels.layout(omodels.reg_1d_bnn(50), "normal")[1:] == (2701, 5402)      # config C1
    assert omodels.layout(omodels.gaussian(10), "delta")[1:] == (10, 10)


def test_gaussian_delta_closed_form():
    om = omodels.gaussian(7)
    psi = np.random.default_rng(0).standard_normal((4, 7))
    G, E = oelbo.elbo_grads(om, psi, None, None, tau=0.25, guide="delta", dtype=torch.float64)
    np.testing.assert_allclose(G, -0.25 * psi, rtol=1e-12)
    np.testing.assert_allclose(E, 0.25 * (-0.5 * (psi ** 2).sum(1) - 3.5 * math.log(2 * math.pi)), rtol=1e-12)


def test_single_component_normal_guide_is_plain_elbo():
    """m = 1: log q is the entropy term of one Gaussian; gradient wrt loc of E[log p] - E[log q]."""
    om = omodels.gaussian(3)
    rng = np.random.default_rng(1)
    psi = rng.standard_normal((1, 6))
    eps = rng.standard_normal((1, 50, 3))
    G, E = oelbo.elbo_grads(om, psi, eps, None, guide="normal", dtype=torch.float64)
    rows, L, d = omodels.layout(om, "normal")
    loc, rho = psi[0, rows[0]["loc"]:rows[0]["loc"] + 3], psi[0, rows[0]["rho"]:rows[0]["rho"] + 3]
    sig = np.log1p(np.exp(rho))
    u = loc + sig * eps[0]
    # d/dloc mean_s [-0.5 u^2] = -mean_s u ; entropy does not depend on loc
    np.testing.assert_allclose(G[0, rows[0]["loc"]:rows[0]["loc"] + 3], -u.mean(0), rtol=1e-10)
    # d/dsig: mean_s[-u eps] + 1/sig ; chain through softplus
    dsig = (-u * eps[0]).mean(0) + 1.0 / sig
    np.testing.assert_allclose(G[0, rows[0]["rho"]:rows[0]["rho"] + 3], dsig / (1 + np.exp(-rho)), rtol=1e-10)


def test_mixture_elbo_gradient_finite_differences():
    rng = np.random.default_rng(2)
    for om, kind in [(omodels.uci_bnn(2), "reg"), (omodels.reg_1d_bnn(3), "reg1"), (omodels.mnist_mlp(4, 3, 3), "cls")]:
        rows, L, d = omodels.layout(om, "normal")
        m, S, N, B = 3, 2, 9, 4
        psi = 0.3 * rng.standard_normal((m, d))
        eps = rng.standard_normal((m, S, L))
        idx = np.stack([rng.permutation(N)[:B] for _ in range(m * S)]).reshape(m, S, B)
        x = rng.standard_normal((N, om.in_dim))
        y = rng.integers(0, 3, N) if kind == "cls" else rng.standard_normal(N)
        G, E = oelbo.elbo_grads(om, psi, eps, idx, x, y, N, 0.7, "normal", torch.float64)
        psi_t = torch.tensor(psi)
        xt = torch.tensor(x)
        yt = torch.tensor(y) if kind != "cls" else torch.tensor(y).long()
        i = 1

        def f(v):
            return float(oelbo.particle_elbo(om, psi_t, i, torch.tensor(eps[i]), torch.tensor(idx[i]).long(), xt, yt,
                                             N, 0.7, "normal", psi_i=torch.tensor(v)))
        np.testing.assert_allclose(f(psi[i]), E[i], rtol=1e-12)
        fd = _fd_grad(f, psi[i].copy())
        np.testing.assert_allclose(G[i], fd, rtol=2e-5, atol=2e-6)


def test_other_particles_are_constants():
    """The gradient of particle i's objective ignores the dependence of q_j (j != i) on psi_j."""
    om = omodels.gaussian(2)
    rng = np.random.default_rng(3)
    psi = 0.1 * rng.standard_normal((3, 4))
    eps = rng.standard_normal((3, 5, 2))
    G1, _ = oelbo.elbo_grads(om, psi, eps, None, guide="normal", dtype=torch.float64)
    psi2 = psi.copy()
    psi2[2] += 0.05       # moving particle 2 changes G[0] only through the constants
    G2, _ = oelbo.elbo_grads(om, psi2, eps, None, guide="normal", dtype=torch.float64)
    assert np.abs(G1[0] - G2[0]).max() > 1e-6 and G1.shape == (3, 4)
