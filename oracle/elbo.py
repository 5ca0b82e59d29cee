"""Oracle: per-particle mixture-ELBO value and gradient (torch-CPU autograd).

Restates numpyro/contrib/einstein/stein_loss.py (SteinLoss.single_particle_loss /
particle_loss) and the autoguides of numpyro/infer/autoguide.py:

  u        = loc_i + softplus(rho_i) * eps                (AutoNormal; AutoDelta: u = loc_i)
  theta    = T(u)                                          (exp for positive sites)
  log q_j  = sum_k Normal(u_k; loc_jk, softplus(rho_jk)) - log|dT/du|
  log q    = logsumexp_j log q_j - log m                   (j == i differentiable, j != i constant)
  ELBO_is  = tau * [log prior(theta) + (N/B) sum_n log lik] - log q
  g_i      = d/d psi_i mean_s ELBO_is

Engine construction sites in the reference: experiments/uci_benchmark.py:188-195,
experiments/img_classifier.py:170-177, experiments/variance_collapse.py:97-123.
Gradients come from torch autograd so they are independent of the hand-derived CUDA backward.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import math

import numpy as np
import torch
import torch.nn.functional as Fnn

from .models import LOG_2PI, T_EXP, layout, log_joint


def _gather_index(rows, key):
    """Index vector mapping latent position -> position inside psi of leaf `key`."""
    idx = []
    for r in rows:
        idx.extend(range(r[key], r[key] + r["size"]))
    return torch.tensor(idx, dtype=torch.long)


def particle_elbo(spec, psi, i, eps_i, idx_i, x, y, n_total, tau=1.0, guide="normal",
                  psi_i=None):
    """mean_s ELBO_is for particle i.
    psi (m, d) all particles (constants); psi_i optional (d,) leaf that replaces row i and
    carries the gradient.  eps_i (S, L) noise (ignored for guide="delta": S = 1).
    idx_i (S, B) int subsample indices or None (full batch / no data)."""
    rows, L, d = layout(spec, guide)
    m = psi.shape[0]
    dt = psi.dtype
    loc_ix = _gather_index(rows, "loc")
    me = psi[i] if psi_i is None else psi_i
    loc_i = me[loc_ix]
    tmask = torch.tensor([r["transform"] == T_EXP for r in rows for _ in range(r["size"])])
    if guide == "normal":
        rho_ix = _gather_index(rows, "rho")
        sig_i = Fnn.softplus(me[rho_ix])
        u = loc_i[None, :] + sig_i[None, :] * eps_i                      # (S, L)
        others_loc = psi[:, loc_ix].detach()
        others_sig = Fnn.softplus(psi[:, rho_ix]).detach()
        sel = (torch.arange(m) == i)[:, None]
        locs = torch.where(sel, loc_i[None, :], others_loc)              # (m, L)
        sigs = torch.where(sel, sig_i[None, :], others_sig)
        zs = (u[:, None, :] - locs[None]) / sigs[None]                   # (S, m, L)
        lq_j = (-0.5 * zs * zs - torch.log(sigs)[None] - 0.5 * LOG_2PI).sum(-1)   # (S, m)
        logjac = (u * tmask.to(dt)).sum(-1)                              # log|d exp(u)/du| = u
        lq = torch.logsumexp(lq_j, dim=1) - logjac - math.log(m)
    else:
        u = loc_i[None, :]
        lq = torch.zeros(1, dtype=dt)                                    # Delta guide: log q = 0
    theta = torch.where(tmask[None, :], torch.exp(u), u)
    if spec.kind == "gauss":
        lp = log_joint(spec, theta)
    else:
        S = u.shape[0]
        if idx_i is None:
            xb = x[None].expand(S, *x.shape)
            yb = y[None].expand(S, *y.shape)
            scale = 1.0
        else:
            xb, yb = x[idx_i], y[idx_i]
            scale = float(n_total) / idx_i.shape[1]
        lp = log_joint(spec, theta, xb, yb, scale)
    return (tau * lp - lq).mean()


def elbo_grads(spec, psi, eps, idx, x=None, y=None, n_total=None, tau=1.0, guide="normal",
               dtype=torch.float32, rows=None):
    """G (m, d) = vmap_i grad_{psi_i} mean_s ELBO_is  and  elbo (m,).
    psi (m, d); eps (m, S, L) or None; idx (m, S, B) or None.
    rows = (row0, nrows): only those particles' gradients (a row shard); eps / idx then hold
    the shard's rows only and the results have nrows rows."""
    psi_t = torch.as_tensor(np.asarray(psi)).to(dtype)
    m, d = psi_t.shape
    x_t = None if x is None else torch.as_tensor(np.asarray(x)).to(dtype)
    if y is None:
        y_t = None
    else:
        y_np = np.asarray(y)
        y_t = torch.as_tensor(y_np).to(dtype) if spec.kind != "bnn_cls" else torch.as_tensor(y_np).long()
    row0, nrows = (0, m) if rows is None else rows
    G = torch.zeros(nrows, d, dtype=dtype)
    E = torch.zeros(nrows, dtype=dtype)
    for r in range(nrows):
        i = row0 + r
        leaf = psi_t[i].clone().requires_grad_(True)
        eps_i = None if eps is None else torch.as_tensor(np.asarray(eps[r])).to(dtype)
        idx_i = None if idx is None else torch.as_tensor(np.array(idx[r])).long()
        val = particle_elbo(spec, psi_t, i, eps_i, idx_i, x_t, y_t, n_total, tau, guide, psi_i=leaf)
        (g,) = torch.autograd.grad(val, leaf)
        G[r] = g
        E[r] = val.detach()
    return G.numpy(), E.numpy()
