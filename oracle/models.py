"""Oracle: the reference's model log-joints and parameter layouts (torch-CPU).

Restates the NumPyro model functions of the reference:
  src/models/uci_bnn.py:21-68        UCI BNN  in->50->50->1 ReLU, prec ~ Gamma(1, 0.1)
  src/models/reg_1d_bnn.py:22-67     1-D BNN  1->h->h->1 tanh, fixed noise 0.1
  src/models/mlp.py:14-65            MNIST MLP 784->100->100->10 tanh, Categorical
  experiments/variance_collapse.py:70-74   MVN(0, I_d)
and the AutoNormal / AutoDelta parameter naming + ravel order of
numpyro/infer/autoguide.py and jax.flatten_util (sorted names).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import math
from dataclasses import dataclass

import numpy as np
import torch

LOG_2PI = math.log(2.0 * math.pi)

T_IDENTITY = 0
T_EXP = 1


@dataclass(frozen=True)
class Site:
    name: str
    shape: tuple
    transform: int = T_IDENTITY

    @property
    def size(self):
        return int(np.prod(self.shape)) if len(self.shape) else 1


@dataclass(frozen=True)
class ModelSpec:
    kind: str            # "bnn_reg" | "bnn_cls" | "gauss"
    in_dim: int = 1
    hidden: int = 50
    out_dim: int = 1
    act: str = "relu"    # "relu" | "tanh"
    noise: str = "gamma"  # "gamma" (prec ~ Gamma(1, .1)) | "fixed"
    sigma: float = 0.1
    prefix: str = ""     # "nn_" for reg_1d_bnn

    def sites(self):
        """Latent sites in MODEL order (order of the sample statements)."""
        if self.kind == "gauss":
            return [Site("x", (self.in_dim,))]
        p, i, h, o = self.prefix, self.in_dim, self.hidden, self.out_dim
        if self.kind == "bnn_reg":
            s = [Site("prec", (), T_EXP)] if self.noise == "gamma" else []
            return s + [Site(p + "w1", (i, h)), Site(p + "b1", (h,)), Site(p + "w2", (h, h)),
                        Site(p + "b2", (h,)), Site(p + "w3", (h,)), Site(p + "b3", ())]
        if self.kind == "bnn_cls":
            # flax Dense params as lifted by random_flax_module("mlp", ...) [NP?: names]
            return [Site("mlp/Dense_0.bias", (h,)), Site("mlp/Dense_0.kernel", (i, h)),
                    Site("mlp/Dense_1.bias", (h,)), Site("mlp/Dense_1.kernel", (h, h)),
                    Site("mlp/Dense_2.bias", (o,)), Site("mlp/Dense_2.kernel", (h, o))]
        raise ValueError(self.kind)

    @property
    def latent_dim(self):
        return sum(s.size for s in self.sites())


def uci_bnn(in_dim=13):
    return ModelSpec("bnn_reg", in_dim, 50, 1, "relu", "gamma")


def reg_1d_bnn(hidden):
    return ModelSpec("bnn_reg", 1, hidden, 1, "tanh", "fixed", 0.1, "nn_")


def mnist_mlp(in_dim=784, hidden=100, out_dim=10):
    return ModelSpec("bnn_cls", in_dim, hidden, out_dim, "tanh")


def gaussian(d):
    return ModelSpec("gauss", d)


def layout(spec, guide="normal"):
    """Site table: for every site (model order) its offset in the latent vector and the
    offsets of its loc / rho leaves inside the raveled Stein particle (sorted names).
    Returns (rows, L, d); rows = list of dict(name, size, transform, lat, loc, rho)."""
    sites = spec.sites()
    lat, off = {}, 0
    for s in sites:
        lat[s.name] = off
        off += s.size
    L = off
    names = []
    for s in sites:
        names.append((s.name + "_auto_loc", s, "loc"))
        if guide == "normal":
            names.append((s.name + "_auto_scale", s, "rho"))
    names.sort(key=lambda t: t[0])
    pos, off = {}, 0
    for nm, s, which in names:
        pos[(s.name, which)] = off
        off += s.size
    d = off
    rows = [dict(name=s.name, size=s.size, transform=s.transform, lat=lat[s.name],
                 loc=pos[(s.name, "loc")], rho=pos.get((s.name, "rho"), -1)) for s in sites]
    return rows, L, d


def _act(spec, t):
    return torch.relu(t) if spec.act == "relu" else torch.tanh(t)


def _unpack(spec, theta):
    """theta (..., L) constrained latent -> dict of named tensors with site shapes."""
    out, off = {}, 0
    for s in spec.sites():
        out[s.name] = theta[..., off:off + s.size].reshape(theta.shape[:-1] + tuple(s.shape))
        off += s.size
    return out


def forward(spec, theta, xb):
    """Network output for latent batch theta (S, L) and inputs xb (S, B, in) -> (S, B[, out])."""
    t = _unpack(spec, theta)
    p = spec.prefix
    if spec.kind == "bnn_reg":
        h1 = _act(spec, torch.einsum("sbi,sih->sbh", xb, t[p + "w1"]) + t[p + "b1"][:, None, :])
        h2 = _act(spec, torch.einsum("sbh,shk->sbk", h1, t[p + "w2"]) + t[p + "b2"][:, None, :])
        return torch.einsum("sbh,sh->sb", h2, t[p + "w3"]) + t[p + "b3"][:, None]
    if spec.kind == "bnn_cls":
        h1 = _act(spec, torch.einsum("sbi,sih->sbh", xb, t["mlp/Dense_0.kernel"]) + t["mlp/Dense_0.bias"][:, None, :])
        h2 = _act(spec, torch.einsum("sbh,shk->sbk", h1, t["mlp/Dense_1.kernel"]) + t["mlp/Dense_1.bias"][:, None, :])
        return torch.einsum("sbh,sho->sbo", h2, t["mlp/Dense_2.kernel"]) + t["mlp/Dense_2.bias"][:, None, :]
    raise ValueError(spec.kind)


def log_joint(spec, theta, xb=None, yb=None, lik_scale=1.0):
    """log p(theta, y_b | x_b) for a batch of constrained latents theta (S, L) -> (S,).
    Likelihood terms are multiplied by lik_scale = N/B (plate subsample scaling)."""
    if spec.kind == "gauss":
        return -0.5 * (theta * theta).sum(-1) - 0.5 * spec.in_dim * LOG_2PI
    lp = 0.0
    off = 0
    for s in spec.sites():
        v = theta[..., off:off + s.size]
        if s.transform == T_EXP:        # prec ~ Gamma(concentration 1, rate 0.1)
            lp = lp + (math.log(0.1) - 0.1 * v).sum(-1)
        else:                           # Normal(0, 1)
            lp = lp + (-0.5 * v * v - 0.5 * LOG_2PI).sum(-1)
        off += s.size
    out = forward(spec, theta, xb)
    if spec.kind == "bnn_reg":
        if spec.noise == "gamma":
            prec = theta[..., 0:1]
            ll = -0.5 * prec * (yb - out) ** 2 + 0.5 * torch.log(prec) - 0.5 * LOG_2PI
        else:
            ll = -0.5 * ((yb - out) / spec.sigma) ** 2 - math.log(spec.sigma) - 0.5 * LOG_2PI
        return lp + lik_scale * ll.sum(-1)
    lsm = torch.log_softmax(out, dim=-1)
    ll = torch.gather(lsm, -1, yb[..., None].long()).squeeze(-1)
    return lp + lik_scale * ll.sum(-1)
