"""Model and guide descriptors: the log-joints the reference's experiments feed to SteinVI.

NumPyro is not available, so a model is not an arbitrary Python callable but a descriptor of
one of the reference's model functions; the CUDA kernels implement their log-joints:

  BNNRegression(in_dim, hidden, act, noise)  src/models/uci_bnn.py:21-68 (ReLU, Gamma(1, .1) precision)
                                             src/models/reg_1d_bnn.py:22-67 (tanh, fixed sigma 0.1)
  BNNClassifier(in_dim, hidden, out_dim)     src/models/mlp.py:14-65 (tanh MLP, Categorical)
  GaussianTarget(d)                          experiments/variance_collapse.py:70-74

Guides: AutoNormal (numpyro.infer.autoguide.AutoNormal: per-site `<site>_auto_loc`,
`<site>_auto_scale` = softplus(rho)) and AutoDelta (`<site>_auto_loc` only).  A Stein particle
is the concatenation of the unconstrained leaves in lexicographically sorted name order
(jax.flatten_util.ravel_pytree over a dict).
"""

import math
from dataclasses import dataclass, field

import numpy as np

from . import _lib


@dataclass(frozen=True)
class Site:
    name: str
    shape: tuple
    transform: int = _lib.T_IDENTITY

    @property
    def size(self):
        return int(np.prod(self.shape)) if len(self.shape) else 1


class init_to_uniform:
    """numpyro.infer.init_to_uniform(radius): U(-radius, radius) in unconstrained space
    (reference experiments/uci_benchmark.py:179, experiments/variance_collapse.py:91-93)."""

    def __init__(self, radius=2.0):
        self.radius = float(radius)


class AutoNormal:
    """Mean-field Normal guide (reference experiments/uci_benchmark.py:179,
    experiments/img_classifier.py:170-172)."""

    kind = _lib.GUIDE_NORMAL

    def __init__(self, model=None, init_loc_fn=None, init_scale=0.1, **_ignored):
        self.model = model
        self.init_loc_fn = init_loc_fn if init_loc_fn is not None else init_to_uniform(2.0)
        self.init_scale = float(init_scale)


class AutoDelta:
    """Point-mass guide used by SVGD / ASVGD (reference src/methods/svgd.py:15-22)."""

    kind = _lib.GUIDE_DELTA

    def __init__(self, model=None, init_loc_fn=None, **_ignored):
        self.model = model
        self.init_loc_fn = init_loc_fn if init_loc_fn is not None else init_to_uniform(2.0)
        self.init_scale = None


@dataclass(frozen=True)
class Model:
    """Base descriptor.  `sites()` lists the latent sites in MODEL order."""

    def sites(self):
        raise NotImplementedError

    @property
    def latent_dim(self):
        return sum(s.size for s in self.sites())

    # data handling: map the reference's call-site kwargs to (x, y, n_total, B)
    def bind(self, args, kwargs):
        raise NotImplementedError


@dataclass(frozen=True)
class BNNRegression(Model):
    in_dim: int = 13
    hidden: int = 50
    act: str = "relu"
    noise: str = "gamma"   # "gamma": prec ~ Gamma(1, 0.1); "fixed": sigma
    sigma: float = 0.1
    prefix: str = ""       # "nn_" in reg_1d_bnn.py

    kind = _lib.MODEL_BNN_REG
    out_dim = 1

    def sites(self):
        p, i, h = self.prefix, self.in_dim, self.hidden
        s = [Site("prec", (), _lib.T_EXP)] if self.noise == "gamma" else []
        return s + [Site(p + "w1", (i, h)), Site(p + "b1", (h,)), Site(p + "w2", (h, h)),
                    Site(p + "b2", (h,)), Site(p + "w3", (h,)), Site(p + "b3", ())]


@dataclass(frozen=True)
class BNNClassifier(Model):
    in_dim: int = 784
    hidden: int = 100
    out_dim: int = 10
    act: str = "tanh"

    kind = _lib.MODEL_BNN_CLS
    noise = "fixed"
    sigma = 1.0

    def sites(self):
        i, h, o = self.in_dim, self.hidden, self.out_dim
        return [Site("mlp/Dense_0.bias", (h,)), Site("mlp/Dense_0.kernel", (i, h)),
                Site("mlp/Dense_1.bias", (h,)), Site("mlp/Dense_1.kernel", (h, h)),
                Site("mlp/Dense_2.bias", (o,)), Site("mlp/Dense_2.kernel", (h, o))]


@dataclass(frozen=True)
class GaussianTarget(Model):
    dim: int = 10

    kind = _lib.MODEL_GAUSS
    in_dim = property(lambda self: self.dim)
    hidden = 0
    out_dim = 0
    act = "relu"
    noise = "fixed"
    sigma = 1.0

    def sites(self):
        return [Site("x", (self.dim,))]


def uci_bnn(in_dim=13):
    """reference src/models/uci_bnn.py:21-68"""
    return BNNRegression(in_dim, 50, "relu", "gamma")


def reg_1d_bnn(hidden=5):
    """reference src/models/reg_1d_bnn.py:14-19 (low = 5, high = 100)"""
    return BNNRegression(1, hidden, "tanh", "fixed", 0.1, "nn_")


def mnist_mlp(in_dim=784, hidden=100, out_dim=10):
    """reference src/models/mlp.py:33-65"""
    return BNNClassifier(in_dim, hidden, out_dim)


@dataclass
class Layout:
    """Where every site lives: latent vector (model order) and raveled particle (sorted names)."""
    rows: list
    latent_dim: int
    particle_dim: int
    names: list = field(default_factory=list)   # sorted leaf names
    leaves: dict = field(default_factory=dict)  # name -> (offset, shape, site, which)


def layout(model, guide_kind):
    sites = model.sites()
    if len(sites) > _lib.SMI_MAX_SITES:
        raise ValueError("too many sites")
    lat, off = {}, 0
    for s in sites:
        lat[s.name] = off
        off += s.size
    L = off
    leaves = []
    for s in sites:
        leaves.append((s.name + "_auto_loc", s, "loc"))
        if guide_kind == _lib.GUIDE_NORMAL:
            leaves.append((s.name + "_auto_scale", s, "rho"))
    leaves.sort(key=lambda t: t[0])
    pos, table, off = {}, {}, 0
    for nm, s, which in leaves:
        pos[(s.name, which)] = off
        table[nm] = (off, s.shape, s, which)
        off += s.size
    rows = [dict(name=s.name, size=s.size, transform=s.transform, lat=lat[s.name],
                 loc=pos[(s.name, "loc")], rho=pos.get((s.name, "rho"), -1), shape=s.shape)
            for s in sites]
    return Layout(rows, L, off, [t[0] for t in leaves], table)


def make_desc(model, guide_kind):
    """SmiModelDesc for libsmi (include/smi.h)."""
    lay = layout(model, guide_kind)
    md = _lib.SmiModelDesc()
    md.kind = model.kind
    md.guide = guide_kind
    md.in_dim = int(model.in_dim)
    md.hidden = int(model.hidden)
    md.out_dim = int(model.out_dim)
    md.act = _lib.ACT_RELU if model.act == "relu" else _lib.ACT_TANH
    md.noise = _lib.NOISE_GAMMA_PREC if model.noise == "gamma" else _lib.NOISE_FIXED
    md.sigma = float(model.sigma)
    md.n_sites = len(lay.rows)
    md.latent_dim = lay.latent_dim
    md.particle_dim = lay.particle_dim
    for t, r in enumerate(lay.rows):
        md.sites[t].size = r["size"]
        md.sites[t].transform = r["transform"]
        md.sites[t].lat = r["lat"]
        md.sites[t].loc = r["loc"]
        md.sites[t].rho = r["rho"]
    return md, lay


def softplus_inv(v):
    return math.log(math.expm1(v))
