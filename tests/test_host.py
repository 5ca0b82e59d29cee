"""Host-side logic of the engine mirror that needs no GPU: layouts, descriptors, schedules."""
from functools import partial

import numpy as np
import pytest

from oracle import models as omodels


def test_layout_matches_oracle_for_every_reference_model():
    from smi_experiments_b200 import _lib, models
    pairs = [(models.uci_bnn(13), omodels.uci_bnn(13)), (models.reg_1d_bnn(5), omodels.reg_1d_bnn(5)),
             (models.reg_1d_bnn(100), omodels.reg_1d_bnn(100)), (models.mnist_mlp(), omodels.mnist_mlp()),
             (models.GaussianTarget(1000), omodels.gaussian(1000))]
    for pm, om in pairs:
        for gk, gname in ((_lib.GUIDE_NORMAL, "normal"), (_lib.GUIDE_DELTA, "delta")):
            md, lay = models.make_desc(pm, gk)
            rows, L, d = omodels.layout(om, gname)
            assert (md.latent_dim, md.particle_dim) == (L, d) == (lay.latent_dim, lay.particle_dim)
            for t, r in enumerate(rows):
                s = md.sites[t]
                assert (s.size, s.transform, s.lat, s.loc, s.rho) == (r["size"], r["transform"], r["lat"], r["loc"], r["rho"])


def test_config_dimensions():
    """SURVEY.md section 8 config table."""
    from smi_experiments_b200 import _lib, models
    dims = lambda m: models.make_desc(m, _lib.GUIDE_NORMAL)[1]
    assert (dims(models.uci_bnn(13)).latent_dim, dims(models.uci_bnn(13)).particle_dim) == (3302, 6604)
    assert dims(models.mnist_mlp()).particle_dim == 179220
    assert dims(models.reg_1d_bnn(50)).particle_dim == 5402
    names = dims(models.uci_bnn(13)).names
    assert names == sorted(names) and names[0] == "b1_auto_loc" and names[-1] == "w3_auto_scale"


def test_init_radius_forms():
    from smi_experiments_b200 import init_to_uniform
    from smi_experiments_b200.steinvi import _radius
    assert _radius(partial(init_to_uniform, radius=0.1)) == 0.1       # experiments/uci_benchmark.py:179
    assert _radius(init_to_uniform(radius=20)) == 20.0
    assert _radius(init_to_uniform) == 2.0 and _radius(None) == 2.0


def test_cyclical_annealing_schedule():
    """ASVGD._cyclical_annealing mirrored at src/engine_run.py:16-18."""
    from smi_experiments_b200 import ASVGD
    T, Cn, p = 100, 10, 10
    cyc = ASVGD._cyclical_annealing(T, Cn, p)
    norm, L = (T + 1) / Cn, T // Cn
    for t in (1, 5, 10, 11, 55, 89):
        assert cyc(t) == pytest.approx(((((t - 1) % L) + 1) / norm) ** p)
    assert cyc(90) == 1 and cyc(100) == 1          # last cycle holds the temperature at 1
    assert 0 < cyc(1) < cyc(5) < cyc(10) <= 1


def test_engine_constructor_surface():
    """Constructors keep the reference's signatures; nothing touches the GPU before init()."""
    import smi_experiments_b200 as smi
    model = smi.uci_bnn(13)
    guide = smi.AutoNormal(model, init_loc_fn=partial(smi.init_to_uniform, radius=0.1))
    e = smi.SteinVI(model, guide, smi.Adam(5e-3), smi.RBFKernel(), num_stein_particles=5, num_elbo_particles=100,
                    device="cpu")
    assert e.num_stein_particles == 5 and e.loss_temperature == 1.0 and e.repulsion_temperature == 1.0
    assert e.model is model and e.guide is guide and len(e.guide_sites) == 14
    s = smi.SVGD(model, smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=20, repulsion_temperature=10.0,
                 guide_kwargs={"init_loc_fn": partial(smi.init_to_uniform, radius=0.1)}, device="cpu")
    assert s.num_elbo_particles == 1 and s.loss_temperature == pytest.approx(1 / 20) and s.repulsion_temperature == 10.0
    assert s.layout.particle_dim == 3302
    b1 = smi.SVGD(model, smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=20, compat_b1=True, device="cpu")
    assert b1.num_stein_particles == 10             # the wrapper's bug B1 (src/methods/svgd.py:15-22)
    a = smi.ASVGD(model, smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=5, num_cycles=4, trans_speed=7, device="cpu")
    assert (a.num_cycles, a.trans_speed, a.num_stein_particles) == (4, 7, 5)
    with pytest.raises(TypeError):
        smi.SteinVI(lambda x: x, guide, smi.Adam(1e-3), smi.RBFKernel())
    with pytest.raises(NotImplementedError):
        smi.RBFKernel(mode="vector")


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: operators raise on host tensors instead of computing there."""
    import torch
    from smi_experiments_b200 import ops
    with pytest.raises(ValueError):
        ops.stein_force(torch.zeros(4, 4), torch.zeros(4, 4))
