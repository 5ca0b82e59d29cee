"""Pins for the PRNG and optimizer oracles.  threefry2x32: Random123 known-answer vectors."""
import numpy as np
import torch

from oracle import optim as ooptim
from oracle import prng
from oracle import step as ostep
from oracle import models as omodels


def _tf(k0, k1, c0, c1):
    o0, o1 = prng.threefry2x32((np.uint32(k0), np.uint32(k1)), np.array([c0], np.uint32), np.array([c1], np.uint32))
    return int(o0[0]), int(o1[0])


def test_threefry_random123_kat():
    assert _tf(0, 0, 0, 0) == (0x6B200159, 0x99BA4EFE)
    assert _tf(0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF) == (0x1CB996FC, 0xBB002BE7)
    assert _tf(0x13198A2E, 0x03707344, 0x243F6A88, 0x85A308D3) == (0xC4923A9C, 0x483DF7A0)


def test_split_layouts_and_shapes():
    key = np.array([0, 42], np.uint32)
    for lay in (prng.LAYOUT_LEGACY, prng.LAYOUT_PARTITIONABLE):
        ks = prng.split(key, 5, lay)
        assert ks.shape == (5, 2) and ks.dtype == np.uint32
        assert len({tuple(k) for k in ks}) == 5
    # partitionable: split(key, n)[j] does not depend on n
    assert (prng.split(key, 3)[:2] == prng.split(key, 2)).all()


def test_uniform_and_normal_statistics():
    key = np.array([1, 2], np.uint32)
    u = prng.uniform(key, 200_000, -0.1, 0.1)
    assert u.min() >= -0.1 and u.max() < 0.1 and abs(u.mean()) < 1e-3
    z = prng.normal(key, 200_000)
    assert abs(z.mean()) < 1e-2 and abs(z.std() - 1) < 1e-2 and np.isfinite(z).all()
    # erfinv polynomial against scipy
    from scipy.special import erfinv
    x = np.linspace(-0.999, 0.999, 1001).astype(np.float32)
    np.testing.assert_allclose(prng.erfinv32(x), erfinv(x.astype(np.float64)), rtol=2e-6, atol=1e-7)


def test_subsample_is_partial_permutation():
    key = np.array([3, 4], np.uint32)
    idx = ostep.subsample_indices(key, 57, 10)
    assert idx.shape == (10,) and len(set(idx.tolist())) == 10 and idx.min() >= 0 and idx.max() < 57
    assert (ostep.subsample_indices(key, 57, 57).tolist() != list(range(57)))
    assert sorted(ostep.subsample_indices(key, 57, 57).tolist()) == list(range(57))


def test_draw_noise_shapes_and_key_chain():
    om = omodels.uci_bnn(3)
    key = np.array([0, 7], np.uint32)
    nk, att = ostep.draw_keys(key)
    assert (nk == prng.split(key, 3)[0]).all()
    eps, idx = ostep.draw_noise(att, 2, 3, om, "normal", 20, 5)
    assert eps.shape == (2, 3, om.latent_dim) and idx.shape == (2, 3, 5)
    eps2, _ = ostep.draw_noise(att, 2, 3, om, "normal", None, None)
    assert not np.allclose(eps, eps2)      # the plate consumes the first subkey


def test_adam_matches_torch_optim():
    rng = np.random.default_rng(0)
    x0 = rng.standard_normal(50)
    xt = torch.tensor(x0, dtype=torch.float64, requires_grad=True)
    opt = torch.optim.Adam([xt], lr=1e-2, betas=(0.9, 0.999), eps=1e-8)
    x, m, v = x0.copy(), np.zeros(50), np.zeros(50)
    for it in range(5):
        g = rng.standard_normal(50)
        xt.grad = torch.tensor(g)
        opt.step()
        x, m, v = ooptim.adam_update(it, g, x, m, v, 1e-2)
    # torch puts eps outside the bias-corrected sqrt like jax: sqrt(vhat) + eps
    np.testing.assert_allclose(x, xt.detach().numpy(), rtol=1e-10, atol=1e-12)


def test_adagrad_hand_pin():
    x, acc, mom = np.array([1.0, 2.0, 3.0]), np.zeros(3), np.zeros(3)
    g = np.array([0.5, 0.0, -2.0])
    x, acc, mom = ooptim.adagrad_update(0, g, x, acc, mom, 0.8)
    # acc = g^2; mom = 0.1 * g / |g| (0 where g == 0); x -= 0.8 * mom
    np.testing.assert_allclose(acc, [0.25, 0.0, 4.0])
    np.testing.assert_allclose(mom, [0.1, 0.0, -0.1])
    np.testing.assert_allclose(x, [1.0 - 0.08, 2.0, 3.0 + 0.08])
    x, acc, mom = ooptim.adagrad_update(1, g, x, acc, mom, 0.8)
    np.testing.assert_allclose(mom, [0.1 * 0.5 / np.sqrt(0.5) + 0.09, 0.0, -(0.1 * 2 / np.sqrt(8.0) + 0.09)])
