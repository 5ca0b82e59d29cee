"""Optimizer (row a8) and PRNG (row a9) kernels against the oracle, through the C ABI."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import optim as ooptim
from oracle import prng as oprng
from oracle import step as ostep
from oracle import models as omodels

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 3, 4, 1023, 6604 * 20, 1_000_003])
def test_adam_matches_oracle(cuda, n):
    from smi_experiments_b200 import ops
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    m0 = np.zeros(n, np.float32)
    v0 = np.zeros(n, np.float32)
    xt, mt, vt = (torch.from_numpy(a.copy()).to(cuda) for a in (x, m0, v0))
    step = torch.zeros(1, dtype=torch.int32, device=cuda)
    xo, mo, vo = x.copy(), m0.copy(), v0.copy()
    for it in range(3):
        F = rng.standard_normal(n).astype(np.float32)
        ops.adam(xt, mt, vt, torch.from_numpy(F).to(cuda), step, 1e-2)
        xo, mo, vo = ooptim.adam_update(it, -F, xo, mo, vo, 1e-2)
    assert int(step.item()) == 3
    # elementwise: device FMA contraction vs numpy's separate roundings -> a few ulp of the larger
    # operand (atol), 1e-5 relative otherwise (north_star tolerance)
    np.testing.assert_allclose(xt.cpu().numpy(), xo, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(mt.cpu().numpy(), mo, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(vt.cpu().numpy(), vo, rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("n", [2, 4097, 179220 * 10])
def test_adagrad_matches_oracle(cuda, n):
    from smi_experiments_b200 import ops
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    xt = torch.from_numpy(x.copy()).to(cuda)
    at, mt = torch.zeros_like(xt), torch.zeros_like(xt)
    step = torch.zeros(1, dtype=torch.int32, device=cuda)
    xo, ao, mo = x.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)
    for it in range(3):
        F = rng.standard_normal(n).astype(np.float32)
        F[::7] = 0.0     # zero gradients exercise the where(acc > 0, ., 0) branch
        ops.adagrad(xt, at, mt, torch.from_numpy(F).to(cuda), step, 0.8)
        xo, ao, mo = ooptim.adagrad_update(it, -F, xo, ao, mo, 0.8)
    np.testing.assert_allclose(xt.cpu().numpy(), xo, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(at.cpu().numpy(), ao, rtol=1e-5)
    np.testing.assert_allclose(mt.cpu().numpy(), mo, rtol=1e-5, atol=1e-7)


def _key(dev, k):
    return torch.from_numpy(np.asarray(k, np.uint32).view(np.int32).copy()).to(dev)


@pytest.fixture(params=["partitionable", "legacy"])
def layout(request):
    """Both jax_threefry_partitionable stream layouts (row n2); restores the default afterwards."""
    from smi_experiments_b200 import ops
    prev = ops.set_prng_layout(request.param)
    yield oprng.LAYOUT_PARTITIONABLE if request.param == "partitionable" else oprng.LAYOUT_LEGACY
    ops.set_prng_layout(prev)


def test_prng_primitives_bit_exact(cuda, lib, layout):
    from smi_experiments_b200 import ops
    key = np.array([0x13198a2e, 0x03707344], np.uint32)
    kt = _key(cuda, key)
    out = torch.empty(7, 2, dtype=torch.int32, device=cuda)
    ops.call("smi_prng_split", ops._ptr(kt), 7, ops._ptr(out), ops._stream())
    assert (out.cpu().numpy().view(np.uint32) == oprng.split(key, 7, layout)).all()
    n = 100_003
    bits = torch.empty(n, dtype=torch.int32, device=cuda)
    ops.call("smi_prng_bits_u32", ops._ptr(kt), n, ops._ptr(bits), ops._stream())
    assert (bits.cpu().numpy().view(np.uint32) == oprng.random_bits(key, n, layout)).all()
    u = torch.empty(n, dtype=torch.float32, device=cuda)
    ops.call("smi_prng_uniform_f32", ops._ptr(kt), n, -0.1, 0.1, ops._ptr(u), ops._stream())
    assert (u.cpu().numpy() == oprng.uniform(key, n, -0.1, 0.1, layout)).all()
    z = torch.empty(n, dtype=torch.float32, device=cuda)
    ops.call("smi_prng_normal_f32", ops._ptr(kt), n, ops._ptr(z), ops._stream())
    # erfinv polynomial: device uses fused multiply-adds, numpy does not -> a few ulp
    np.testing.assert_allclose(z.cpu().numpy(), oprng.normal(key, n, layout), rtol=2e-6, atol=2e-7)
    # an even word count takes the other branch of the legacy halves
    ops.call("smi_prng_bits_u32", ops._ptr(kt), 1000, ops._ptr(bits), ops._stream())
    assert (bits[:1000].cpu().numpy().view(np.uint32) == oprng.random_bits(key, 1000, layout)).all()


@pytest.mark.parametrize("case", ["uci", "gauss_delta", "cls"])
def test_prng_step_chain(cuda, case, layout):
    """Key chain of one update: indices bit-exact, normals to a few ulp, next key bit-exact."""
    from smi_experiments_b200 import models, ops, _lib
    if case == "uci":
        pm, om, guide, gk = models.uci_bnn(3), omodels.uci_bnn(3), "normal", _lib.GUIDE_NORMAL
        m, S, N, B = 3, 4, 57, 10
    elif case == "cls":
        pm, om, guide, gk = models.mnist_mlp(6, 4, 3), omodels.mnist_mlp(6, 4, 3), "normal", _lib.GUIDE_NORMAL
        m, S, N, B = 2, 3, 1000, None
    else:
        pm, om, guide, gk = models.GaussianTarget(5), omodels.gaussian(5), "delta", _lib.GUIDE_DELTA
        m, S, N, B = 4, 1, 0, None
    md, lay = models.make_desc(pm, gk)
    key = np.array([7, 42], np.uint32)
    next_key, attractive = ostep.draw_keys(key, layout)
    kin, kout = _key(cuda, key), torch.zeros(2, dtype=torch.int32, device=cuda)
    eps = torch.empty(m, S, lay.latent_dim, dtype=torch.float32, device=cuda) if guide == "normal" else None
    idx = torch.empty(m, S, B, dtype=torch.int32, device=cuda) if B else None
    ops.prng_step(md, kin, kout, m, S, N, B or 0, 0, m, eps, idx)
    assert (kout.cpu().numpy().view(np.uint32) == next_key).all()
    if case == "cls":
        # the plate consumes a subkey although no indices are produced: emulate with b = 1
        e_ref, _ = ostep.draw_noise(attractive, m, S, om, guide, N, 1, layout)
    else:
        e_ref, i_ref = ostep.draw_noise(attractive, m, S, om, guide, N if B else None, B, layout)
    if idx is not None:
        assert (idx.cpu().numpy() == i_ref).all()
        assert all(len(set(r)) == B for r in idx.cpu().numpy().reshape(-1, B))
    if eps is not None:
        np.testing.assert_allclose(eps.cpu().numpy(), e_ref, rtol=2e-6, atol=2e-7)


def test_prng_step_row_sharding(cuda):
    """Rows [row0, row0+nrows) of a sharded call equal the same rows of the full call."""
    from smi_experiments_b200 import models, ops, _lib
    md, lay = models.make_desc(models.uci_bnn(3), _lib.GUIDE_NORMAL)
    m, S, N, B = 5, 3, 40, 8
    kin, kout = _key(cuda, [1, 2]), torch.zeros(2, dtype=torch.int32, device=cuda)
    eps = torch.empty(m, S, lay.latent_dim, device=cuda)
    idx = torch.empty(m, S, B, dtype=torch.int32, device=cuda)
    ops.prng_step(md, kin, kout, m, S, N, B, 0, m, eps, idx)
    e2 = torch.empty(2, S, lay.latent_dim, device=cuda)
    i2 = torch.empty(2, S, B, dtype=torch.int32, device=cuda)
    ops.prng_step(md, kin, kout, m, S, N, B, 2, 2, e2, i2)
    assert torch.equal(e2, eps[2:4]) and torch.equal(i2, idx[2:4])


@pytest.mark.parametrize("rule,N,B", [("jax", 91, 17), ("choice", 91, 17), ("choice", 3000, 64)])
@pytest.mark.parametrize("layout_name", ["partitionable", "legacy"])
def test_subsample_rules(cuda, layout_name, rule, N, B):
    """Switches: j_t drawn as jax.random.randint does (two words, modular) instead of one-word multiply-shift;
    sort-based permutation (random.choice without replacement; 3000 rows need two sort rounds)."""
    from smi_experiments_b200 import models, ops, _lib
    pm, om_ = models.uci_bnn(3), omodels.uci_bnn(3)
    md, lay = models.make_desc(pm, _lib.GUIDE_NORMAL)
    m, S = 2, 3
    key = np.array([5, 77], np.uint32)
    lm = oprng.LAYOUT_PARTITIONABLE if layout_name == "partitionable" else oprng.LAYOUT_LEGACY
    prev_l, prev_r = ops.set_prng_layout(layout_name), ops.set_randint_rule(rule)
    try:
        kin, kout = _key(cuda, key), torch.zeros(2, dtype=torch.int32, device=cuda)
        idx = torch.empty(m, S, B, dtype=torch.int32, device=cuda)
        ops.prng_step(md, kin, kout, m, S, N, B, 0, m, None, idx)
        _, attractive = ostep.draw_keys(key, lm)
        _, i_ref = ostep.draw_noise(attractive, m, S, om_, "normal", N, B, lm, randint=rule)
        got = idx.cpu().numpy()
        assert (got == i_ref).all()
        assert all(len(set(r)) == B and 0 <= min(r) and max(r) < N for r in got.reshape(-1, B))
        _, i_ms = ostep.draw_noise(attractive, m, S, om_, "normal", N, B, lm)
        assert (got != i_ms).any()
    finally:
        ops.set_prng_layout(prev_l)
        ops.set_randint_rule(prev_r)


# ---- external pin: the DEVICE PRNG against values printed in JAX's documentation (tests/jax_published.py) ----
def _dev_split(cuda, key, n):
    from smi_experiments_b200 import ops
    out = torch.empty(n, 2, dtype=torch.int32, device=cuda)
    ops.call("smi_prng_split", ops._ptr(_key(cuda, key)), n, ops._ptr(out), ops._stream())
    return out.cpu().numpy().view(np.uint32)


def _with_layout(code):
    from smi_experiments_b200 import ops
    return ops.set_prng_layout("partitionable" if code == 1 else "legacy")


def test_device_prng_matches_jax_docs(cuda, lib):
    import jax_published as jp
    from smi_experiments_b200 import ops
    prev = ops.set_prng_layout(None)
    try:
        for layout, seed, n, want in jp.SPLITS:
            _with_layout(layout)
            assert _dev_split(cuda, jp.key_of(seed), n).tolist() == want
        for layout, seed, path, n, want in jp.NORMALS:
            _with_layout(layout)
            key = jp.key_of(seed)
            for nsplit, i in path:
                key = _dev_split(cuda, key, nsplit)[i]
            z = torch.empty(n, dtype=torch.float32, device=cuda)
            ops.call("smi_prng_normal_f32", ops._ptr(_key(cuda, key)), n, ops._ptr(z), ops._stream())
            np.testing.assert_allclose(z.cpu().numpy(), np.asarray(want, np.float32), rtol=jp.NORMAL_RTOL, atol=0)
        for layout, seed, want in jp.UNIFORMS:
            _with_layout(layout)
            u = torch.empty(1, dtype=torch.float32, device=cuda)
            ops.call("smi_prng_uniform_f32", ops._ptr(_key(cuda, jp.key_of(seed))), 1, 0.0, 1.0, ops._ptr(u), ops._stream())
            assert float(u.item()) == float(np.float32(want))
    finally:
        ops.set_prng_layout(prev)
