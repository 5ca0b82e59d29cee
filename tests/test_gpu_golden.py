"""CUDA path against the committed golden vectors (float64 oracle outputs) -- one per BASELINE config.
Nothing here reads /root/reference; the vectors travel with the repo."""
import os

import numpy as np
import pytest
import torch

from test_golden import CASES, G, _rel, load_case

pytestmark = pytest.mark.gpu
RTOL = 1e-5     # north_star: <= 1e-5 relative (fp32) for forces, ELBO and updated particles


def _models():
    import smi_experiments_b200 as smi
    return {"c1_sine_bnn": smi.reg_1d_bnn(5), "c2_uci_bnn": smi.uci_bnn(13), "c3_gauss_svgd": smi.GaussianTarget(30),
            "c3_gauss_smi": smi.GaussianTarget(30), "c4_mlp_cls": smi.mnist_mlp(36, 12, 5)}


@pytest.mark.parametrize("name", sorted(CASES))
def test_stages_against_golden(cuda, name):
    """prng_step -> elbo_grad -> stein_force -> optimizer, each stage fed the golden inputs."""
    from smi_experiments_b200 import _lib, models, ops
    _, guide, optimizer = CASES[name]
    z, c = load_case(name)
    gk = _lib.GUIDE_NORMAL if guide == "normal" else _lib.GUIDE_DELTA
    md, lay = models.make_desc(_models()[name], gk)
    t = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    m, S, B = c["m"], c["S"], c["B"]
    # PRNG: key chain, noise, indices
    key = t(z["key"].view(np.int32))
    kout = torch.zeros(2, dtype=torch.int32, device=cuda)
    eps = torch.empty(m, S, c["L"], device=cuda) if guide == "normal" else None
    plate_only = name == "c4_mlp_cls"
    idx = torch.empty(m, S, B, dtype=torch.int32, device=cuda) if (B and not plate_only) else None
    ops.prng_step(md, key, kout, m, S, c["N"] if B else 0, B if idx is not None else 0, 0, m, eps, idx)
    assert (kout.cpu().numpy().view(np.uint32) == z["next_key"]).all()
    if idx is not None:
        assert (idx.cpu().numpy() == z["idx"]).all()                       # bit-exact particle indexing
    if eps is not None:
        np.testing.assert_allclose(eps.cpu().numpy(), z["eps"], rtol=2e-6, atol=2e-7)
    # ELBO gradients on the GOLDEN noise (identical inputs)
    lik = (c["N"] / B) if B else 1.0
    Gd, Ed = ops.elbo_grad(md, t(z["psi"]), t(c["eps"]), None if plate_only else t(z["idx"] if z["idx"].size else None),
                           t(c["x"]), t(c["y"]), S, B or (c["N"] or 0), lik, float(z["tau"]))
    assert _rel(Gd.cpu().numpy(), z["G"]) <= RTOL
    assert np.max(np.abs(Ed.cpu().numpy() - z["elbo"]) / np.maximum(np.abs(z["elbo"]), 1.0)) <= RTOL
    # Stein force on the golden scores
    F, loss, h = ops.stein_force(t(z["psi"]), t(z["G"].astype(np.float32)), alpha=float(z["alpha"]))
    assert abs(float(h) - float(z["h"])) <= RTOL * float(z["h"])
    assert _rel(F.cpu().numpy(), z["F"]) <= RTOL
    assert abs(float(loss) - float(z["loss"])) <= RTOL * float(z["loss"])
    # optimizer on the golden force
    x, a, b = t(z["psi"]), t(z["opt_a"]), t(z["opt_b"])
    step = torch.full((1,), int(z["step"]), dtype=torch.int32, device=cuda)
    Fg = t(z["F"].astype(np.float32))
    (ops.adam if optimizer == "adam" else ops.adagrad)(x, a, b, Fg, step, float(z["lr"]))
    assert _rel(x.cpu().numpy(), z["psi_new"]) <= RTOL
    assert _rel(a.cpu().numpy(), z["opt_a_new"]) <= RTOL and _rel(b.cpu().numpy(), z["opt_b_new"]) <= 2 * RTOL
    assert int(step.item()) == int(z["step"]) + 1


def test_stein_force_golden(cuda):
    from smi_experiments_b200 import ops
    z = np.load(os.path.join(G, "c5_stein_force.npz"))
    X, Gs = torch.from_numpy(z["X"]).to(cuda), torch.from_numpy(z["G"]).to(cuda)
    F, loss, h = ops.stein_force(X, Gs, alpha=float(z["alpha"]), path=ops.PATH_TC)
    assert abs(float(h) - float(z["h"])) <= RTOL * float(z["h"])
    assert _rel(F.cpu().numpy(), z["F"]) <= RTOL
    assert abs(float(loss) - float(z["loss"])) <= RTOL * float(z["loss"])
    # bit-exact bandwidth selection on the device's own distance matrix
    D2 = ops.sqdist(X, path=ops.PATH_TC)
    _, ranks = ops.median_bandwidth(D2)
    a = np.sort(D2.cpu().numpy().ravel())
    lo, hi = (int(v) for v in z["ranks"])
    assert (ranks.cpu().numpy().view(np.uint32) == np.array([a[lo], a[hi]], np.float32).view(np.uint32)).all()


def test_prng_golden(cuda):
    from smi_experiments_b200 import ops
    z = np.load(os.path.join(G, "prng.npz"))
    key = torch.from_numpy(z["key"].view(np.int32).copy()).to(cuda)
    out = torch.empty(7, 2, dtype=torch.int32, device=cuda)
    ops.call("smi_prng_split", ops._ptr(key), 7, ops._ptr(out), ops._stream())
    assert (out.cpu().numpy().view(np.uint32) == z["split7"]).all()
    bits = torch.empty(257, dtype=torch.int32, device=cuda)
    ops.call("smi_prng_bits_u32", ops._ptr(key), 257, ops._ptr(bits), ops._stream())
    assert (bits.cpu().numpy().view(np.uint32) == z["bits"]).all()
    u = torch.empty(257, device=cuda)
    ops.call("smi_prng_uniform_f32", ops._ptr(key), 257, -0.1, 0.1, ops._ptr(u), ops._stream())
    assert np.array_equal(u.cpu().numpy(), z["uniform"])
