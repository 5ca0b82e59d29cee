"""CPU pins of the predictive / metrics restatement (oracle/predictive.py) against independent
formulas (reference src/metrics.py:222-268)."""
import numpy as np
from scipy.special import logsumexp

from oracle import models as om
from oracle import predictive as op


def _setup(guide="normal", I=4, m=3, n=7):
    spec = om.uci_bnn(I)
    rows, L, d = om.layout(spec, guide)
    rng = np.random.default_rng(0)
    psi = rng.normal(size=(m, d)) * 0.1
    for r in rows:
        if r["rho"] >= 0:
            psi[:, r["rho"]:r["rho"] + r["size"]] = np.log(np.expm1(0.1))
    return spec, rows, L, psi, rng.normal(size=(n, I)), rng.normal(size=n)


def test_metrics_match_scipy():
    spec, rows, L, psi, x, y = _setup()
    o = op.predictive_reg(spec, psi, [0, 5], 200, x, y)
    ll = o["loglik"]
    ref = logsumexp(ll, axis=0) - np.log(ll.shape[0])
    assert np.allclose(op.lpd(ll), ref, rtol=1e-12)
    assert abs(op.nll_reg(ll) + ref.mean()) < 1e-12 and abs(op.lppd(ll) - ref.sum()) < 1e-10
    assert abs(op.rmse(o["y_loc"], y) - np.sqrt(((y[None] - o["y_loc"]) ** 2).mean())) < 1e-12
    # log-likelihood is the Normal density with sigma = prec^-1/2 (src/models/uci_bnn.py:63-67)
    prec = o["latents"][:, rows[0]["lat"]]
    z = (y[None] - o["y_loc"]) * np.sqrt(prec)[:, None]
    assert np.allclose(ll, -0.5 * z * z + 0.5 * np.log(prec)[:, None] - 0.5 * np.log(2 * np.pi))


def test_draws_come_from_the_assigned_particle():
    spec, rows, L, psi, x, y = _setup()
    o = op.predictive_reg(spec, psi, [3, 9], 3000, x)
    a = o["assign"]
    assert a.min() == 0 and a.max() == 2
    assert np.all(np.abs(np.bincount(a, minlength=3) / 3000.0 - 1 / 3) < 0.03)        # uniform components
    w1 = rows[1]
    for j in range(3):   # unconstrained site: draws of component j centre on its loc with scale 0.1
        sel = o["latents"][a == j][:, w1["lat"]:w1["lat"] + w1["size"]]
        assert np.abs(sel.mean(0) - psi[j, w1["loc"]:w1["loc"] + w1["size"]]).max() < 0.02
        assert abs(sel.std(0).mean() - 0.1) < 0.005
    assert o["loglik"] is None
    # deterministic in the key, different across keys
    o2 = op.predictive_reg(spec, psi, [3, 9], 50, x)
    o3 = op.predictive_reg(spec, psi, [3, 9], 50, x)
    assert np.array_equal(o2["y_loc"], o3["y_loc"])
    assert not np.array_equal(o2["y_loc"], op.predictive_reg(spec, psi, [3, 10], 50, x)["y_loc"])


def test_delta_guide_is_one_draw_per_particle():
    spec, rows, L, psi, x, y = _setup("delta")
    o = op.predictive_reg(spec, psi, [1, 2], 3, x, y, guide="delta")
    assert list(o["assign"]) == [0, 1, 2]
    for r in rows:
        u = psi[:, r["loc"]:r["loc"] + r["size"]]
        want = np.exp(u) if r["transform"] == om.T_EXP else u
        assert np.allclose(o["latents"][:, r["lat"]:r["lat"] + r["size"]], want)
