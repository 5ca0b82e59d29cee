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


def test_binned_conf_acc_transcription_and_closed_form():
    """The literal transcription of binned_conf_acc (src/metrics.py:270-321) agrees with the per-bin
    closed form the device kernel uses: plain bin means for occupied bins that follow an occupied bin,
    and the cumulative-sum artefacts around empty bins."""
    rng = np.random.default_rng(0)
    for n, C, scale in ((50, 3, 1.0), (400, 10, 3.0), (7, 4, 0.2)):
        logits = rng.standard_normal((n, C)) * scale
        y = rng.integers(0, C, n)
        mc, ma = op.binned_conf_acc(logits, y, 100)
        sm = np.exp(logits - logits.max(-1, keepdims=True))
        sm /= sm.sum(-1, keepdims=True)
        conf, acc = sm.max(-1), (sm.argmax(-1) == y).astype(float)
        edges = (np.arange(101, dtype=np.float32) / np.float32(100)).astype(np.float64)
        k = np.clip(np.searchsorted(edges, conf, side="right") - 1, 0, 99)
        order = np.lexsort((np.arange(n), conf))
        T_c = T_a = prev_c = prev_a = 0.0
        for b in range(100):
            sel = k == b
            T_c, T_a = T_c + conf[sel].sum(), T_a + acc[sel].sum()
            cum_c, cum_a = T_c, T_a
            if not sel.any():
                later = [i for i in order if k[i] > b]
                if later:
                    cum_c, cum_a = cum_c + conf[later[0]], cum_a + acc[later[0]]
            want_c, want_a = cum_c - prev_c, cum_a - prev_a
            if sel.any():
                want_c, want_a = want_c / sel.sum(), want_a / sel.sum()
            prev_c, prev_a = cum_c, cum_a
            assert abs(mc[b] - want_c) < 1e-9 and abs(ma[b] - want_a) < 1e-9, (n, b)
    # all rows in one bin and correct: perfectly calibrated except for the empty-bin artefact before it
    logits = np.tile(np.array([[10.0, 0.0, 0.0]]), (20, 1))
    m = op.classification_metrics(logits[None], np.zeros(20, int), np.zeros((1, 20), int))
    assert abs(m[0, 1] - 1.0) < 1e-12 and m[0, 0] < 1e-3 and m[0, 2] < 1e-6


def test_classifier_predictive_oracle_sanity():
    spec = om.mnist_mlp(12, 8, 3)
    rows, L, d = om.layout(spec, "normal")
    rng = np.random.default_rng(1)
    psi = rng.normal(size=(4, d)) * 0.3
    for r in rows:
        psi[:, r["rho"]:r["rho"] + r["size"]] = np.log(np.expm1(0.05))
    x = rng.normal(size=(200, 12))
    o = op.predictive_cls(spec, psi, [0, 3], 6, x)
    assert o["logits"].shape == (6, 200, 3) and o["y"].shape == (6, 200) and o["y"].min() >= 0 and o["y"].max() <= 2
    # the Gumbel-max draw follows softmax(logits): agreement with argmax well above chance, below 1
    agree = (o["y"] == o["logits"].argmax(-1)).mean()
    sm = np.exp(o["logits"] - o["logits"].max(-1, keepdims=True))
    sm /= sm.sum(-1, keepdims=True)
    assert abs(agree - sm.max(-1).mean()) < 0.06
    o2 = op.predictive_cls(spec, psi, [0, 3], 6, x)
    assert np.array_equal(o["y"], o2["y"]) and np.array_equal(o["logits"], o2["logits"])
