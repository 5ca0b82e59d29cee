"""MixtureGuidePredictive + regression metrics on the device against the CPU
restatement of reference src/metrics.py:216-268.  Integer outputs bit-exact; floats <= 1e-5."""
import numpy as np
import pytest
import torch

from oracle import models as om
from oracle import predictive as op

pytestmark = pytest.mark.gpu
TOL = 1e-5


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _particles(smi, model, guide, m, seed):
    md, lay = smi.models.make_desc(model, guide.kind)
    g = torch.Generator().manual_seed(seed)
    psi = (torch.rand(m, lay.particle_dim, generator=g) - 0.5) * 0.4
    for r in lay.rows:
        if r["rho"] >= 0:
            psi[:, r["rho"]:r["rho"] + r["size"]] = float(np.log(np.expm1(0.1)))
    return psi


@pytest.mark.parametrize("which,S,n,layout", [("uci", 64, 23, "partitionable"), ("uci", 7, 1, "partitionable"),
                                              ("sine", 40, 40, "partitionable"), ("uci", 64, 23, "legacy"),
                                              ("sine", 7, 40, "legacy")])
def test_mixture_predictive_matches_oracle(cuda, which, S, n, layout):
    import smi_experiments_b200 as smi
    from oracle import prng as oprng
    prev = smi.set_prng_layout(layout)
    try:
        _mixture_predictive_case(cuda, smi, which, S, n,
                                 oprng.LAYOUT_LEGACY if layout == "legacy" else oprng.LAYOUT_PARTITIONABLE)
    finally:
        smi.set_prng_layout(prev)


def _mixture_predictive_case(cuda, smi, which, S, n, layout_mode):
    if which == "uci":
        model, spec, I = smi.uci_bnn(6), om.uci_bnn(6), 6
    else:
        model, spec, I = smi.reg_1d_bnn(5), om.reg_1d_bnn(5), 1
    guide = smi.AutoNormal(model)
    psi = _particles(smi, model, guide, 5, 1)
    rng = np.random.default_rng(2)
    x = rng.standard_normal((n, I)).astype(np.float32)
    y = rng.standard_normal(n).astype(np.float32)

    class St:  # state-like carrier of the raveled particles
        class optim_state:
            pass
    St.optim_state.x = psi
    pred = smi.MixtureGuidePredictive(model, guide, St, num_samples=S)
    out = pred(smi.PRNGKey(11), x, y)
    ref = op.predictive_reg(spec, psi.numpy().astype(np.float64), [0, 11], S, x.astype(np.float64),
                            y.astype(np.float64), layout_mode=layout_mode)
    assert np.array_equal(out["_assign"].cpu().numpy(), ref["assign"])                    # bit-exact
    lat = torch.cat([out[s.name].reshape(S, -1) for s in model.sites()], 1).cpu().numpy()
    assert rel(lat, ref["latents"]) <= TOL
    assert rel(out["y_loc"].cpu().numpy(), ref["y_loc"]) <= TOL
    assert rel(out["y"].cpu().numpy(), ref["y"]) <= TOL
    assert rel(out["_loglik"].cpu().numpy(), ref["loglik"]) <= TOL
    assert abs(float(smi.rmse_fn(out, y)) - op.rmse(ref["y_loc"], y)) <= TOL * op.rmse(ref["y_loc"], y)
    nll, nll_ref = float(smi.nll_reg_fn(out)), op.nll_reg(ref["loglik"])
    assert abs(nll - nll_ref) <= TOL * max(abs(nll_ref), 1.0)
    lp, lp_ref = float(smi.lppd_fn(out)), op.lppd(ref["loglik"])
    assert abs(lp - lp_ref) <= TOL * max(abs(lp_ref), 1.0)


def test_delta_predictive_and_params_roundtrip(cuda):
    """svgd_sampler (src/metrics.py:227-237): one draw per particle; params given by name as
    engine.get_params returns them (constrained)."""
    import smi_experiments_b200 as smi
    model, spec = smi.uci_bnn(4), om.uci_bnn(4)
    eng = smi.SVGD(model, smi.Adam(1e-2), smi.RBFKernel(), num_stein_particles=6)
    rng = np.random.default_rng(0)
    x = rng.standard_normal((30, 4)).astype(np.float32)
    y = rng.standard_normal(30).astype(np.float32)
    st = eng.init(smi.PRNGKey(0), x=x, y=y, subsample=10)
    for _ in range(3):
        st, _ = eng.update(st, x=x, y=y, subsample=10)
    pred = smi.Predictive(model, eng.guide, eng.get_params(st), num_samples=1, return_sites=["y_loc", "y"])
    out = pred(smi.PRNGKey(5), x, y)
    assert set(out) == {"y_loc", "y", "_assign", "_loglik"} and out["y_loc"].shape == (6, 30)
    ref = op.predictive_reg(spec, st.optim_state.x.cpu().numpy().astype(np.float64), [0, 5], 6, x.astype(np.float64),
                            y.astype(np.float64), guide="delta")
    assert rel(out["y_loc"].cpu().numpy(), ref["y_loc"]) <= TOL
    assert rel(out["_loglik"].cpu().numpy(), ref["loglik"]) <= TOL


def test_metrics_at_reference_size_against_torch(cuda):
    """5000 draws (experiments/uci_benchmark.py:104) x 51 test rows: the fixed-order device
    reductions agree with float64 torch reductions of the same device outputs."""
    import smi_experiments_b200 as smi
    model = smi.uci_bnn(13)
    guide = smi.AutoNormal(model)
    psi = _particles(smi, model, guide, 20, 3)

    class St:
        class optim_state:
            pass
    St.optim_state.x = psi
    rng = np.random.default_rng(4)
    x = rng.standard_normal((51, 13)).astype(np.float32)
    y = rng.standard_normal(51).astype(np.float32)
    out = smi.MixtureGuidePredictive(model, guide, St, num_samples=5000)(smi.PRNGKey(1), x, y)
    yl, ll = out["y_loc"].double(), out["_loglik"].double()
    yt = torch.from_numpy(y).to(yl.device).double()
    r_ref = float(torch.sqrt(((yt[None] - yl) ** 2).mean()))
    assert abs(float(smi.rmse_fn(out, y)) - r_ref) <= 1e-6 * r_ref
    lpd = torch.logsumexp(ll, 0) - np.log(5000.0)
    assert abs(float(smi.nll_reg_fn(out)) + float(lpd.mean())) <= 1e-6 * abs(float(lpd.mean()))
    assert abs(float(smi.lppd_fn(out)) - float(lpd.sum())) <= 1e-6 * abs(float(lpd.sum()))
    counts = torch.bincount(out["_assign"].long(), minlength=20).cpu().numpy()
    assert counts.sum() == 5000 and counts.min() > 150          # uniform over the 20 components


@pytest.mark.parametrize("P,n,C,scale", [(3, 7, 4, 0.3), (5, 400, 10, 2.0), (2, 10000, 10, 4.0)])
def test_classification_metrics_match_oracle(cuda, P, n, C, scale):
    """nll / acc / brier / ece / mce / conf (src/metrics.py:17-160) incl. binned_conf_acc's empty-bin
    behaviour; 10000 x 10 is the MNIST test-set shape."""
    import smi_experiments_b200 as smi
    rng = np.random.default_rng(n + C)
    logits = (rng.standard_normal((P, n, C)) * scale).astype(np.float32)
    y = rng.integers(0, C, n).astype(np.int32)
    ypred = rng.integers(0, C, (P, n)).astype(np.int32)
    got = smi.classification_metrics({"logits": torch.from_numpy(logits).to(cuda), "y": torch.from_numpy(ypred)}, y)
    ref = op.classification_metrics(logits.astype(np.float64), y, ypred)
    for i, k in enumerate(("nll", "acc", "brier", "ece", "mce", "conf")):
        g = got[k].cpu().numpy().astype(np.float64)
        assert np.all(np.abs(g - ref[:, i]) <= TOL * np.maximum(np.abs(ref[:, i]), 1e-2)), (k, g, ref[:, i])
    # without the sampled y site: acc / conf undefined, the rest unchanged
    got2 = smi.classification_metrics({"logits": torch.from_numpy(logits).to(cuda)}, y)
    assert torch.isnan(got2["acc"]).all() and torch.equal(got2["ece"], got["ece"])


@pytest.mark.parametrize("I,H,O,S,n,m", [(12, 8, 3, 5, 37, 3), (70, 20, 5, 4, 300, 2), (784, 100, 10, 3, 130, 2)])
def test_classifier_predictive_matches_oracle(cuda, I, H, O, S, n, m):
    """logits through the stacked tcgen05 first-layer GEMM + per-draw head; y by Gumbel-max.  Integer
    outputs (assign, y) bit-exact wherever the winning margin exceeds the fp32 logit error."""
    import smi_experiments_b200 as smi
    model, spec = smi.mnist_mlp(I, H, O), om.mnist_mlp(I, H, O)
    guide = smi.AutoNormal(model)
    psi = _particles(smi, model, guide, m, 7)

    class St:
        class optim_state:
            pass
    St.optim_state.x = psi
    rng = np.random.default_rng(5)
    x = rng.standard_normal((n, I)).astype(np.float32)
    out = smi.MixtureGuidePredictive(model, guide, St, num_samples=S, return_sites=["logits", "y"])(smi.PRNGKey(9), x)
    ref = op.predictive_cls(spec, psi.numpy().astype(np.float64), [0, 9], S, x.astype(np.float64))
    assert np.array_equal(out["_assign"].cpu().numpy(), ref["assign"])
    lg = out["logits"].cpu().numpy()
    assert lg.shape == (S, n, O) and rel(lg, ref["logits"]) <= TOL
    # Gumbel-max draw: recompute the float64 perturbed scores; rows whose top-2 margin is above 1e-4 must agree
    yg = out["y"].cpu().numpy()
    assert yg.shape == (S, n) and yg.dtype == np.int32
    from oracle import prng as oprng
    km = oprng.split(np.asarray([0, 9], np.uint32), 3)[2]
    nk = oprng.split(km, S)
    tiny = np.float32(1.17549435e-38)
    for s in range(S):
        f = oprng.bits_to_unit_float(oprng.random_bits(nk[s], n * O)).astype(np.float32)
        g = -np.log(-np.log(np.maximum(tiny, f + tiny).astype(np.float64))).reshape(n, O)
        sc = ref["logits"][s] + g
        top2 = np.sort(sc, -1)[:, -2:]
        clear = (top2[:, 1] - top2[:, 0]) > 1e-4
        assert clear.mean() > 0.99
        assert np.array_equal(yg[s][clear], ref["y"][s][clear])
    # metrics straight from the predictive output
    y_true = rng.integers(0, O, n).astype(np.int32)
    met = smi.classification_metrics(out, y_true)
    refm = op.classification_metrics(lg.astype(np.float64), y_true, yg)
    for i, k in enumerate(("nll", "acc", "brier", "ece", "mce", "conf")):
        assert np.all(np.abs(met[k].cpu().numpy() - refm[:, i]) <= 1e-5 * np.maximum(np.abs(refm[:, i]), 1e-2)), k
