"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's posterior predictive and
regression metrics.

Follows the reference's samplers and metric functions:
  src/metrics.py:216-224  smi_sampler  -> MixtureGuidePredictive(model, guide, params, num_samples,
                                          return_sites, guide_sites)        [numpyro, not in the tree]
  src/metrics.py:227-237  svgd_sampler -> Predictive(..., num_samples=1, batch_ndims=1)
  src/metrics.py:249-253  rmse_fn      sqrt(mean((y[None] - y_loc) ** 2))
  src/metrics.py:256-263  nll_reg_fn   -(logsumexp(lls, 0) - log S).mean()
  src/metrics.py:222-246  lppd_fn      sum_i (logsumexp_s ll_si - log S)
MixtureGuidePredictive lives in numpyro.contrib.einstein (floating master, absent here): its
published algorithm -- draw a component uniformly, sample that particle's guide, run the model --
is restated with the key chain documented in smi_experiments_b200/csrc/predictive.cu.
Parity unpinned (no reference fixture exists for predictive draws).
"""
import numpy as np
import torch

from . import models, prng


def softplus(x):
    return np.maximum(x, 0) + np.log1p(np.exp(-np.abs(x)))


def predictive_reg(spec, psi, key, num_samples, x, y=None, guide="normal", dtype=np.float64,
                   layout_mode=prng.LAYOUT_PARTITIONABLE):
    """dict(assign (S,), latents (S, L), y_loc (S, n), y (S, n), loglik (S, n) | None)."""
    rows, L, d = models.layout(spec, guide)
    psi = np.asarray(psi, dtype)
    m, S, n = psi.shape[0], int(num_samples), x.shape[0]
    ka, kg, km = prng.split(np.asarray(key, np.uint32), 3, layout_mode)
    if guide == "normal":
        assign = prng.randint_bits_to_index(prng.random_bits(ka, S, layout_mode), m)
    else:
        assign = (np.arange(S) % m).astype(np.int32)
    dkeys, nkeys = prng.split(kg, S, layout_mode), prng.split(km, S, layout_mode)
    lat = np.zeros((S, L), dtype)
    noise = np.zeros((S, n), dtype)
    for s in range(S):
        kk = dkeys[s]
        p = psi[assign[s]]
        for r in rows:
            kk, sub = prng.split(kk, 2, layout_mode)
            u = p[r["loc"]:r["loc"] + r["size"]]
            if guide == "normal":
                eps = prng.normal(sub, r["size"], layout_mode).astype(dtype)
                u = u + softplus(p[r["rho"]:r["rho"] + r["size"]]) * eps
            lat[s, r["lat"]:r["lat"] + r["size"]] = np.exp(u) if r["transform"] == models.T_EXP else u
        noise[s] = prng.normal(nkeys[s], n, layout_mode).astype(dtype)
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    theta = torch.from_numpy(lat).to(tdt)
    xb = torch.from_numpy(np.asarray(x, dtype)).to(tdt).reshape(n, -1)[None].expand(S, n, -1)
    y_loc = models.forward(spec, theta, xb).numpy()
    if spec.noise == "gamma":
        sigma = np.sqrt(1.0 / lat[:, rows[0]["lat"]])[:, None]
    else:
        sigma = np.full((S, 1), spec.sigma, dtype)
    out = dict(assign=assign, latents=lat, y_loc=y_loc, y=y_loc + sigma * noise, loglik=None)
    if y is not None:
        z = (np.asarray(y, dtype)[None] - y_loc) / sigma
        out["loglik"] = -0.5 * z * z - np.log(sigma) - 0.5 * np.log(2 * np.pi)
    return out


def rmse(y_loc, y):
    return float(np.sqrt(np.mean((np.asarray(y)[None] - y_loc) ** 2)))


def lpd(loglik):
    mx = loglik.max(0)
    return mx + np.log(np.exp(loglik - mx).sum(0)) - np.log(loglik.shape[0])


def nll_reg(loglik):
    return float(-lpd(loglik).mean())


def lppd(loglik):
    return float(lpd(loglik).sum())


# ---- classification metrics: literal numpy transcription of reference src/metrics.py ----------
def _softmax(l):
    e = np.exp(l - l.max(-1, keepdims=True))
    return e / e.sum(-1, keepdims=True)


def binned_conf_acc(logits, targets, bins=100):
    """src/metrics.py:270-321, statement by statement (jnp gather clamps out-of-range indices;
    jnp.histogram edges = linspace(0, 1, bins + 1) in float32, last bin right-closed)."""
    confs_all = _softmax(np.asarray(logits, np.float64))
    preds = confs_all.argmax(-1)
    confs = confs_all.max(-1)
    accs = preds == targets
    idx = np.argsort(confs, kind="stable")
    confs, accs = confs[idx], accs[idx]
    edges = (np.arange(bins + 1, dtype=np.float32) / np.float32(bins)).astype(np.float64)
    counts, _ = np.histogram(confs, bins=edges)
    counts = counts.astype(int)

    def compute_bins_mean(arr, counts):
        bins_end = counts.cumsum()
        bins_end = np.where(counts > 0, bins_end - 1, bins_end)
        cs = arr.cumsum()
        bin_cumsum = cs[np.clip(bins_end, 0, len(cs) - 1)]
        prev = np.concatenate([np.array([0], cs.dtype), bin_cumsum[:-1]])
        bin_sum = bin_cumsum - prev
        return np.where(counts > 0, bin_sum / np.maximum(counts, 1), bin_sum)

    return compute_bins_mean(confs, counts), compute_bins_mean(accs.astype(np.int64), counts)


def classification_metrics(logits, y, ypred=None, bins=100):
    """(P, 6) array {nll, acc, brier, ece, mce, conf} per posterior draw (nll_fn, acc_fn, brier_fn,
    ece_fn, mce_fn, conf_fn of src/metrics.py:17-160); acc / conf need the sampled y site."""
    logits = np.asarray(logits, np.float64)
    P, n, C = logits.shape
    out = np.full((P, 6), np.nan)
    for p in range(P):
        sm = _softmax(logits[p])
        lsm = np.log(sm)
        out[p, 0] = -lsm[np.arange(n), y].mean()
        out[p, 2] = ((sm - np.eye(C)[y]) ** 2).sum(1).mean()
        mc, ma = binned_conf_acc(logits[p], y, bins)
        out[p, 3] = np.abs(mc - ma).mean()
        out[p, 4] = np.abs(mc - ma).max()
        if ypred is not None:
            out[p, 1] = (y == ypred[p]).mean()
            out[p, 5] = np.exp(lsm[np.arange(n), ypred[p]]).mean()
    return out


def predictive_cls(spec, psi, key, num_samples, x, guide="normal", dtype=np.float64,
                   layout_mode=prng.LAYOUT_PARTITIONABLE):
    """Classifier counterpart of predictive_reg (reference src/models/mlp.py:14-65 under
    MixtureGuidePredictive): dict(assign, latents (S, L), logits (S, n, O), y (S, n)).
    y = argmax(logits + Gumbel), Gumbel = -log(-log(uniform(key_s, (n, O), minval=tiny)))
    (jax.random.categorical, as numpyro's Categorical.sample draws)."""
    rows, L, d = models.layout(spec, guide)
    psi = np.asarray(psi, dtype)
    m, S, n = psi.shape[0], int(num_samples), x.shape[0]
    ka, kg, km = prng.split(np.asarray(key, np.uint32), 3, layout_mode)
    assign = (prng.randint_bits_to_index(prng.random_bits(ka, S, layout_mode), m) if guide == "normal"
              else (np.arange(S) % m).astype(np.int32))
    dkeys, nkeys = prng.split(kg, S, layout_mode), prng.split(km, S, layout_mode)
    lat = np.zeros((S, L), dtype)
    for s in range(S):
        kk = dkeys[s]
        p = psi[assign[s]]
        for r in rows:
            kk, sub = prng.split(kk, 2, layout_mode)
            u = p[r["loc"]:r["loc"] + r["size"]]
            if guide == "normal":
                u = u + softplus(p[r["rho"]:r["rho"] + r["size"]]) * prng.normal(sub, r["size"], layout_mode).astype(dtype)
            lat[s, r["lat"]:r["lat"] + r["size"]] = u
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    xb = torch.from_numpy(np.asarray(x, dtype).reshape(n, -1)).to(tdt)[None].expand(S, n, -1)
    logits = models.forward(spec, torch.from_numpy(lat).to(tdt), xb).numpy()
    O = logits.shape[-1]
    tiny = np.float32(1.17549435e-38)
    y = np.zeros((S, n), np.int32)
    for s in range(S):
        f = prng.bits_to_unit_float(prng.random_bits(nkeys[s], n * O, layout_mode)).astype(np.float32)
        u = np.maximum(tiny, f + tiny).astype(np.float64)
        g = -np.log(-np.log(u)).reshape(n, O)
        y[s] = np.argmax(logits[s] + g, axis=-1)
    return dict(assign=assign, latents=lat, logits=logits, y=y)
