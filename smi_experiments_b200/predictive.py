"""MixtureGuidePredictive / Predictive and the regression metrics on the device.

Host mirror of what the reference runs right after the path:
``smi_sampler`` / ``svgd_sampler`` (src/metrics.py:216-237) and ``rmse_fn`` / ``nll_reg_fn`` /
``lppd_fn`` (src/metrics.py:222-268), as called from experiments/uci_benchmark.py:297-316 with
5000 draws.  Every operator is one call into libsmi (csrc/predictive.cu); there is no CPU path.
Supported models: the regression BNN descriptors (uci_bnn, reg_1d_bnn).
"""
import ctypes as C

import torch

from . import _lib, models, ops
from .steinvi import _key_tensor


class MixtureGuidePredictive:
    """``MixtureGuidePredictive(model, guide, params, num_samples, return_sites, guide_sites)``;
    calling it with ``(rng_key, x, y=None, subsample=None)`` returns ``{site: tensor}`` with a
    leading ``num_samples`` axis.  ``params`` is ``engine.get_params(state)`` (or the state)."""

    def __init__(self, model, guide, params, num_samples=1, return_sites=None, guide_sites=None, device=None):
        if not isinstance(model, (models.BNNRegression, models.BNNClassifier)):
            raise NotImplementedError("predictive draws are implemented for the BNN descriptors")
        self.model, self.guide, self.num_samples = model, guide, int(num_samples)
        self.return_sites = list(return_sites) if return_sites is not None else None
        self.guide_sites = guide_sites
        self.md, self.layout = models.make_desc(model, guide.kind)
        self.device = torch.device(device if device is not None else "cuda")
        self.psi = self._ravel(params)

    def _ravel(self, params):
        """Constrained named params (m, *shape) -> raveled unconstrained particles (m, d)."""
        if hasattr(params, "optim_state"):
            return params.optim_state.x.to(self.device).contiguous()
        first = next(iter(params.values()))
        m = first.shape[0]
        psi = torch.empty(m, self.layout.particle_dim, dtype=torch.float32, device=self.device)
        for name in self.layout.names:
            off, shape, site, which = self.layout.leaves[name]
            v = torch.as_tensor(params[name], dtype=torch.float32).to(self.device).reshape(m, -1)
            if which == "rho":
                v = torch.log(torch.expm1(v))                     # softplus^-1 of the guide scale
            elif self.guide.kind == _lib.GUIDE_DELTA and site.transform == _lib.T_EXP:
                v = torch.log(v)
            psi[:, off:off + site.size] = v
        return psi

    def _call_classifier(self, rng_key, x):
        """logits (S, n, O) and the Categorical draw y (S, n): latents by threefry, first layer as one
        stacked tcgen05 GEMM over all draws, layers 2-3 + Gumbel-argmax per draw."""
        dev = self.device
        x = torch.as_tensor(x)
        x = x.to(dev).reshape(x.shape[0], -1).to(torch.float32).contiguous()
        n, L, H, O = x.shape[0], self.layout.latent_dim, self.model.hidden, self.model.out_dim
        S = self.psi.shape[0] if self.guide.kind == _lib.GUIDE_DELTA else self.num_samples
        key = _key_tensor(rng_key, dev)
        lat = torch.empty(S, L, device=dev)
        assign = torch.empty(S, dtype=torch.int32, device=dev)
        _lib.call("smi_predictive_latents_f32", C.byref(self.md), ops._ptr(self.psi), self.psi.shape[0],
                  ops._ptr(key), S, ops._ptr(assign), ops._ptr(lat), ops._stream())
        nbytes = _lib.load().smi_predictive_cls_layer1_workspace_bytes(C.byref(self.md), S, n)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        h1 = torch.empty(S, n, H, device=dev)
        _lib.call("smi_predictive_cls_layer1_f32", C.byref(self.md), ops._ptr(lat), S, ops._ptr(x), n, ops._ptr(h1),
                  ops._ptr(ws), ws.numel(), ops._stream())
        logits = torch.empty(S, n, O, device=dev)
        yd = torch.empty(S, n, dtype=torch.int32, device=dev)
        _lib.call("smi_predictive_cls_head_f32", C.byref(self.md), ops._ptr(lat), ops._ptr(h1), ops._ptr(key), S, n,
                  ops._ptr(logits), ops._ptr(yd), ops._stream())
        res = {"logits": logits, "y": yd}
        off = 0
        for site in self.model.sites():
            res[site.name] = lat[:, off:off + site.size].reshape((S,) + tuple(site.shape))
            off += site.size
        if self.return_sites is not None:
            res = {k: v for k, v in res.items() if k in self.return_sites}
        res["_assign"] = assign
        return res

    def __call__(self, rng_key, x, y=None, subsample=None, **_ignored):
        if isinstance(self.model, models.BNNClassifier):
            return self._call_classifier(rng_key, x)
        dev = self.device
        x = torch.as_tensor(x, dtype=torch.float32).to(dev).reshape(len(x), -1).contiguous()
        n, S, L = x.shape[0], self.num_samples, self.layout.latent_dim
        if self.guide.kind == _lib.GUIDE_DELTA:
            S = self.psi.shape[0]
        yd = None if y is None else torch.as_tensor(y, dtype=torch.float32).to(dev).contiguous()
        key = _key_tensor(rng_key, dev)
        want = self.return_sites
        out = dict(y_loc=torch.empty(S, n, device=dev), y=torch.empty(S, n, device=dev),
                   assign=torch.empty(S, dtype=torch.int32, device=dev),
                   latents=torch.empty(S, L, device=dev),
                   loglik=torch.empty(S, n, device=dev) if yd is not None else None)
        _lib.call("smi_predictive_reg_f32", C.byref(self.md), ops._ptr(self.psi), self.psi.shape[0], ops._ptr(key),
                  S, ops._ptr(x), n, ops._ptr(yd), ops._ptr(out["assign"]), ops._ptr(out["latents"]),
                  ops._ptr(out["y_loc"]), ops._ptr(out["y"]), ops._ptr(out["loglik"]), ops._stream())
        res = {"y_loc": out["y_loc"], "y": out["y"]}
        off = 0
        for site in self.model.sites():
            res[site.name] = out["latents"][:, off:off + site.size].reshape((S,) + tuple(site.shape))
            off += site.size
        if want is not None:
            res = {k: v for k, v in res.items() if k in want}
        res["_assign"], res["_loglik"] = out["assign"], out["loglik"]
        return res


Predictive = MixtureGuidePredictive   # AutoDelta guides: one draw per particle (svgd_sampler)


def _ws(S, n, dev):
    nbytes = _lib.load().smi_metric_workspace_bytes(S, n)
    return torch.empty(nbytes, dtype=torch.uint8, device=dev)


def rmse_fn(posterior_samples, y, batch_ndims=1):
    """src/metrics.py:249-253 -- device scalar."""
    y_loc = posterior_samples["y_loc"].contiguous()
    S, n = y_loc.shape
    yd = torch.as_tensor(y, dtype=torch.float32).to(y_loc.device).contiguous()
    out = torch.empty(1, device=y_loc.device)
    w = _ws(S, n, y_loc.device)
    _lib.call("smi_metric_rmse_f32", ops._ptr(y_loc), ops._ptr(yd), S, n, ops._ptr(out), ops._ptr(w), w.numel(),
              ops._stream())
    return out.reshape(())


def _nll(loglik):
    S, n = loglik.shape
    out = torch.empty(2, device=loglik.device)
    w = _ws(S, n, loglik.device)
    _lib.call("smi_metric_nll_f32", ops._ptr(loglik.contiguous()), S, n, ops._ptr(out), None, ops._ptr(w), w.numel(),
              ops._stream())
    return out


def nll_reg_fn(posterior_samples, model=None, x=None, y=None, batch_ndims=1):
    """src/metrics.py:256-263 from the per-draw log-likelihoods the predictive call produced."""
    ll = posterior_samples["_loglik"]
    if ll is None:
        raise ValueError("call the predictive with y to get log-likelihoods")
    return _nll(ll)[0]


def lppd_fn(posterior_samples, model=None, xs=None, ys=None, batch_ndims=1):
    """src/metrics.py:222-246."""
    ll = posterior_samples["_loglik"]
    if ll is None:
        raise ValueError("call the predictive with y to get log-likelihoods")
    return _nll(ll)[1]


def classification_metrics(posterior_samples, y, bins=100):
    """{"nll", "acc", "brier", "ece", "mce", "conf"}: per-draw device vectors, the six metrics of
    INDIST_EVAL_METRICS (src/constants.py:26-33; src/metrics.py:17-160) from ``posterior_samples["logits"]``
    (S, n, C) and, for acc / conf, the sampled ``posterior_samples["y"]`` (S, n)."""
    logits = torch.as_tensor(posterior_samples["logits"], dtype=torch.float32)
    if not logits.is_cuda:
        logits = logits.cuda()
    logits = logits.contiguous()
    if logits.dim() == 4:           # flatten_post_dim (src/metrics.py:324-345)
        logits = logits.reshape(-1, *logits.shape[2:])
    S, n, C = logits.shape
    dev = logits.device
    yd = torch.as_tensor(y).to(dev).to(torch.int32).contiguous()
    yp = posterior_samples.get("y")
    if yp is not None:
        yp = torch.as_tensor(yp).to(dev).to(torch.int32).reshape(S, n).contiguous()
    out = torch.empty(S, 6, device=dev)
    _lib.call("smi_metric_classification_f32", ops._ptr(logits), ops._ptr(yd), ops._ptr(yp), S, n, C, int(bins),
              ops._ptr(out), ops._stream())
    return {k: out[:, i] for i, k in enumerate(("nll", "acc", "brier", "ece", "mce", "conf"))}
