"""Oracle: one full SteinVI.update -- PRNG chain, ELBO grads, Stein force, optimizer.

Restates numpyro/contrib/einstein/steinvi.py (SteinVI.update, _svgd_loss_and_grads;
SURVEY.md Appendix A.4-A.6) plus the seed-handler key chain of numpyro/handlers.py and
numpyro/primitives.py::_subsample_fn (CPU branch) (Appendix A.5).  Reference callers:
src/engine_run.py:25,30,41-48; src/model_select/early_stopping.py:42.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import numpy as np

from . import optim as ooptim
from . import prng
from . import stein
from .elbo import elbo_grads
from .models import layout


def draw_keys(state_key, layout_mode=prng.LAYOUT_PARTITIONABLE):
    """state key -> (next_key, step_key) ; step_key -> attractive_key  (Appendix A.5)."""
    k3 = prng.split(state_key, 3, layout_mode)
    next_key, step_key = k3[0], k3[2]
    attractive_key = prng.split(step_key, 2, layout_mode)[0]
    return next_key, attractive_key


def subsample_indices(key, n_total, b, layout_mode=prng.LAYOUT_PARTITIONABLE, randint="mulshift"):
    """Partial Fisher-Yates over arange(n_total): slot t swaps positions (n-1-t, j_t),
    j_t uniform in [0, n-t); returns the last b entries.  j_t is drawn from ONE 32-bit word of
    split(key, b)[t] by multiply-shift (device rule) -- jax.random.randint's two-word
    modular construction is NOT reproduced [NP?]."""
    if randint == "choice":
        # the branch _subsample_fn takes off the CPU backend: random.choice(key, n, (b,), replace=False)
        # = permutation(key, n)[:b]; jax/_src/random.py::_shuffle: rounds of a stable sort by fresh words
        rounds = int(np.ceil(3 * np.log(max(1, n_total)) / np.log(np.iinfo(np.uint32).max)))
        x = np.arange(n_total, dtype=np.int32)
        for _ in range(rounds):
            pair = prng.split(key, 2, layout_mode)
            key, sub = pair[0], pair[1]
            x = x[np.argsort(prng.random_bits(sub, n_total, layout_mode), kind="stable")]
        return x[:b].copy()
    keys = prng.split(key, b, layout_mode)
    val = np.arange(n_total, dtype=np.int32)
    for t in range(b):
        span = n_total - t
        if randint == "jax":   # the switch for jax.random.randint's own construction
            j = prng.randint_jax(keys[t], span, layout_mode)
        else:
            bits = prng.random_bits(keys[t], 1, layout_mode)
            j = int(prng.randint_bits_to_index(bits, span)[0])
        i = span - 1
        val[i], val[j] = val[j], val[i]
    return val[n_total - b:].copy()


def draw_noise(attractive_key, m, S, spec, guide="normal", n_total=None, b=None,
               layout_mode=prng.LAYOUT_PARTITIONABLE, randint="mulshift"):
    """eps (m, S, L) float32 and idx (m, S, b) int32 for one step.
    Chain: split(attractive, m)[i] -> split(., S)[s] -> split(., 2)[0] = guide key; the seed
    handler then peels one subkey per RNG-consuming statement: the data plate (if subsampled)
    first, then every latent site in model order."""
    rows, L, _ = layout(spec, guide)
    eps = np.zeros((m, S, L), np.float32) if guide == "normal" else None
    idx = np.zeros((m, S, b), np.int32) if b is not None else None
    pkeys = prng.split(attractive_key, m, layout_mode)
    for i in range(m):
        dkeys = prng.split(pkeys[i], S, layout_mode)
        for s in range(S):
            key = prng.split(dkeys[s], 2, layout_mode)[0]
            if b is not None:
                pair = prng.split(key, 2, layout_mode)
                key, sub = pair[0], pair[1]
                idx[i, s] = subsample_indices(sub, n_total, b, layout_mode, randint)
            if guide == "normal":
                for r in rows:
                    pair = prng.split(key, 2, layout_mode)
                    key, sub = pair[0], pair[1]
                    eps[i, s, r["lat"]:r["lat"] + r["size"]] = prng.normal(sub, r["size"], layout_mode)
    return eps, idx


def smi_step(spec, psi, opt_state, step_i, eps, idx, x=None, y=None, n_total=None,
             guide="normal", tau=1.0, alpha=1.0, c_att=1.0, optimizer="adam", lr=1e-3,
             median_mode=stein.MEDIAN_JAX_FP32_INDEX, dtype=np.float32):
    """One update given explicit noise.  Returns dict(psi, opt_state, loss, h, G, F, elbo)."""
    import torch
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    psi = np.asarray(psi, dtype)
    G, elbo = elbo_grads(spec, psi, eps, idx, x, y, n_total, tau, guide, tdt)
    F, loss, h, _ = stein.stein_force(psi, G.astype(dtype), alpha, c_att, median_mode, dtype)
    grad = -F
    a, b = opt_state
    if optimizer == "adam":
        new_psi, a, b = ooptim.adam_update(step_i, grad, psi, a, b, lr)
    else:
        new_psi, a, b = ooptim.adagrad_update(step_i, grad, psi, a, b, lr)
    return dict(psi=new_psi, opt_state=(a, b), loss=loss, h=h, G=G, F=F, elbo=elbo)
