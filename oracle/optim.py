"""Oracle: numpyro.optim.Adam / Adagrad update rules.

Restates jax.example_libraries.optimizers.adam / adagrad as wrapped by numpyro/optim.py
(SURVEY.md Appendix A.4).  Reference call sites: ``Adam(lr)`` experiments/uci_benchmark.py:191,
experiments/synthetic/in_between_uncertainty.py:122; ``Adagrad(0.8)``
experiments/img_classifier.py:161-173; ``Adagrad(LR)`` experiments/variance_collapse.py:100.

The optimizer DESCENDS on the gradient it is handed; the Stein step hands it ``-F``.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import numpy as np


def adam_init(x):
    return np.zeros_like(x), np.zeros_like(x)


def adam_update(step, g, x, m, v, lr, b1=0.9, b2=0.999, eps=1e-8):
    """step is the 0-based count of updates applied so far."""
    dt = x.dtype.type
    m = dt(1 - b1) * g + dt(b1) * m
    v = dt(1 - b2) * (g * g) + dt(b2) * v
    mhat = m / dt(1 - np.asarray(b1, dt) ** dt(step + 1))
    vhat = v / dt(1 - np.asarray(b2, dt) ** dt(step + 1))
    x = x - dt(lr) * mhat / (np.sqrt(vhat) + dt(eps))
    return x.astype(dt), m.astype(dt), v.astype(dt)


def adagrad_init(x):
    return np.zeros_like(x), np.zeros_like(x)


def adagrad_update(step, g, x, acc, mom, lr, momentum=0.9):
    dt = x.dtype.type
    acc = acc + g * g
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = np.where(acc > 0, dt(1) / np.sqrt(acc), dt(0)).astype(dt)
    mom = dt(1.0 - momentum) * (g * inv) + dt(momentum) * mom
    x = x - dt(lr) * mom
    return x.astype(dt), acc.astype(dt), mom.astype(dt)
