"""numpyro.optim.Adam / Adagrad mirrors.

Reference construction sites: ``Adam(lr)`` experiments/uci_benchmark.py:191,
experiments/synthetic/in_between_uncertainty.py:122; ``Adagrad(0.8)``
experiments/img_classifier.py:161-173; ``Adagrad(LR)`` experiments/variance_collapse.py:100.
The update itself is one fused CUDA pass over the raveled particles (csrc/optim.cu).
"""

from collections import namedtuple

import torch

from . import ops

# step: device int32[1]; x: raveled unconstrained particles (m, d); a, b: optimizer slots.
OptimState = namedtuple("OptimState", ["step", "x", "a", "b"])


class _Optim:
    def init(self, x):
        return OptimState(torch.zeros(1, dtype=torch.int32, device=x.device), x,
                          torch.zeros_like(x), torch.zeros_like(x))

    def get_params(self, state):
        return state.x


class Adam(_Optim):
    """jax.example_libraries.optimizers.adam as wrapped by numpyro.optim.Adam."""

    name = "adam"

    def __init__(self, step_size, b1=0.9, b2=0.999, eps=1e-8):
        self.step_size, self.b1, self.b2, self.eps = float(step_size), float(b1), float(b2), float(eps)

    def update(self, F, state, grad_sign=-1.0):
        """In place: descends on grad_sign * F (the Stein step hands the optimizer -F)."""
        ops.adam(state.x, state.a, state.b, F, state.step, self.step_size, grad_sign, self.b1, self.b2,
                 self.eps)
        return state

    def update_gated(self, F, state, grad_sign=-1.0, skip=None):
        """As update, but the step counter is left to ops.step_commit and the whole pass is a
        no-op on the device once *skip is set (early-stopped run)."""
        ops.adam_gated(state.x, state.a, state.b, F, state.step, self.step_size, grad_sign, self.b1, self.b2,
                       self.eps, skip)
        return state


class Adagrad(_Optim):
    """jax.example_libraries.optimizers.adagrad (momentum 0.9) as wrapped by numpyro.optim.Adagrad."""

    name = "adagrad"

    def __init__(self, step_size, momentum=0.9):
        self.step_size, self.momentum = float(step_size), float(momentum)

    def update(self, F, state, grad_sign=-1.0):
        ops.adagrad(state.x, state.a, state.b, F, state.step, self.step_size, grad_sign, self.momentum)
        return state

    def update_gated(self, F, state, grad_sign=-1.0, skip=None):
        ops.adagrad_gated(state.x, state.a, state.b, F, self.step_size, grad_sign, self.momentum, skip)
        return state
