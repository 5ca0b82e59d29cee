"""Stein kernels (numpyro.contrib.einstein.stein_kernels).

Only what the reference constructs: ``RBFKernel()`` (experiments/uci_benchmark.py:168,
experiments/img_classifier.py:161-174, experiments/variance_collapse.py:83,
experiments/synthetic/in_between_uncertainty.py:113, experiments/synthetic/smi_recovery.py:81,86):
mode "norm", bandwidth h = median(pairwise distance)^2 / ln(m) + eps32.
"""

from . import _lib


class RBFKernel:
    def __init__(self, mode="norm", bandwidth_factor=None, median_mode=_lib.MEDIAN_JAX_FP32_INDEX,
                 bandwidth=None):
        if mode != "norm":
            raise NotImplementedError("only RBFKernel(mode='norm') is on the reference's path")
        if bandwidth_factor is not None:
            raise NotImplementedError("the reference uses the default bandwidth_factor n -> 1/ln n")
        self.mode = mode
        self.median_mode = int(median_mode)
        self.h_fixed = 0.0 if bandwidth is None else float(bandwidth)
