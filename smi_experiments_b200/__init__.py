"""smi_experiments_b200: B200-native Stein Mixture Inference / SVGD update step.

Drop-in for the NumPyro SteinVI path that aleatory-science/smi_experiments drives
(SURVEY.md section 8): `SteinVI`, `SVGD`, `ASVGD`, `RBFKernel`, `Adam`, `Adagrad`, AutoNormal /
AutoDelta guides and descriptors of the reference's models, over hand-written sm_100a CUDA
kernels behind the C ABI of include/smi.h (libsmi.so).  No CPU fallback.
"""

from . import _lib, models, ops  # noqa: F401
from .kernels import RBFKernel  # noqa: F401
from .models import (AutoDelta, AutoNormal, BNNClassifier, BNNRegression, GaussianTarget,  # noqa: F401
                     init_to_uniform, mnist_mlp, reg_1d_bnn, uci_bnn)
from .predictive import classification_metrics  # noqa: F401
from .predictive import MixtureGuidePredictive, Predictive, lppd_fn, nll_reg_fn, rmse_fn  # noqa: F401
from .optim import Adagrad, Adam  # noqa: F401
from .early_stop import OnlineEarlyStop  # noqa: F401
from .steinvi import ASVGD, SVGD, PRNGKey, SteinVI, SteinVIRunResult, SteinVIState  # noqa: F401

__version__ = "0.1.0"
from .logger import ClassExpLogger, ExpLogger  # noqa: F401
from .ops import set_prng_layout, set_randint_rule  # noqa: F401
