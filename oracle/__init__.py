"""CPU oracle for the SMI / SVGD update step.  TEST INFRASTRUCTURE ONLY.

This package restates, in NumPy / torch-CPU, the algorithm of the Stein step that
aleatory-science/smi_experiments drives through NumPyro's ``SteinVI`` (call sites:
reference ``src/engine_run.py:25,30``, ``src/model_select/early_stopping.py:42``,
``experiments/uci_benchmark.py:188-195``).  The arithmetic itself lives in the
third-party dependency ``numpyro @ git+https://github.com/pyro-ppl/numpyro.git@master``
(reference ``requirements.txt:12`` -- a floating branch, between ~0.15.2 and ~0.17.0
by the evidence in SURVEY.md section 8c) and in unpinned ``jax``; neither is installed
on this image and neither can be fetched.

PARITY UNPINNED: the reference holds no tests, golden vectors or fixtures for this
path, and the reference itself cannot be executed here.  The oracle is therefore
pinned only by independent checks (closed forms, finite differences, numpy.sort,
torch.optim, Random123 threefry known-answer vectors) -- see tests/test_oracle_*.py.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
``smi_experiments_b200`` never does.
"""

from . import stein, optim, prng, models, elbo, step  # noqa: F401
