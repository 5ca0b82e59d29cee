"""ExpLogger-compatible experiment logs.

Same call surface and the same files on disk as the reference's two loggers, so the reference's
``experiments/*.py`` and its plotting scripts read what a run against this engine wrote:

  regression      src/logger/reg_logger.py:34-199    entries keyed model_type / method / kernel / dataset
                  (ExpLogger("uci_std" | "uci_gap" | "lppd_1d" | "lppd_recover" | "var"),
                  used at experiments/uci_benchmark.py:443-461)
  classification  src/logger/class_logger.py:35-174  entries keyed model_type / method / dataset
                  (ExpLogger(name in src/constants.py:3-12))

Layout under ``<logs_dir>/<experiment dir>/<timestamp>/``:
  logs.json                          nested dict of lists, one appended value per write_entry call
  exp_config.json                    write_exp_config(**kw)
  <dataset>/<model_type>/<method>[/<kernel>]/<measure>/<n>.npy     list-valued measures
  <dataset>/<model_type>/<method>[/<kernel>]/artifact/<n>.dill    write_artifact(...)
list-valued measures and artifacts are stored beside logs.json and entered by their path relative
to the run directory.  Device tensors are brought to the host (numpy) before they are written.

Differences, on purpose: the log directory is an argument (default ``./logs``) instead of the
package's parent directory, and every logger owns its log dict and timestamp (the reference keeps
both as class attributes shared by all instances).
"""

import json
import pickle
from datetime import datetime
from pathlib import Path

import numpy as np

try:  # the reference pickles artifacts with dill (closures in model / guide objects)
    import dill as _pickler
except ImportError:  # pragma: no cover
    _pickler = pickle

REG_EXPERIMENTS = {          # src/logger/reg_logger.py:55-83
    "lppd_1d": ("lppd", "1d"),
    "lppd_recover": ("lppd", "recover"),
    "uci_std": ("uci", "std"),
    "uci_gap": ("uci", "gap"),
    "var": ("var",),
}
CLASS_EXPERIMENTS = ("mnist", "collapse", "uci", "1d-reg", "recover")   # src/constants.py:3-12


def _to_host(v):
    """Tensors -> numpy, containers walked; everything else unchanged."""
    try:
        import torch
        if isinstance(v, torch.Tensor):
            return v.detach().cpu().numpy()
    except ImportError:  # pragma: no cover
        pass
    if isinstance(v, dict):
        return {k: _to_host(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return type(v)(_to_host(x) for x in v)
    return v


def _jsonable(v):
    v = _to_host(v)
    if isinstance(v, np.ndarray):
        return v.tolist()
    if isinstance(v, np.generic):
        return v.item()
    if isinstance(v, dict):
        return {k: _jsonable(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [_jsonable(x) for x in v]
    return v


class _BaseLogger:
    nkeys = 0   # number of name levels above the measures

    def __init__(self, base_dir):
        self.base_dir = Path(base_dir)
        self.base_dir.mkdir(parents=True, exist_ok=True)
        self.timestamp = str(datetime.now())
        self.root = self.base_dir / self.timestamp
        self.logs = {}

    # ---- entries ------------------------------------------------------------------------
    def _entry(self, names):
        assert len(names) == self.nkeys
        node = self.logs
        for n in names:
            node = node.setdefault(n, {})
        return node

    def _dir(self, names):
        # <dataset>/<the other names in order>
        return self.root.joinpath(names[-1], *names[:-1])

    def _write_entry(self, names, measures):
        node = self._entry(names)
        for k, v in measures.items():
            if isinstance(v, list):
                d = self._dir(names) / k
                d.mkdir(parents=True, exist_ok=True)
                fp = d / ("%d.npy" % len(node.get(k, [])))
                np.save(fp, np.array(_to_host(v)))
                v = str(fp.relative_to(self.root))
            node.setdefault(k, []).append(_jsonable(v))

    def _write_artifact(self, names, guide, model, params, hyper_params, post_draws, summary):
        node = self._entry(names)
        d = self._dir(names) / "artifact"
        d.mkdir(parents=True, exist_ok=True)
        fp = d / ("%d.dill" % len(node.get("artifact", [])))
        with open(fp, "wb") as f:
            _pickler.dump({"model": model, "guide": guide, "params": _to_host(params), "hyper_params": hyper_params,
                           "post_draws": _to_host(post_draws), "summary": _to_host(summary)}, f)
        node.setdefault("artifact", []).append(str(fp.relative_to(self.root)))

    # ---- config -------------------------------------------------------------------------
    def write_exp_config(self, **kwarg):
        self.root.mkdir(parents=True, exist_ok=True)
        with (self.root / "exp_config.json").open("w") as f:
            json.dump(_jsonable(kwarg), f)
        self.logs["exp_config"] = "exp_config.json"

    def load_exp_config(self):
        with (self.root / self.logs["exp_config"]).open() as f:
            return json.load(f)

    def merge(self, logger):
        for mt, methods in logger.get_logs().items():
            if mt == "exp_config":
                continue
            for method, res in methods.items():
                self.logs.setdefault(mt, {})[method] = res

    # ---- persistence --------------------------------------------------------------------
    def get_logs(self):
        return json.loads(json.dumps(self.logs))   # plain nested dicts / lists, detached

    def save_logs(self):
        self.root.mkdir(parents=True, exist_ok=True)
        with (self.root / "logs.json").open("w") as f:
            json.dump(self.logs, f)

    def load_logs(self, timestamp):
        self.timestamp = timestamp
        self.root = self.base_dir / timestamp
        with (self.root / "logs.json").open() as f:
            self.logs = json.load(f)

    def load_latest_logs(self):
        runs = sorted(p.name for p in self.base_dir.glob("*") if p.is_dir())
        if not runs:
            raise FileNotFoundError("no logged runs under %s" % self.base_dir)
        self.load_logs(runs[-1])

    def load_measure(self, rel):
        """Array (or artifact dict) behind a path stored in the logs."""
        fp = self.root / rel
        if fp.suffix == ".npy":
            return np.load(fp, allow_pickle=False)
        with open(fp, "rb") as f:
            return _pickler.load(f)


class ExpLogger(_BaseLogger):
    """Regression logger: write_entry(model_type, method, kernel, dataset, **measures)."""
    nkeys = 4

    def __init__(self, experiment_name, logs_dir="logs"):
        if experiment_name not in REG_EXPERIMENTS:
            raise AssertionError("Unknown experiment.")
        super().__init__(Path(logs_dir).joinpath(*REG_EXPERIMENTS[experiment_name]))

    def write_entry(self, model_type, method, kernel, dataset, **measures):
        self._write_entry((model_type, method, kernel, dataset), measures)

    def write_artifact(self, model_type, method, kernel, dataset, guide, model, params, hyper_params, post_draws,
                       summary):
        self._write_artifact((model_type, method, kernel, dataset), guide, model, params, hyper_params, post_draws,
                             summary)


class ClassExpLogger(_BaseLogger):
    """Classification logger: write_entry(model_type, method, dataset, **measures)."""
    nkeys = 3

    def __init__(self, experiment_name, logs_dir="logs"):
        if experiment_name not in CLASS_EXPERIMENTS:
            raise AssertionError("Unknown experiment.")
        super().__init__(Path(logs_dir) / experiment_name)

    def write_entry(self, model_type, method, dataset, **measures):
        self._write_entry((model_type, method, dataset), measures)

    def write_artifact(self, model_type, method, dataset, guide, model, params, hyper_params, post_draws, summary):
        self._write_artifact((model_type, method, dataset), guide, model, params, hyper_params, post_draws, summary)
