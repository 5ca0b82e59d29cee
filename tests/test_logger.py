"""ExpLogger shim (row n4): call surface and on-disk layout of src/logger/{reg,class}_logger.py."""
import json

import numpy as np
import pytest
import torch

from smi_experiments_b200.logger import ClassExpLogger, ExpLogger


def test_regression_logger_layout(tmp_path):
    lg = ExpLogger("uci_std", logs_dir=tmp_path)
    assert lg.base_dir == tmp_path / "uci" / "std"
    lg.write_exp_config(repeat=2, methods=["smi"], lr=np.float32(5e-3))
    for rep in range(2):
        lg.write_entry("bnn", "smi", "rbf", "boston", rmse=2.5 + rep, nll=np.float64(2.6), loss=[3.0, 2.0, 1.0 + rep],
                       time=torch.tensor(0.25))
    lg.write_artifact("bnn", "smi", "rbf", "boston", guide="AutoNormal", model="uci_bnn",
                      params={"w1_auto_loc": torch.ones(2, 3)}, hyper_params={"lr": 5e-3}, post_draws={"y": np.zeros(4)},
                      summary={"rmse": 2.5})
    lg.save_logs()
    logs = json.load(open(lg.root / "logs.json"))
    e = logs["bnn"]["smi"]["rbf"]["boston"]
    assert e["rmse"] == [2.5, 3.5] and e["nll"] == [2.6, 2.6] and e["time"] == [0.25, 0.25]
    assert e["loss"] == ["boston/bnn/smi/rbf/loss/0.npy", "boston/bnn/smi/rbf/loss/1.npy"]
    assert e["artifact"] == ["boston/bnn/smi/rbf/artifact/0.dill"]
    assert np.array_equal(np.load(lg.root / e["loss"][1]), np.array([3.0, 2.0, 2.0]))
    art = lg.load_measure(e["artifact"][0])
    assert set(art) == {"model", "guide", "params", "hyper_params", "post_draws", "summary"}
    assert isinstance(art["params"]["w1_auto_loc"], np.ndarray) and art["params"]["w1_auto_loc"].shape == (2, 3)
    assert logs["exp_config"] == "exp_config.json"
    assert lg.load_exp_config() == {"repeat": 2, "methods": ["smi"], "lr": pytest.approx(5e-3)}
    # a fresh logger finds the run again
    lg2 = ExpLogger("uci_std", logs_dir=tmp_path)
    lg2.load_latest_logs()
    assert lg2.get_logs() == logs and lg2.root == lg.root


def test_loggers_do_not_share_state(tmp_path):
    a, b = ExpLogger("var", logs_dir=tmp_path), ExpLogger("var", logs_dir=tmp_path)
    a.write_entry("gauss", "svgd", "rbf", "d10", var=0.5)
    assert b.get_logs() == {}
    b.write_entry("gauss", "smi", "rbf", "d10", var=0.9)
    a.merge(b)
    assert set(a.get_logs()["gauss"]) == {"svgd", "smi"}


def test_classification_logger_layout(tmp_path):
    lg = ClassExpLogger("mnist", logs_dir=tmp_path)
    lg.write_entry("mlp", "smi", "mnist", acc=0.95, conf=[0.1, 0.9])
    lg.save_logs()
    e = json.load(open(lg.root / "logs.json"))["mlp"]["smi"]["mnist"]
    assert e["acc"] == [0.95] and e["conf"] == ["mnist/mlp/smi/conf/0.npy"]
    assert (lg.root / "mnist" / "mlp" / "smi" / "conf" / "0.npy").exists()


def test_unknown_experiment():
    with pytest.raises(AssertionError):
        ExpLogger("nope")
    with pytest.raises(AssertionError):
        ClassExpLogger("uci_std")
