"""The committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py from the
float64 oracle) still match what the oracle computes, in float64 (tight) and float32 (the
tolerance-calibration run).  Guards the checker itself against drift."""
import os

import numpy as np
import pytest

from oracle import models as om
from oracle import prng, stein
from oracle import step as ostep

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

CASES = {
    "c1_sine_bnn": (om.reg_1d_bnn(5), "normal", "adam"),
    "c2_uci_bnn": (om.uci_bnn(13), "normal", "adam"),
    "c3_gauss_svgd": (om.gaussian(30), "delta", "adam"),
    "c3_gauss_smi": (om.gaussian(30), "normal", "adagrad"),
    "c4_mlp_cls": (om.mnist_mlp(36, 12, 5), "normal", "adagrad"),
}


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def load_case(name):
    z = np.load(os.path.join(G, name + ".npz"))
    m, S, N, B, L, d = (int(v) for v in z["meta"])
    eps = z["eps"] if z["eps"].size else None
    x = z["x"] if z["x"].size else None
    y = z["y"] if z["y"].size else None
    if name == "c4_mlp_cls":
        idx = np.broadcast_to(np.arange(B, dtype=np.int32), (m, S, B))
    else:
        idx = z["idx"] if z["idx"].size else None
    return z, dict(m=m, S=S, N=N or None, B=B or None, L=L, d=d, eps=eps, x=x, y=y, idx=idx)


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 2e-4)])
def test_oracle_reproduces_golden_step(name, dtype, tol):
    spec, guide, optimizer = CASES[name]
    z, c = load_case(name)
    out = ostep.smi_step(spec, z["psi"].astype(dtype), (z["opt_a"].astype(dtype), z["opt_b"].astype(dtype)),
                         int(z["step"]), c["eps"], c["idx"], c["x"], c["y"], c["N"], guide, float(z["tau"]),
                         float(z["alpha"]), 1.0, optimizer, float(z["lr"]), dtype=dtype)
    assert _rel(out["G"], z["G"]) <= tol
    assert _rel(out["F"], z["F"]) <= tol
    assert abs(float(out["loss"]) - float(z["loss"])) <= tol * float(z["loss"])
    assert abs(float(out["h"]) - float(z["h"])) <= max(tol, 1e-4 if dtype == np.float32 else 0) * float(z["h"])
    assert _rel(out["psi"], z["psi_new"]) <= tol


@pytest.mark.parametrize("name", sorted(CASES))
def test_golden_noise_comes_from_the_key_chain(name):
    spec, guide, _ = CASES[name]
    z, c = load_case(name)
    nk, att = ostep.draw_keys(z["key"])
    assert (nk == z["next_key"]).all()
    if name == "c4_mlp_cls":
        eps, _ = ostep.draw_noise(att, c["m"], c["S"], spec, guide, c["N"], 1)
    else:
        eps, idx = ostep.draw_noise(att, c["m"], c["S"], spec, guide, c["N"] if c["B"] else None, c["B"])
        if c["idx"] is not None:
            assert (idx == z["idx"]).all()
    if eps is not None:
        assert np.array_equal(eps, z["eps"])


def test_golden_stein_force():
    z = np.load(os.path.join(G, "c5_stein_force.npz"))
    F, loss, h, _ = stein.stein_force(z["X"], z["G"], alpha=float(z["alpha"]), dtype=np.float64)
    assert _rel(F, z["F"]) <= 1e-12 and abs(loss - z["loss"]) <= 1e-12 * z["loss"] and h == z["h"]
    F32, loss32, h32, _ = stein.stein_force(z["X"], z["G"], alpha=float(z["alpha"]), dtype=np.float32)
    assert _rel(F32, z["F"]) <= 1e-5 and abs(float(h32) - float(z["h"])) <= 1e-5 * float(z["h"])
    assert tuple(z["ranks"]) == stein.median_ranks(130 * 130)


def test_golden_prng():
    z = np.load(os.path.join(G, "prng.npz"))
    key = z["key"]
    assert (prng.split(key, 7) == z["split7"]).all()
    assert (prng.split(key, 7, prng.LAYOUT_LEGACY) == z["split7_legacy"]).all()
    assert (prng.random_bits(key, 257) == z["bits"]).all()
    assert np.array_equal(prng.uniform(key, 257, -0.1, 0.1), z["uniform"])
    assert np.array_equal(prng.normal(key, 257), z["normal"])
    assert (ostep.subsample_indices(key, 57, 10) == z["subsample"]).all()
