"""Generate the golden vectors under tests/golden/ from the float64 oracle.

    python tests/golden/make_golden.py

PARITY UNPINNED: the reference (aleatory-science/smi_experiments) ships no tests or fixtures for
the SteinVI step and its engine (numpyro @ master, jax unpinned) cannot be installed here, so
these vectors come from the repo's own CPU restatement (oracle/), evaluated in float64.  They
freeze the oracle (tests/test_golden.py, CPU) and are what the CUDA path is compared with on
the GPU box (tests/test_gpu_golden.py), where /root/reference does not exist.

One .npz per BASELINE config, at sizes that keep the files small:
  c1_sine_bnn      SMI, 1-D tanh BNN (src/models/reg_1d_bnn.py), sine-cluster data regenerated
                   as src/sine_wave_data.py:15-37 does (legacy seed 52), full batch, Adam
  c2_uci_bnn       SMI, UCI-shape ReLU BNN with Gamma precision, subsampled plate, Adam
  c3_gauss_svgd    SVGD on N(0, I), alpha = 10, Adam      (experiments/variance_collapse.py)
  c3_gauss_smi     SMI  on N(0, I), Adagrad
  c4_mlp_cls       SMI, tanh MLP classifier, pre-sliced minibatch, n / batch_size scaling, Adagrad
  c5_stein_force   Stein force alone at m = 130, d = 257 (tensor-core path shape, ragged tiles)
  prng             threefry2x32 split / bits / uniform / normal / subsample / key chain
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import models as om  # noqa: E402
from oracle import prng, stein  # noqa: E402
from oracle import step as ostep  # noqa: E402


def sine_clusters(nc1=20, nc2=20):
    """src/sine_wave_data.py:29-37 (sample_clusters) with its legacy global seeding."""
    np.random.seed(52)
    c1 = np.random.uniform(-1.5, -0.5, size=(nc1,))
    c2 = np.random.uniform(1.3, 1.7, size=(nc2,))
    x = np.sort(np.concatenate([c1, c2]))
    np.random.seed(52)
    y = 1.5 * np.sin(2 * np.pi * x + 2 / 3 * np.pi) + 3.0 * x + 1.0 + np.random.randn(x.shape[0]) * 0.1
    return x.reshape(-1, 1).astype(np.float32), y.astype(np.float32)


def init_particles(spec, guide, m, radius, rng, scale=0.1, jitter=0.2):
    rows, L, d = om.layout(spec, guide)
    psi = rng.uniform(-radius, radius, (m, d)).astype(np.float32)
    if guide == "normal":
        for r in rows:
            psi[:, r["rho"]:r["rho"] + r["size"]] = (np.log(np.expm1(scale)) +
                                                     jitter * rng.standard_normal((m, r["size"]))).astype(np.float32)
    return psi, L, d


def one_step(name, spec, guide, m, S, seed, x=None, y=None, N=None, B=None, tau=1.0, alpha=1.0, optimizer="adam",
             lr=1e-3, radius=0.1, key=(0, 1), plate_only=False, step_i=0):
    rng = np.random.default_rng(seed)
    psi, L, d = init_particles(spec, guide, m, radius, rng)
    a0 = (0.01 * rng.standard_normal((m, d))).astype(np.float32)
    b0 = (0.01 * rng.standard_normal((m, d)) ** 2).astype(np.float32)
    if optimizer == "adagrad":      # slot a = squared-gradient accumulator (>= 0), slot b = momentum
        a0, b0 = b0 + np.float32(1e-3), a0
    key = np.array(key, np.uint32)
    next_key, attractive = ostep.draw_keys(key)
    if plate_only:
        eps, _ = ostep.draw_noise(attractive, m, S, spec, guide, N, 1)
        idx = np.broadcast_to(np.arange(B, dtype=np.int32), (m, S, B)).copy()
        idx_out = np.zeros((0,), np.int32)
    else:
        eps, idx = ostep.draw_noise(attractive, m, S, spec, guide, N if B else None, B)
        idx_out = idx if idx is not None else np.zeros((0,), np.int32)
    out = ostep.smi_step(spec, psi.astype(np.float64), (a0.astype(np.float64), b0.astype(np.float64)), step_i, eps,
                         idx, x, y, N, guide, tau, alpha, 1.0, optimizer, lr, dtype=np.float64)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), psi=psi, opt_a=a0, opt_b=b0, step=np.int32(step_i), key=key,
        next_key=next_key, eps=eps if eps is not None else np.zeros((0,), np.float32), idx=idx_out,
        x=x if x is not None else np.zeros((0,), np.float32), y=y if y is not None else np.zeros((0,), np.float32),
        G=out["G"], elbo=out["elbo"], F=out["F"], loss=np.float64(out["loss"]), h=np.float64(out["h"]),
        psi_new=out["psi"], opt_a_new=out["opt_state"][0], opt_b_new=out["opt_state"][1],
        meta=np.array([m, S, N or 0, B or 0, L, d], np.int64), tau=np.float64(tau), alpha=np.float64(alpha),
        lr=np.float64(lr))
    print("%-16s m=%d S=%d L=%d d=%d loss=%.6f h=%.6f" % (name, m, S, L, d, out["loss"], out["h"]))


def main():
    x, y = sine_clusters()
    one_step("c1_sine_bnn", om.reg_1d_bnn(5), "normal", 5, 8, seed=0, x=x, y=y, N=40, radius=2.0, lr=1e-3, key=(0, 10))
    rng = np.random.default_rng(1)
    xu = rng.standard_normal((60, 13)).astype(np.float32)
    yu = (xu @ rng.standard_normal(13) * 0.3 + 0.1 * rng.standard_normal(60)).astype(np.float32)
    one_step("c2_uci_bnn", om.uci_bnn(13), "normal", 5, 6, seed=1, x=xu, y=yu, N=60, B=10, lr=5e-3, key=(0, 11),
             step_i=3)
    one_step("c3_gauss_svgd", om.gaussian(30), "delta", 20, 1, seed=2, tau=1.0 / 20, alpha=10.0, lr=0.05, radius=2.0,
             key=(0, 12))
    one_step("c3_gauss_smi", om.gaussian(30), "normal", 20, 10, seed=2, alpha=0.1, optimizer="adagrad", lr=0.05,
             radius=2.0, key=(0, 13), step_i=1)
    rng = np.random.default_rng(3)
    xc = rng.integers(0, 256, (16, 36)).astype(np.float32) / np.float32(255.0)
    yc = rng.integers(0, 5, 16).astype(np.int32)
    one_step("c4_mlp_cls", om.mnist_mlp(36, 12, 5), "normal", 4, 5, seed=3, x=xc, y=yc, N=60000, B=16,
             optimizer="adagrad", lr=0.8, radius=0.5, key=(0, 14), plate_only=True)

    rng = np.random.default_rng(4)
    m, d = 130, 257
    X = (rng.standard_normal((m, d)) + 0.5).astype(np.float32)
    G = rng.standard_normal((m, d)).astype(np.float32)
    F, loss, h, K = stein.stein_force(X, G, alpha=2.0, dtype=np.float64)
    D = stein.all_pairs_eucl_dist(X, np.float64)
    np.fill_diagonal(D, 0.0)
    a = np.sort(D.ravel())
    lo, hi = stein.median_ranks(m * m, stein.MEDIAN_JAX_FP32_INDEX)
    np.savez_compressed(os.path.join(HERE, "c5_stein_force.npz"), X=X, G=G, F=F, loss=np.float64(loss),
                        h=np.float64(h), alpha=np.float64(2.0), ranks=np.array([lo, hi], np.int64),
                        median_pair=np.array([a[lo], a[hi]]))
    print("c5_stein_force   m=%d d=%d loss=%.6f h=%.6f" % (m, d, loss, h))

    key = np.array([0x13198A2E, 0x03707344], np.uint32)
    spec = om.uci_bnn(2)
    nk, att = ostep.draw_keys(key)
    eps, idx = ostep.draw_noise(att, 2, 3, spec, "normal", 23, 5)
    np.savez_compressed(os.path.join(HERE, "prng.npz"), key=key, split7=prng.split(key, 7), bits=prng.random_bits(key, 257),
                        uniform=prng.uniform(key, 257, -0.1, 0.1), normal=prng.normal(key, 257),
                        subsample=ostep.subsample_indices(key, 57, 10), next_key=nk, attractive=att, eps=eps, idx=idx,
                        split7_legacy=prng.split(key, 7, prng.LAYOUT_LEGACY))
    print("prng             ok")


if __name__ == "__main__":
    main()
