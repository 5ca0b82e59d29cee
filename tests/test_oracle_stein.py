"""Pins for the Stein-force oracle (oracle/stein.py).  PARITY UNPINNED: the reference holds no
tests for this path; these are the independent checks of SURVEY.md 8(c)."""
import math

import numpy as np
import pytest

from oracle import stein


def test_force_matches_double_loop_definition():
    rng = np.random.default_rng(0)
    for m, d, alpha, c_att in [(3, 4, 1.0, 1.0), (7, 11, 10.0, 1.0), (5, 2, 0.1, 0.5)]:
        x = rng.standard_normal((m, d))
        g = rng.standard_normal((m, d))
        f, loss, h, k = stein.stein_force(x, g, alpha, c_att, dtype=np.float64)
        fd = stein.stein_force_direct(x, g, float(h), alpha, c_att)
        np.testing.assert_allclose(f, fd, rtol=1e-12, atol=1e-14)
        assert abs(loss - np.linalg.norm(fd)) <= 1e-12 * loss
        np.testing.assert_allclose(k, k.T, rtol=0, atol=1e-15)
        np.testing.assert_allclose(np.diag(k), 1.0)


def test_repulsive_term_is_gradient_of_kernel_finite_difference():
    rng = np.random.default_rng(1)
    m, d, h = 4, 3, 1.7
    x = rng.standard_normal((m, d))
    g = np.zeros((m, d))
    f = stein.stein_force_direct(x, g, h, alpha=1.0)
    e = 1e-6
    for i in range(m):
        num = np.zeros(d)
        for j in range(m):
            for k in range(d):
                xp, xm = x[j].copy(), x[j].copy()
                xp[k] += e
                xm[k] -= e
                kp = math.exp(-np.sum((xp - x[i]) ** 2) / h) if j != i else 1.0
                km = math.exp(-np.sum((xm - x[i]) ** 2) / h) if j != i else 1.0
                num[k] += (kp - km) / (2 * e) / m
        np.testing.assert_allclose(f[i], num, rtol=1e-6, atol=1e-9)


def test_gaussian_target_svgd_closed_form():
    """SVGD on N(0, I): score = -x * tau; with h -> infinity K = 1 and F_i = tau * mean-free sum."""
    rng = np.random.default_rng(2)
    m, d = 6, 5
    x = rng.standard_normal((m, d))
    g = -x / m
    f, _, _, _ = stein.stein_force(x, g, alpha=1.0, dtype=np.float64, h=1e12)
    np.testing.assert_allclose(f, np.broadcast_to(g.sum(0), (m, d)), rtol=1e-9, atol=1e-12)


def test_single_particle_edge():
    x = np.array([[0.3, -1.0, 2.0]], np.float32)
    g = np.array([[1.0, 2.0, 3.0]], np.float32)
    f, loss, h, k = stein.stein_force(x, g)
    assert h == np.float32(1.0) and k.shape == (1, 1) and k[0, 0] == 1.0
    np.testing.assert_array_equal(f, g)
    assert stein.median_bandwidth(x) == np.float32(1.0)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 25, 400, 4096 ** 2, 4097 ** 2])
def test_median_ranks_agree_below_2_pow_24(n):
    if n <= 2 ** 24:
        assert stein.median_ranks(n, stein.MEDIAN_JAX_FP32_INDEX) == stein.median_ranks(n, stein.MEDIAN_EXACT_MIDPOINT)
    lo, hi = stein.median_ranks(n, stein.MEDIAN_EXACT_MIDPOINT)
    assert lo == (n - 1) // 2 and hi == n // 2


def test_median_fp32_index_edges():
    # float32(n - 1) rounds to a power of two: a single element is selected (Appendix A.3)
    assert stein.median_ranks(8192 ** 2, stein.MEDIAN_JAX_FP32_INDEX) == (33554432, 33554432)
    assert stein.median_ranks(16384 ** 2, stein.MEDIAN_JAX_FP32_INDEX) == (2 ** 27, 2 ** 27)
    assert stein.median_ranks(8192 ** 2, stein.MEDIAN_EXACT_MIDPOINT) == (33554431, 33554432)


def test_median_matches_numpy():
    rng = np.random.default_rng(3)
    for n in (1, 2, 9, 16, 101, 400):
        a = rng.standard_normal(n).astype(np.float32)
        assert stein.median(a) == np.float32(np.median(a))
        assert stein.median(a, stein.MEDIAN_EXACT_MIDPOINT) == np.float32(np.median(a))


def test_bandwidth_formula_and_distance_form():
    rng = np.random.default_rng(4)
    m, d = 9, 6
    x = rng.standard_normal((m, d)).astype(np.float32)
    D = stein.all_pairs_eucl_dist(x)
    ref = np.sqrt(((x[:, None, :].astype(np.float64) - x[None].astype(np.float64)) ** 2).sum(-1))
    np.testing.assert_allclose(D, ref, atol=2e-3)      # Gram form: diagonal is rounding noise, not 0
    h = stein.median_bandwidth(x)
    want = np.float32(np.median(D)) ** 2 * np.float32(1 / math.log(m)) + np.float32(stein.BANDWIDTH_FLOOR)
    assert h == np.float32(want)
