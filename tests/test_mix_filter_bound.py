"""The certificate behind the tensor-core mixture filter (csrc/elbo_mix_tc.cu), checked in NumPy: the expanded form
a_sj = [u^2, u mu'_i, u] . [q_j, 2 q_j, -2 mu'_j q_j] + C_ij equals the quadratic form, the sum of the absolute terms of
that dot product is bounded by P + 2 sqrt(P) (sqrt(D_ij) + sqrt(D_jj)) with P <= bmax^2 |eps|^2 (Cauchy-Schwarz), a
TF32-split (hi + lo, lo*lo dropped) evaluation stays inside eta = 2^-17 of that sum, and the resulting threshold never
drops a component whose float32 responsibility is non-zero."""
import numpy as np


def _split(x):
    hi = (x.astype(np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)
    return hi, (x.astype(np.float32) - hi)


def _case(rng, m, L, spread, S=6):
    mu = rng.uniform(-spread, spread, (m, L))
    sig = 0.1 * np.exp(0.3 * rng.standard_normal((m, L)))
    eps = rng.standard_normal((S, L))
    return mu, sig, eps


def test_expanded_form_bound_and_threshold():
    rng = np.random.default_rng(0)
    eta = 2.0 ** -17
    for m, L, spread in [(12, 400, 0.1), (12, 400, 2.0), (8, 3000, 0.1), (6, 50, 0.01), (10, 1200, 0.5)]:
        mu, sig, eps = _case(rng, m, L, spread)
        q = 1.0 / sig ** 2
        cent = mu.mean(0)
        for i in range(min(m, 3)):
            u = sig[i] * eps                                   # (S, L)
            mi = mu[i] - cent
            exact = (((u[:, None, :] + mu[i] - mu[None]) ** 2) * q[None]).sum(-1)            # (S, m)
            A = np.concatenate([u * u, u * mi, u], axis=1)                                      # (S, 3L)
            B = np.concatenate([q, 2 * q, -2 * (mu - cent) * q], axis=1)                        # (m, 3L)
            C = (q * (mu[i] - mu) ** 2).sum(1)                                                  # (m,)
            np.testing.assert_allclose(A @ B.T + C, exact, rtol=1e-9, atol=1e-9)
            # sum of absolute terms <= the bound assembled from per-pair constants and |eps|^2
            T = np.abs(A[:, None, :] * B[None]).sum(-1)
            E = (eps ** 2).sum(1)
            bmax = (sig[i] ** 2 * q).max(1)
            Dij, Djj = (mi ** 2 * q).sum(1), ((mu - cent) ** 2 * q).sum(1)
            P = bmax[None] * E[:, None]
            bound = P + 2 * np.sqrt(P) * (np.sqrt(Dij) + np.sqrt(Djj))[None]
            assert (T <= bound * (1 + 1e-12)).all()
            # 3xTF32 (lo*lo dropped), float32 accumulation: error within eta * T
            Ah, Al = _split(A)
            Bh, Bl = _split(B)
            approx = (Al @ Bh.T + Ah @ Bl.T + Ah @ Bh.T).astype(np.float64) + C
            assert (np.abs(approx - exact) <= eta * T + 1e-6 * np.abs(exact)).all()
            # the threshold keeps every component with a non-zero float32 responsibility
            ld = np.log(sig).sum(1)
            lq, lq_apx = -0.5 * exact - ld, -0.5 * approx - ld
            delta = 0.5 * eta * bound + 1e-6 * np.abs(lq_apx)
            thr = lq_apx.max(1) - 104.0 - 2.0 * delta.max(1) - 1.0
            keep = lq_apx >= thr[:, None]
            w = np.exp((lq - lq.max(1, keepdims=True)).astype(np.float32))
            assert not ((w != 0) & ~keep).any()
            assert keep[:, i].all()                           # the own component always survives
