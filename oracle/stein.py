"""Oracle: RBF kernel with median-heuristic bandwidth and the Stein force.

Restates (from the published NumPyro algorithm):
  numpyro/contrib/einstein/stein_util.py   all_pairs_eucl_dist, median_bandwidth
  numpyro/contrib/einstein/stein_kernels.py RBFKernel (mode="norm")
  numpyro/contrib/einstein/steinvi.py      _apply_kernel, _kernel_grad, force sum
  jax/_src/numpy/reductions.py             median -> quantile(method="midpoint")
Reference call sites: ``RBFKernel()`` at experiments/uci_benchmark.py:168,
experiments/variance_collapse.py:83; ``repulsion_temperature`` at
experiments/variance_collapse.py:104,113,122.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Parity unpinned.
"""

import math

import numpy as np

EPS32 = float(np.finfo(np.float32).eps)  # 1.1920929e-07
# numpyro stein_util.median_bandwidth adds "10 * jnp.finfo(dists.dtype).eps" to median^2 * factor [NP?]: the ONE place
# this constant lives in the oracle (device side: kSmiBandwidthFloor in csrc/common.cuh).  Recalled, not verifiable here.
BANDWIDTH_FLOOR = 10.0 * EPS32

MEDIAN_EXACT_MIDPOINT = 0  # textbook midpoint of the two middle order statistics
MEDIAN_JAX_FP32_INDEX = 1  # q = 0.5f * float32(n - 1) computed in float32 [NP?]


def all_pairs_eucl_dist(x, dtype=np.float32):
    """D_ij = sqrt(max(|x_i|^2 + |x_j|^2 - 2 x_i.x_j, 0)), Gram form, full m x m."""
    x = np.asarray(x, dtype=dtype)
    sq = np.sum(x * x, axis=1, dtype=dtype)
    gram = (x @ x.T).astype(dtype)
    d2 = (sq[None, :] + sq[:, None]) - dtype(2) * gram
    return np.sqrt(np.maximum(d2, dtype(0))).astype(dtype)


def median_ranks(n, mode=MEDIAN_JAX_FP32_INDEX):
    """(lo, hi) 0-based order statistics averaged by jnp.median over n values."""
    if mode == MEDIAN_JAX_FP32_INDEX:
        q = np.float32(0.5) * np.float32(n - 1)
        return int(np.floor(q)), int(np.ceil(q))
    return (n - 1) // 2, n // 2


def median(values, mode=MEDIAN_JAX_FP32_INDEX):
    """jnp.median semantics on a flat array: (a[lo] + a[hi]) * 0.5 in the array dtype."""
    a = np.sort(np.asarray(values).ravel())
    lo, hi = median_ranks(a.size, mode)
    return ((a[lo] + a[hi]) * a.dtype.type(0.5)).astype(a.dtype)


def bandwidth_from_dists(dists, m, mode=MEDIAN_JAX_FP32_INDEX):
    """h = median(D)^2 * (1/ln m) + 10 eps32 ; m == 1 => 1.0."""
    dt = np.asarray(dists).dtype.type
    if m == 1:
        return dt(1.0)
    med = median(dists, mode)
    factor = dt(1.0 / math.log(m))
    return dt(med * med * factor + dt(BANDWIDTH_FLOOR))


def median_bandwidth(x, mode=MEDIAN_JAX_FP32_INDEX, dtype=np.float32):
    x = np.asarray(x, dtype=dtype)
    m = x.shape[0]
    if m == 1:
        return dtype(1.0)
    return bandwidth_from_dists(all_pairs_eucl_dist(x, dtype), m, mode)


def rbf_matrix(x, h, dtype=np.float32):
    """K_ij = exp(-|x_i - x_j|^2 / h), evaluated from differences (reference form)."""
    x = np.asarray(x, dtype=dtype)
    m = x.shape[0]
    k = np.empty((m, m), dtype=dtype)
    for i in range(m):
        diff = x - x[i]
        k[i] = np.exp(-np.sum(diff * diff, axis=1, dtype=dtype) / dtype(h))
    return k


def stein_force(x, g, alpha=1.0, c_att=1.0, mode=MEDIAN_JAX_FP32_INDEX,
                dtype=np.float32, h=None):
    """F = A + R, loss = |F|_F, with
         A_i = c_att * sum_j K_ij g_j
         R_i = alpha * mean_j grad_{x_j} k(x_j, x_i) = (2 alpha/(h m)) (x_i sum_j K_ij - sum_j K_ij x_j)
    Returns (F, loss, h, K)."""
    x = np.asarray(x, dtype=dtype)
    g = np.asarray(g, dtype=dtype)
    m = x.shape[0]
    if h is None:
        h = median_bandwidth(x, mode, dtype)
    h = dtype(h)
    k = rbf_matrix(x, h, dtype)
    att = dtype(c_att) * (k @ g)
    c = dtype(2.0 * alpha) / (h * dtype(m))
    rep = c * (x * np.sum(k, axis=1, dtype=dtype)[:, None] - k @ x)
    f = (att + rep).astype(dtype)
    loss = dtype(np.sqrt(np.sum(f.astype(np.float64) ** 2)))
    return f, loss, h, k


def stein_force_direct(x, g, h, alpha=1.0, c_att=1.0):
    """Float64 double-loop evaluation straight from the definitions (no algebra):
    used to pin stein_force() itself.  O(m^2 d) python loops: small cases only."""
    x = np.asarray(x, dtype=np.float64)
    g = np.asarray(g, dtype=np.float64)
    m, d = x.shape
    f = np.zeros((m, d))
    for i in range(m):
        for j in range(m):
            diff = x[j] - x[i]
            kij = math.exp(-float(diff @ diff) / h)
            f[i] += c_att * kij * g[j]
            # grad wrt first argument x_j of k(x_j, x_i): -2 (x_j - x_i)/h * k
            f[i] += alpha / m * (-2.0 * diff / h * kij)
    return f


def stein_force_torch(x, g, alpha=1.0, c_att=1.0, mode=MEDIAN_JAX_FP32_INDEX, threads=None, chunk=32):
    """The same algorithm as stein_force() in float32 on ALL host threads (torch CPU ops): Gram-form distances +
    full sort for the bandwidth, the kernel from DIFFERENCES (the m x m x d broadcast the reference's vmapped
    `kernel(x, y)` performs, in row chunks), two products for the force.  bench.py's CPU leg for the Stein stage
    (config C5) -- never the product path.  Returns (F, loss, h) as numpy / floats."""
    import torch
    if threads:
        torch.set_num_threads(threads)
    xt = torch.as_tensor(np.asarray(x, np.float32))
    gt = torch.as_tensor(np.asarray(g, np.float32))
    m = xt.shape[0]
    if m == 1:
        h = 1.0
    else:
        sq = (xt * xt).sum(1)
        d = torch.sqrt(torch.clamp((sq[None, :] + sq[:, None]) - 2.0 * (xt @ xt.T), min=0.0))
        a = torch.sort(d.reshape(-1)).values
        lo, hi = median_ranks(m * m, mode)
        med = (a[lo] + a[hi]) * 0.5
        h = float(np.float32(med.item()) * np.float32(med.item()) * np.float32(1.0 / math.log(m)) + np.float32(BANDWIDTH_FLOOR))
    k = torch.empty(m, m, dtype=torch.float32)
    for i0 in range(0, m, chunk):
        diff = xt[None, :, :] - xt[i0:i0 + chunk, None, :]
        k[i0:i0 + chunk] = torch.exp(-(diff * diff).sum(-1) / h)
    c = 2.0 * alpha / (h * m)
    f = c_att * (k @ gt) + c * (xt * k.sum(1)[:, None] - k @ xt)
    loss = float(torch.sqrt((f.double() ** 2).sum()))
    return f.numpy(), loss, h
