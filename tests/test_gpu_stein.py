"""Parity of the CUDA Stein-force stage (rows a4-a7) against the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import stein as ostein

pytestmark = pytest.mark.gpu

# north_star tolerance: <= 1e-5 relative (fp32) for forces / bandwidth / loss
RTOL = 1e-5


def _rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _mk(m, d, seed, scale=1.0, collapse=False):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((m, d)).astype(np.float32) * np.float32(scale)
    if collapse:  # variance-collapse regime: particles nearly coincide around an offset
        x = (np.float32(3.0) + np.float32(1e-2) * x).astype(np.float32)
    g = rng.standard_normal((m, d)).astype(np.float32)
    return x, g


@pytest.mark.parametrize("m,d", [(1, 7), (2, 1), (5, 92), (5, 5402), (10, 179220), (20, 6604), (20, 20),
                                 (33, 1000), (64, 513)])
@pytest.mark.parametrize("alpha", [1.0, 10.0])
def test_small_path_matches_oracle(cuda, m, d, alpha):
    from smi_experiments_b200 import ops
    x, g = _mk(m, d, seed=m * 1000 + d)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), alpha=alpha,
                                 path=ops.PATH_SMALL)
    # oracle in float64 is the truth; the fp32 oracle calibrates what fp32 can reach
    F64, loss64, h64, _ = ostein.stein_force(x, g, alpha, 1.0, ostein.MEDIAN_JAX_FP32_INDEX, np.float64)
    assert abs(float(h.item()) - float(h64)) <= RTOL * float(h64)
    assert _rel(F.cpu().numpy(), F64) <= RTOL
    assert abs(float(loss.item()) - float(loss64)) <= RTOL * float(loss64)


@pytest.mark.parametrize("m,d", [(65, 300), (80, 6604), (160, 6604), (97, 1), (250, 130), (256, 1027)])
@pytest.mark.parametrize("alpha,mode", [(1.0, 1), (10.0, 0)])
def test_mid_path_matches_oracle(cuda, m, d, alpha, mode):
    """64 < m <= 256 (weak-scaling particle counts): CUDA-core difference-form path, exact radix-select median."""
    from smi_experiments_b200 import ops
    x, g = _mk(m, d, seed=m * 1000 + d)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), alpha=alpha,
                                 median_mode=mode, path=ops.PATH_SMALL)
    F64, loss64, h64, _ = ostein.stein_force(x, g, alpha, 1.0, mode, np.float64)
    assert abs(float(h.item()) - float(h64)) <= RTOL * float(h64)
    assert _rel(F.cpu().numpy(), F64) <= RTOL
    assert abs(float(loss.item()) - float(loss64)) <= RTOL * float(loss64)
    # AUTO takes the same path for these sizes, bit for bit
    F2, loss2, h2 = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), alpha=alpha,
                                    median_mode=mode)
    assert torch.equal(F, F2) and float(h2.item()) == float(h.item()) and float(loss2.item()) == float(loss.item())


def test_mid_path_collapsed_and_fixed_bandwidth(cuda):
    from smi_experiments_b200 import ops
    x, g = _mk(120, 700, seed=11, collapse=True)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda))
    hv = float(h.item())
    F64, loss64, _, _ = ostein.stein_force(x, g, 1.0, 1.0, dtype=np.float64, h=hv)
    assert _rel(F.cpu().numpy(), F64) <= RTOL
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), h_fixed=0.25)
    assert float(h.item()) == 0.25
    F64, loss64, _, _ = ostein.stein_force(x, g, dtype=np.float64, h=0.25)
    assert _rel(F.cpu().numpy(), F64) <= RTOL and abs(float(loss.item()) - float(loss64)) <= RTOL * float(loss64)


def test_small_path_collapsed_particles(cuda):
    from smi_experiments_b200 import ops
    x, g = _mk(20, 2000, seed=7, collapse=True)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda))
    # bandwidth comes from the Gram form in fp32 (as in the reference): compare h against the fp32
    # oracle loosely, and the force against the float64 oracle evaluated at the SAME bandwidth.
    hv = float(h.item())
    F64, loss64, _, _ = ostein.stein_force(x, g, 1.0, 1.0, dtype=np.float64, h=hv)
    assert _rel(F.cpu().numpy(), F64) <= RTOL
    assert abs(float(loss.item()) - float(loss64)) <= RTOL * float(loss64)


def test_fixed_bandwidth_and_m1(cuda):
    from smi_experiments_b200 import ops
    x, g = _mk(1, 50, seed=3)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda))
    assert float(h.item()) == 1.0                      # m == 1 => h = 1 (Appendix A.3)
    assert _rel(F.cpu().numpy(), g) <= 1e-6            # K = 1, no repulsion
    x, g = _mk(8, 64, seed=4)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), h_fixed=3.5)
    assert float(h.item()) == 3.5
    F64, _, _, _ = ostein.stein_force(x, g, dtype=np.float64, h=3.5)
    assert _rel(F.cpu().numpy(), F64) <= RTOL


@pytest.mark.parametrize("m", [2, 3, 5, 20, 64, 100, 257, 1000, 2048])
@pytest.mark.parametrize("mode", [0, 1])
def test_select_is_bit_exact(cuda, m, mode):
    """Bandwidth selection: the two order statistics picked on the device are exactly those
    numpy.sort picks on the same matrix (integer-exact radix select)."""
    from smi_experiments_b200 import ops
    rng = np.random.default_rng(m)
    d2 = (rng.standard_normal((m, m)).astype(np.float32) ** 2) * np.float32(50.0)
    if m > 3:
        d2[rng.integers(0, m, 5), rng.integers(0, m, 5)] = 0.0     # ties at zero
        d2[0, :m // 2] = d2[1, :m // 2]                              # duplicated values
    h, ranks = ops.median_bandwidth(torch.from_numpy(d2).to(cuda), mode)
    a = np.sort(d2.ravel())
    lo, hi = ostein.median_ranks(a.size, mode)
    got = ranks.cpu().numpy().view(np.uint32)
    want = np.array([a[lo], a[hi]], np.float32).view(np.uint32)
    assert (got == want).all()
    med = np.float32((np.sqrt(a[lo]) + np.sqrt(a[hi])) * np.float32(0.5))
    hw = np.float32(med * med * np.float32(1.0 / np.log(m)) + np.float32(ostein.BANDWIDTH_FLOOR))
    assert abs(float(h.item()) - float(hw)) <= 2e-7 * float(hw)


def test_select_fp32_index_edge(cuda):
    """m = 8192: jnp.median's float32 index arithmetic picks a single element (Appendix A.3)."""
    from smi_experiments_b200 import ops
    m = 8192
    gen = torch.Generator(device=cuda).manual_seed(5)
    d2 = torch.rand(m, m, device=cuda, generator=gen) * 100.0
    a = torch.sort(d2.reshape(-1)).values
    for mode in (0, 1):
        lo, hi = ostein.median_ranks(m * m, mode)
        _, ranks = ops.median_bandwidth(d2, mode)
        want = torch.stack([a[lo], a[hi]]).view(torch.int32)
        assert torch.equal(ranks, want)
    assert ostein.median_ranks(m * m, 1) == (33554432, 33554432)


# (700, 2100), (1024, 4100), (1500, 2500), (2048, 2304): split-K Gram (segments of the K loop as separate work
# units + gram_finish_kernel), ragged tile edges and an odd number of 128-row tiles (phantom half of a CTA pair)
@pytest.mark.parametrize("m,d", [(128, 64), (200, 100), (256, 1000), (1000, 333), (1024, 10000), (700, 2100),
                                 (1024, 4100), (1500, 2500), (2048, 2304)])
def test_tc_sqdist(cuda, m, d):
    from smi_experiments_b200 import ops
    rng = np.random.default_rng(m + d)
    x = rng.standard_normal((m, d)).astype(np.float32)
    D2 = ops.sqdist(torch.from_numpy(x).to(cuda), path=ops.PATH_TC).cpu().numpy()
    x64 = x.astype(np.float64)
    sq = (x64 * x64).sum(1)
    ref = np.maximum(sq[:, None] + sq[None, :] - 2.0 * (x64 @ x64.T), 0.0)
    off = ~np.eye(m, dtype=bool)
    # 3xTF32 keeps ~fp32 accuracy: error relative to the magnitude of the terms that cancel
    scale = (sq[:, None] + sq[None, :])
    assert np.max(np.abs(D2 - ref)[off] / scale[off]) <= 2e-6
    assert np.max(np.abs(D2)[~off] / scale[~off]) <= 2e-6
    # the reference's difference form gives an exactly symmetric matrix with an exactly zero diagonal
    assert (D2 == D2.T).all() and (np.diag(D2) == 0.0).all()


def test_tc_pair_kernel_matches_one_cta_kernel(cuda):
    """The CTA-pair GEMMs (cta_group::2) and the one-CTA kernel accumulate every element in the same order:
    with split-K off, squared distances, bandwidth and force agree bit for bit (subprocess: the switch is
    read once per process)."""
    import os, subprocess, sys, tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np, torch\n"
        "sys.path.insert(0, %r)\n"
        "from smi_experiments_b200 import ops\n"
        "rng = np.random.default_rng(5)\n"
        "x = torch.from_numpy(rng.standard_normal((777, 1300)).astype(np.float32)).cuda()\n"
        "g = torch.from_numpy(rng.standard_normal((777, 1300)).astype(np.float32)).cuda()\n"
        "D2 = ops.sqdist(x, path=ops.PATH_TC)\n"
        "F, loss, h = ops.stein_force(x, g, path=ops.PATH_TC)\n"
        "np.savez(sys.argv[1], D2=D2.cpu().numpy(), F=F.cpu().numpy(), loss=loss.cpu().numpy(), h=h.cpu().numpy())\n"
    ) % root
    outs = []
    with tempfile.TemporaryDirectory() as td:
        for pair in ("1", "0"):
            f = os.path.join(td, "o%s.npz" % pair)
            env = dict(os.environ, SMI_TC_PAIR=pair, SMI_GRAM_KSPLIT="1")
            subprocess.run([sys.executable, "-c", code, f], check=True, env=env, cwd=root, timeout=300)
            outs.append(dict(np.load(f)))
    for k in ("D2", "F", "loss", "h"):
        assert (outs[0][k].view(np.uint32) == outs[1][k].view(np.uint32)).all(), k


@pytest.mark.parametrize("m,d", [(65, 40), (128, 64), (200, 100), (512, 2000), (1000, 333), (1024, 10000)])
@pytest.mark.parametrize("alpha", [1.0])
def test_tc_path_matches_oracle(cuda, m, d, alpha):
    from smi_experiments_b200 import ops
    x, g = _mk(m, d, seed=m + d)
    F, loss, h = ops.stein_force(torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda), alpha=alpha,
                                 path=ops.PATH_TC)
    x64, g64 = x.astype(np.float64), g.astype(np.float64)
    sq = (x64 * x64).sum(1)
    d2 = np.maximum(sq[:, None] + sq[None, :] - 2.0 * (x64 @ x64.T), 0.0)
    a = np.sort(np.sqrt(d2).ravel())
    lo, hi = ostein.median_ranks(m * m, 1)
    h64 = (0.5 * (a[lo] + a[hi])) ** 2 / np.log(m) + ostein.BANDWIDTH_FLOOR
    assert abs(float(h.item()) - h64) <= RTOL * h64
    K = np.exp(-d2 / h64)
    c = 2.0 * alpha / (h64 * m)
    F64 = K @ (g64 - c * x64) + c * K.sum(1)[:, None] * x64
    assert _rel(F.cpu().numpy(), F64) <= RTOL
    assert abs(float(loss.item()) - np.linalg.norm(F64)) <= RTOL * np.linalg.norm(F64)


def test_tc_and_small_paths_agree(cuda):
    from smi_experiments_b200 import ops
    x, g = _mk(64, 3000, seed=11)
    xt, gt = torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda)
    F1, l1, h1 = ops.stein_force(xt, gt, path=ops.PATH_SMALL)
    F2, l2, h2 = ops.stein_force(xt, gt, path=ops.PATH_TC)
    assert abs(float(h1.item()) - float(h2.item())) <= RTOL * float(h1.item())
    assert _rel(F2.cpu().numpy(), F1.cpu().numpy()) <= RTOL


def test_full_size_sweep_point_properties(cuda):
    """BASELINE configs[4] at its full size (m = 16384 particles x d = 10000) through size-independent
    properties: exact symmetry / zero diagonal of the squared distances, sampled rows against float64, the
    two order statistics of the radix select against torch.sort on the same matrix (bit-exact), linearity of
    the force in the scores, and sampled force rows against the float64 definition at the device's bandwidth."""
    from smi_experiments_b200 import ops
    m, d = 16384, 10000
    gen = torch.Generator(device=cuda).manual_seed(4)
    X = torch.randn(m, d, device=cuda, generator=gen)
    G = torch.randn(m, d, device=cuda, generator=gen)
    D2 = ops.sqdist(X, path=ops.PATH_TC)
    assert bool((D2.diagonal() == 0).all())
    for r0 in range(0, m, 4096):                       # D2 == D2.T blockwise (keeps the temporaries small)
        assert bool((D2[r0:r0 + 4096] == D2[:, r0:r0 + 4096].T).all())
    rows = torch.tensor([0, 1, 127, 128, 255, 256, 4095, 8191, 8192, 12345, 16383], device=cuda)
    X64 = X.double()
    sq = (X64 * X64).sum(1)
    ref = (sq[rows, None] + sq[None, :] - 2.0 * (X64[rows] @ X64.T)).clamp_min(0)
    scale = sq[rows, None] + sq[None, :]
    err = ((D2[rows].double() - ref).abs() / scale)
    err[torch.arange(len(rows), device=cuda), rows] = 0
    assert float(err.max()) <= 2e-6
    # bandwidth: integer-exact order statistics of the m^2 values
    h, ranks = ops.median_bandwidth(D2, 1)
    keys = D2.reshape(-1).view(torch.int32)            # D2 >= 0: the bit patterns order like the values
    lo, hi = ostein.median_ranks(m * m, 1)
    srt = torch.sort(keys).values
    want = torch.stack([srt[lo], srt[hi]])
    del srt, keys
    assert bool((ranks.view(torch.int32) == want).all())
    # force: linear in the scores at fixed particles (h depends on X only), rows against float64
    F1, _, h1 = ops.stein_force(X, G, path=ops.PATH_TC)
    F0, _, h0 = ops.stein_force(X, torch.zeros_like(G), path=ops.PATH_TC)
    F3, loss3, h3 = ops.stein_force(X, 3.0 * G, path=ops.PATH_TC)
    assert float(h1) == float(h0) == float(h3) == float(h)
    lin = (F3 - F0) - 3.0 * (F1 - F0)
    assert float(lin.double().norm() / (3.0 * (F1 - F0)).double().norm()) <= RTOL
    assert abs(float(loss3) - float(F3.double().norm())) <= RTOL * float(loss3)
    hv = float(h1)
    c = 2.0 / (hv * m)
    Kr = torch.exp(-ref / hv)
    Kr[torch.arange(len(rows), device=cuda), rows] = 1.0
    Fref = Kr @ (G.double() - c * X64) + c * Kr.sum(1, keepdim=True) * X64[rows]
    assert float((F1[rows].double() - Fref).norm() / Fref.norm()) <= RTOL


@pytest.mark.parametrize("m,d", [(64, 40), (200, 100), (1000, 333), (1028, 64), (2052, 32)])
def test_triangle_select_matches_full_select(cuda, m, d):
    """The single-GPU force selects its bandwidth from the upper triangle of the symmetric D2 (off-diagonal elements
    counted twice); the staged entry point reads the whole matrix.  Same integer counts, so the same bits of h --
    also for particle counts that are not multiples of the 256-wide tiles."""
    from smi_experiments_b200 import ops
    x, g = _mk(m, d, seed=3 * m + d)
    xt, gt = torch.from_numpy(x).to(cuda), torch.from_numpy(g).to(cuda)
    _, _, h_tri = ops.stein_force(xt, gt, path=ops.PATH_TC)
    D2 = ops.sqdist(xt, path=ops.PATH_TC)
    assert torch.equal(D2, D2.T)
    h_full, _ = ops.median_bandwidth(D2, 1)
    assert h_tri.view(torch.int32).item() == h_full.view(torch.int32).item()
