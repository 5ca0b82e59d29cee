"""The block-symmetric row-sharded distance / force stages (smi_stein_sym_*, include/smi.h) with the ranks of a ring
EMULATED on one GPU: every "rank" gets its own workspace and runs exactly the stage sequence of
smi_experiments_b200.dist.stein_force_sym; the collectives (all-gather of the column-sum slabs, hand-over of the
transposed D2 blocks, integer all-reduces of the select) are plain tensor copies / sums.  The driver's GPU tier has one
GPU, so this is its record of the sharded kernels: column window that wraps around the matrix, operand orientation,
pack / unpack, slab-ordered mean.  Expected: bandwidth bit-identical to the unsharded path and force rows equal to it
(scripts/check_sharded.py shows the same over NCCL on 2 and 8 GPUs, profiles/r02d_check_sharded_n*.txt)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _emulate(cuda, m, d, world, alpha=2.0):
    from smi_experiments_b200 import _lib, ops
    from smi_experiments_b200.dist import LibsmiRows
    gen = torch.Generator(device=cuda).manual_seed(17 + world)
    X = torch.randn(m, d, device=cuda, generator=gen) + 0.7
    G = torch.randn(m, d, device=cuda, generator=gen)
    n = m // world
    half = world // 2
    right = list(range(1, half + 1))
    sendt = [o for o in right if not (world % 2 == 0 and o == half)]
    bes = [LibsmiRows(m, d, r * n, n, cuda) for r in range(world)]
    assert all(b.sym_supported() for b in bes)
    for b in bes:
        b.sym_setup(len(sendt))
        b.colsum(X)
    for b in bes:                      # "all-gather" of the slab partials
        for o in bes:
            b.part[o.slab0:o.slab1] = o.part[o.slab0:o.slab1]
    for r, b in enumerate(bes):
        b.mean()
        b.split(X, r * n, n)
        for o in right:
            b.split(X, ((r + o) % world) * n, n)
        # own + first right block, then the others: the piecewise window of a pipelined exchange
        b.gram(0, min(2, len(right) + 1))
        if len(right) > 1:
            b.gram(2, len(right) - 1)
        for s, o in enumerate(sendt):
            b.pack_t(((r + o) % world) * n, s)
    for r, b in enumerate(bes):        # hand-over of the transposed blocks to their other owner
        for s, o in enumerate(sendt):
            b.recvbuf[s].copy_(bes[(r - o) % world].sendbuf[s])
            b.unpack(s, ((r - o) % world) * n)
    # exact select with "all-reduced" integer histograms
    mode = _lib.MEDIAN_JAX_FP32_INDEX
    for p in range(3):
        hs = [b.hist(p).clone() for b in bes]
        tot = torch.stack(hs).sum(0).to(torch.int32)
        for b in bes:
            b.hist_buf.copy_(tot)
            b.scan(p, mode)
    tails = [tuple(t.clone() for t in b.tail()) for b in bes]
    cnt = torch.stack([t[0] for t in tails]).sum(0)
    mn = torch.stack([t[1] for t in tails]).min(0).values
    hs, Fs, sq = [], [], 0.0
    for b in bes:
        h = b.bandwidth(cnt.clone(), mn.clone(), mode, alpha, 0.0)
        hs.append(h.clone())
        F, s2 = b.force(X, G, 1.0)
        Fs.append(F.clone())
        sq += float(s2)
    F1, l1, h1 = ops.stein_force(X, G, alpha=alpha, path=ops.PATH_TC)
    return torch.cat(Fs), float(np.sqrt(sq)), hs, F1, float(l1), h1


@pytest.mark.parametrize("world,m,d", [(2, 8192, 200), (4, 8192, 200), (8, 8192, 64), (4, 16384, 100)])   # worlds must divide the 64 mean slabs
def test_block_symmetric_ring_matches_unsharded(cuda, world, m, d):
    F, loss, hs, F1, l1, h1 = _emulate(cuda, m, d, world)
    for h in hs:                                               # integer-exact select on identical distances
        assert h.view(torch.int32).item() == h1.view(torch.int32).item()
    assert torch.equal(F, F1)                                  # same accumulation sequence for every particle pair
    assert abs(loss - l1) <= 1e-6 * l1


def test_sym_rejects_unsupported_shapes(cuda, lib):
    assert lib.smi_stein_sym_supported(8192, 200, 2048) == 1
    assert lib.smi_stein_sym_supported(8192, 200, 1000) == 0       # rows % 256
    assert lib.smi_stein_sym_supported(1024, 10000, 256) == 0      # split-K shape (m <= 4096, long K)
    assert lib.smi_stein_sym_supported(8192, 200, 8192) == 0       # a world of one
