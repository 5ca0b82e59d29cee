"""World-size-2 (and 3) gloo runs of the row-sharded Stein-force choreography on the CPU.

The collectives and the row partition are the product code (smi_experiments_b200.dist.
stein_force_rows); the per-rank compute is a NumPy stand-in for the CUDA stages that follows the
same radix-select protocol (11 + 11 + 10 bit histogram passes, tail count / min).  Checks:
bandwidth bit-identical to the single-process oracle select for every world size, force rows and
loss equal to the unsharded oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class NumpyRows:
    SHIFT, BITS = (21, 10, 0), (11, 11, 10)

    def __init__(self, m, d, row0, rows):
        self.m, self.d, self.row0, self.rows = m, d, row0, rows

    def dist(self, X):
        x = X.numpy().astype(np.float64)
        x = x - x.mean(0)
        sq = (x * x).sum(1)
        r = slice(self.row0, self.row0 + self.rows)
        d2 = np.maximum(sq[r, None] + sq[None, :] - 2.0 * x[r] @ x.T, 0.0).astype(np.float32)
        d2[np.arange(self.rows), np.arange(self.rows) + self.row0] = 0.0
        self.d2 = d2
        self.keys = d2.view(np.uint32).ravel()
        self.prefix, self.rank = 0, None

    # ---- block-symmetric stages (stand-ins for smi_stein_sym_*) ----
    def sym_setup(self, nsend, sendbuf=None):
        m, d, n = self.m, self.d, self.rows
        self.part = torch.zeros(m, d, dtype=torch.float64)          # one slab per row
        self.slab0, self.slab1 = self.row0, self.row0 + n
        self.sendbuf = torch.zeros(max(nsend, 1), n, n)
        self.recvbuf = torch.zeros(max(nsend, 1), n, n)
        self.xc = np.full((m, d), np.nan)
        self.d2 = np.full((n, m), np.nan, np.float32)
        self.prefix, self.rank = 0, None

    def colsum(self, X):
        r = slice(self.row0, self.row0 + self.rows)
        self.part[r] = X[r].double()

    def mean(self):
        self.mu = self.part.numpy().sum(0) / self.m

    def split(self, X, r0, nr):
        self.xc[r0:r0 + nr] = X[r0:r0 + nr].numpy().astype(np.float64) - self.mu

    def gram(self, first, count):
        n, world = self.rows, self.m // self.rows
        me = self.row0 // n
        xa = self.xc[self.row0:self.row0 + n]
        for o in range(first, first + count):
            c0 = ((me + o) % world) * n
            xb = self.xc[c0:c0 + n]
            assert not np.isnan(xb).any()                              # its particles must have arrived and been split
            blk = np.maximum((xa * xa).sum(1)[:, None] + (xb * xb).sum(1)[None, :] - 2.0 * xa @ xb.T, 0.0)
            self.d2[:, c0:c0 + n] = blk.astype(np.float32)
        self.d2[np.arange(n), np.arange(n) + self.row0] = 0.0

    def pack_t(self, col0, slot):
        self.sendbuf[slot] = torch.from_numpy(np.ascontiguousarray(self.d2[:, col0:col0 + self.rows].T))

    def unpack(self, slot, col0):
        self.d2[:, col0:col0 + self.rows] = self.recvbuf[slot].numpy()

    def hist(self, p):
        if p == 0:
            assert not np.isnan(self.d2).any()                         # every block computed or received
            self.keys = self.d2.view(np.uint32).ravel()
        sh, nb = self.SHIFT[p], self.BITS[p]
        k = self.keys
        if p > 0:
            k = k[(k >> np.uint32(sh + nb)) == np.uint32(self.prefix)]
        digit = (k >> np.uint32(sh)) & np.uint32((1 << nb) - 1)
        self.h = torch.from_numpy(np.bincount(digit, minlength=2048).astype(np.int32))
        return self.h

    def scan(self, p, median_mode):
        from oracle import stein
        if p == 0:
            self.rank = stein.median_ranks(self.m * self.m, median_mode)[0]
        c = np.cumsum(self.h.numpy().astype(np.int64))
        b = int(np.searchsorted(c, self.rank, side="right"))
        self.rank -= int(c[b - 1]) if b > 0 else 0
        self.prefix = (self.prefix << self.BITS[p]) | b

    def tail(self):
        a = np.uint32(self.prefix)
        above = self.keys[self.keys > a]
        cnt = torch.tensor([int((self.keys <= a).sum())], dtype=torch.int64)
        mn = torch.tensor([int(above.min()) if above.size else 0xFFFFFFFF], dtype=torch.int64)
        return cnt, mn

    def bandwidth(self, cnt, mn, median_mode, alpha, h_fixed):
        from oracle import stein
        if h_fixed > 0:
            self.hv = np.float32(h_fixed)
        else:
            hi = stein.median_ranks(self.m * self.m, median_mode)[1]
            a = np.array([self.prefix], np.uint32).view(np.float32)[0]
            b = a if hi < int(cnt) else np.array([int(mn)], np.uint32).view(np.float32)[0]
            self.sel = (a, b)
            med = np.float32((np.sqrt(a) + np.sqrt(b)) * np.float32(0.5))
            self.hv = np.float32(med * med * np.float32(1.0 / np.log(self.m)) + np.float32(stein.BANDWIDTH_FLOOR))
        self.alpha = alpha
        return torch.tensor([self.hv])

    def force(self, X, G, c_att):
        x, g = X.numpy().astype(np.float64), G.numpy().astype(np.float64)
        h = float(self.hv)
        K = np.exp(-self.d2.astype(np.float64) / h)
        c = 2.0 * self.alpha / (h * self.m)
        r = slice(self.row0, self.row0 + self.rows)
        F = K @ (c_att * g - c * x) + c * K.sum(1)[:, None] * x[r]
        return torch.from_numpy(F.astype(np.float32)), torch.tensor([float((F * F).sum())], dtype=torch.float64)


def _worker(rank, world, port, m, d, out, sym=False):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from smi_experiments_b200.dist import partition, stein_force_rows, stein_force_sym
    rng = np.random.default_rng(0)
    Xf = rng.standard_normal((m, d)).astype(np.float32) + np.float32(2.0)
    Gf = rng.standard_normal((m, d)).astype(np.float32)
    r0, n = partition(m, world, rank)
    # each rank starts with its own rows only; all-gather in place as the product does
    X, G = torch.zeros(m, d), torch.zeros(m, d)
    X[r0:r0 + n] = torch.from_numpy(Xf[r0:r0 + n])
    G[r0:r0 + n] = torch.from_numpy(Gf[r0:r0 + n])
    be = NumpyRows(m, d, r0, n)
    if sym:       # the exchange is inside: only the local rows are valid on entry
        F, loss, h = stein_force_sym(be, X, G, alpha=2.0)
        assert torch.equal(X, torch.from_numpy(Xf)) and torch.equal(G, torch.from_numpy(Gf))
    else:
        dist.all_gather_into_tensor(X, X[r0:r0 + n].clone())
        dist.all_gather_into_tensor(G, G[r0:r0 + n].clone())
        F, loss, h = stein_force_rows(be, X, G, alpha=2.0)
    torch.save(dict(F=F, loss=loss, h=h, sel=be.sel, r0=r0, n=n), os.path.join(out, "r%d.pt" % rank))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,m,sym", [(2, 24, False), (3, 24, False), (2, 90, False),
                                         # block-symmetric ring: even world (antipodal block computed by both owners),
                                         # odd world, and a world of 4 with one transposed block per rank
                                         (2, 24, True), (3, 24, True), (4, 24, True)])
def test_row_sharded_force_matches_unsharded(tmp_path, world, m, sym):
    from oracle import stein
    d = 40
    mp.spawn(_worker, args=(world, _free_port(), m, d, str(tmp_path), sym), nprocs=world, join=True)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((m, d)).astype(np.float32) + np.float32(2.0)
    G = rng.standard_normal((m, d)).astype(np.float32)
    # single-process reference of the same arithmetic (centred Gram form, exact-zero diagonal)
    one = NumpyRows(m, d, 0, m)
    one.dist(torch.from_numpy(X))
    a = np.sort(one.d2.ravel())
    lo, hi = stein.median_ranks(m * m, stein.MEDIAN_JAX_FP32_INDEX)
    F64, loss64, _, _ = stein.stein_force(X, G, alpha=2.0, dtype=np.float64)
    parts = [torch.load(os.path.join(tmp_path, "r%d.pt" % r), weights_only=False) for r in range(world)]
    hs = {float(p["h"]) for p in parts}
    assert len(hs) == 1                                            # every rank holds the same bandwidth
    for p in parts:
        assert p["sel"][0] == a[lo] and p["sel"][1] == a[hi]       # bit-exact order statistics
        assert abs(float(p["loss"]) - float(loss64)) <= 1e-5 * float(loss64)
        np.testing.assert_allclose(p["F"].numpy(), F64[p["r0"]:p["r0"] + p["n"]], rtol=2e-4, atol=2e-5)
    F = np.concatenate([p["F"].numpy() for p in parts])
    assert np.linalg.norm(F - F64) <= 1e-5 * np.linalg.norm(F64)


def test_partition_requires_even_split():
    from smi_experiments_b200.dist import partition
    assert partition(160, 8, 3) == (60, 20)
    with pytest.raises(ValueError):
        partition(20, 8, 0)
