"""Particle-row sharding across the GPUs of one node.

The reference is single-device; this is the scale-out the north star adds.  One process per GPU
(torch.distributed, NCCL over NVLink).  Rank r owns particle rows [r m/R, (r+1) m/R): their
optimizer state, their ELBO gradients and their rows of D, K and F.  Per step

  1. all-gather particles X                  (m d fp32 across the node)
  2. all-gather scores G  -- on a side stream, overlapped with 3
  3. local rows of the Gram / distance matrix; exact median by radix select:
       3 x (local histogram -> integer all-reduce SUM of 2048 bins -> scan), then a tail pass
       (all-reduce SUM of count(<= a), MIN of min(> a))            -- integers: h is bit-identical
       for every world size
  4. local rows of the force; all-reduce SUM of sum F^2 (double)  -> loss
  5. local optimizer update

`stein_force_rows` holds only the choreography; the per-rank compute comes from a backend
object (LibsmiRows: the CUDA kernels through the C ABI).  The CPU test-suite drives the same
choreography over gloo with a NumPy stand-in backend (tests/test_dist_gloo.py).
"""

import contextlib
import ctypes as C
import math
import os

import torch
import torch.distributed as dist

from . import _lib, ops
from .optim import OptimState
from .steinvi import SteinVI, SteinVIState


def partition(m, world, rank):
    """Contiguous equal row blocks (m must divide evenly: all-gathers are in place)."""
    if m % world != 0:
        raise ValueError("num_stein_particles (%d) must be a multiple of the world size (%d)" % (m, world))
    per = m // world
    return rank * per, per


class LibsmiRows:
    """Per-rank stages of the row-sharded Stein force on the GPU (smi_stein_rows_* in include/smi.h)."""

    def __init__(self, m, d, row0, rows, device):
        self.m, self.d, self.row0, self.rows, self.device = m, d, row0, rows, device
        lib = _lib.load()
        self.ws = torch.empty(int(lib.smi_stein_rows_workspace_bytes(rows, m, d)), dtype=torch.uint8, device=device)
        self.hist_buf = torch.zeros(2048, dtype=torch.int32, device=device)
        self.cnt_min = torch.zeros(2, dtype=torch.int32, device=device)
        self.h = torch.zeros(1, dtype=torch.float32, device=device)
        self.sumsq = torch.zeros(1, dtype=torch.float64, device=device)
        self.F = torch.empty(rows, d, dtype=torch.float32, device=device)
        self.cnt64 = torch.zeros(1, dtype=torch.int64, device=device)
        self.min64 = torch.zeros(1, dtype=torch.int64, device=device)

    def _w(self):
        return ops._ptr(self.ws), self.ws.numel()

    def dist(self, X):
        _lib.call("smi_stein_rows_dist", ops._ptr(X), self.m, self.d, self.row0, self.rows, *self._w(), ops._stream())

    def hist(self, p):
        _lib.call("smi_stein_rows_hist", self.m, self.d, self.row0, self.rows, p, ops._ptr(self.hist_buf), *self._w(),
                  ops._stream())
        return self.hist_buf

    def scan(self, p, median_mode):
        _lib.call("smi_stein_rows_scan", self.m, self.d, self.row0, self.rows, p, median_mode,
                  ops._ptr(self.hist_buf), *self._w(), ops._stream())

    def tail(self):
        """(count(<= a), min(> a)) of the local rows as int64 tensors (0xffffffff = none above)."""
        _lib.call("smi_stein_rows_tail", self.m, self.d, self.row0, self.rows, ops._ptr(self.cnt_min), *self._w(),
                  ops._stream())
        wide = self.cnt_min.to(torch.int64) & 0xFFFFFFFF
        self.cnt64.copy_(wide[0:1])
        self.min64.copy_(wide[1:2])
        return self.cnt64, self.min64

    def bandwidth(self, cnt64, min64, median_mode, alpha, h_fixed):
        if cnt64 is not None:
            self.cnt_min.copy_(torch.cat([cnt64, min64]).to(torch.int32))     # wraps back to the uint32 bits
        _lib.call("smi_stein_rows_bandwidth", self.m, self.d, self.row0, self.rows, median_mode, alpha, h_fixed,
                  ops._ptr(self.cnt_min[0:1]), ops._ptr(self.cnt_min[1:2]), ops._ptr(self.h), *self._w(),
                  ops._stream())
        return self.h

    def force(self, X, G, c_att):
        _lib.call("smi_stein_rows_force", ops._ptr(X), ops._ptr(G), self.m, self.d, self.row0, self.rows, c_att,
                  ops._ptr(self.F), ops._ptr(self.sumsq), *self._w(), ops._stream())
        return self.F, self.sumsq

    # ---- block-symmetric stages (smi_stein_sym_* in include/smi.h) ----
    def sym_supported(self):
        return bool(_lib.load().smi_stein_sym_supported(self.m, self.d, self.rows))

    def sym_setup(self, nsend, sendbuf=None):
        """Buffers of the block-symmetric choreography: the slab partials of the column mean (64 x d doubles, all-gathered
        in place) and the packed transposed D2 blocks that travel to / from their other owner (`sendbuf`: a
        symmetric-memory buffer the peers pull from, PeerPull.ts)."""
        if getattr(self, "part", None) is None:
            self.nslabs = 64
            per = self.m // 64
            self.slab0, self.slab1 = self.row0 // per, (self.row0 + self.rows) // per
            self.part = torch.zeros(64, self.d, dtype=torch.float64, device=self.device)
            self.sendbuf = sendbuf if sendbuf is not None else torch.empty(max(nsend, 1), self.rows, self.rows,
                                                                           dtype=torch.float32, device=self.device)
            self.recvbuf = torch.empty(max(nsend, 1), self.rows, self.rows, dtype=torch.float32, device=self.device)

    def colsum(self, X):
        _lib.call("smi_stein_sym_colsum", ops._ptr(X), self.m, self.d, self.row0, self.rows, ops._ptr(self.part),
                  *self._w(), ops._stream())

    def mean(self):
        _lib.call("smi_stein_sym_mean", ops._ptr(self.part), self.m, self.d, self.row0, self.rows, *self._w(),
                  ops._stream())

    def split(self, X, r0, nr):
        _lib.call("smi_stein_sym_split", ops._ptr(X), self.m, self.d, self.row0, self.rows, r0, nr, *self._w(),
                  ops._stream())

    def gram(self, first, count):
        _lib.call("smi_stein_sym_gram", self.m, self.d, self.row0, self.rows, first, count, *self._w(), ops._stream())

    def pack_t(self, col0, slot):
        _lib.call("smi_stein_sym_pack_t", self.m, self.d, self.row0, self.rows, col0, ops._ptr(self.sendbuf[slot]),
                  *self._w(), ops._stream())

    def unpack(self, slot, col0):
        _lib.call("smi_stein_sym_unpack", ops._ptr(self.recvbuf[slot]), self.m, self.d, self.row0, self.rows, col0,
                  *self._w(), ops._stream())


class _SideStream:
    """A second CUDA stream for the collectives that travel under compute (no-op for CPU tensors: the gloo tests)."""

    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        self.stream = torch.cuda.Stream(device=device) if self.cuda else None

    def fork(self):
        if self.cuda:
            self.stream.wait_stream(torch.cuda.current_stream())

    def ctx(self):
        return torch.cuda.stream(self.stream) if self.cuda else contextlib.nullcontext()

    def event(self):
        if not self.cuda:
            return None
        e = torch.cuda.Event()
        e.record(self.stream)
        return e

    @staticmethod
    def join(e):
        if e is not None:
            torch.cuda.current_stream().wait_event(e)


class PeerPull:
    """Bulk movement of the row-sharded Stein force over NVLink peer memory instead of NCCL: every rank publishes its
    particle rows, score rows and packed transposed D2 blocks in torch symmetric-memory buffers (CUDA IPC mappings),
    and the ranks that need them PULL them with cudaMemcpyAsync on side streams -- copy engines, zero SMs.  NCCL's
    kernels need SMs, and the persistent Gram / force GEMMs hold every SM with one 198 KB CTA: measured on 8 B200s
    (profiles/r02f_sym_n8.json) the grouped ncclSend / ncclRecv of the particle blocks slowed from 1.2 ms to 2.0 ms
    under the Gram and the all-gather of the scores from 1.0 ms to 1.5 ms, while a pull of one 82 MB block runs at
    645 GB/s whatever the SMs do (scripts/probe/symm_probe.py)."""

    def __init__(self, rows, d, nsend, device, group=None, nstreams=1):
        import torch.distributed._symmetric_memory as symm
        g = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.xs = symm.empty(rows, d, dtype=torch.float32, device=device)
        self.gs = symm.empty(rows, d, dtype=torch.float32, device=device)
        self.ts = symm.empty(max(nsend, 1), rows, rows, dtype=torch.float32, device=device)
        hx, hg, ht = symm.rendezvous(self.xs, g), symm.rendezvous(self.gs, g), symm.rendezvous(self.ts, g)
        peers = [p for p in range(self.world) if p != self.rank]
        self.peer_x = {p: hx.get_buffer(p, (rows, d), torch.float32) for p in peers}
        self.peer_g = {p: hg.get_buffer(p, (rows, d), torch.float32) for p in peers}
        self.peer_t = {p: ht.get_buffer(p, (max(nsend, 1), rows, rows), torch.float32) for p in peers}
        self._h = (hx, hg, ht)
        self.streams = [torch.cuda.Stream(device=device) for _ in range(nstreams)]
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)

    def fork(self):
        cur = torch.cuda.current_stream()
        for s in self.streams:
            s.wait_stream(cur)

    def pull(self, jobs):
        """jobs: [(dst, src)] copied round-robin over the side streams; returns one event per stream."""
        for k, (dst, src) in enumerate(jobs):
            with torch.cuda.stream(self.streams[k % len(self.streams)]):
                dst.copy_(src, non_blocking=True)
        evs = []
        for s in self.streams:
            e = torch.cuda.Event()
            e.record(s)
            evs.append(e)
        return evs


def _all_gather_in_place(out, rows, group):
    """all_gather_into_tensor whose input is this rank's slice of the output (NCCL takes the alias; gloo wants a copy)."""
    dist.all_gather_into_tensor(out, rows if out.is_cuda else rows.clone(), group=group)


def _exchange(pairs, group):
    """pairs: [(send_tensor, dst, recv_tensor, src)]: one grouped ncclSend / ncclRecv batch; returns when the current
    stream may use the received data."""
    if not pairs:
        return
    p2p = []
    for snd, dst, rcv, src in pairs:
        p2p.append(dist.P2POp(dist.isend, snd, dst, group))
        p2p.append(dist.P2POp(dist.irecv, rcv, src, group))
    for w in dist.batch_isend_irecv(p2p):
        w.wait()


def stein_force_sym(backend, X, G, group=None, alpha=1.0, c_att=1.0, median_mode=_lib.MEDIAN_JAX_FP32_INDEX,
                    h_fixed=0.0, side=None, trace=None, pull=None, bulk_group=None):
    """Row-sharded Stein force with the exchange INSIDE: on entry only the local rows of X and G are valid.
    Block-symmetric ring (include/smi.h, smi_stein_sym_*): the particle blocks to the right of this rank arrive
    first (grouped ncclSend / ncclRecv on a side stream) and feed the Gram, which computes the own diagonal block and
    world / 2 blocks to its right; the blocks to the left come back as packed transposes from the ranks that
    computed them; the remaining particle blocks and the all-gather of the scores travel under the Gram and are
    only awaited before the force.  Returns (F_rows, loss, h); X and G are complete afterwards.
    bulk_group: a second process group over the same ranks (its own NCCL communicator) for the grouped ncclSend /
    ncclRecv of the right blocks, so that it starts at once instead of queueing behind the small collectives of
    `group` -- the collectives of ONE communicator execute in issue order, whatever stream they are on."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n, r0 = backend.rows, backend.row0
    half = world // 2
    right = list(range(1, half + 1))
    sendt = [o for o in right if not (world % 2 == 0 and o == half)]     # the antipodal block is computed by both owners
    backend.sym_setup(len(sendt))
    side = side if side is not None else _SideStream(X.device)
    def mark(name, stream=None):        # optional timeline (scripts/time_sharded.py): CUDA events at the stage boundaries
        if trace is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream) if stream is not None else e.record()
            trace.append((name, e))
    mark("start")
    blk = lambda t, j: t[(j % world) * n:(j % world) * n + n]
    gr = lambda j: dist.get_global_rank(group, j % world) if group is not None else j % world
    joins = lambda evs: [torch.cuda.current_stream().wait_event(e) for e in evs]
    nn = 0
    if pull is not None:
        # ---- the right blocks the Gram waits for: with a communicator of their own all of them go by grouped
        # ncclSend / ncclRecv at once (every rank sends its own rows: nothing to wait for, and the SMs are idle);
        # otherwise half of them, after the small all-gather below, the other half by copy engine ----
        nn = len(right) if bulk_group is not None else len(right) // 2
        if world <= 3:
            nn = 0      # a single right block: one ncclSend / ncclRecv pair runs at ~250 GB/s, the copy engine at 645
        bg = bulk_group if bulk_group is not None else group
        bgr = lambda j: dist.get_global_rank(bg, j % world) if bg is not None else j % world

        def nccl_right():
            side.fork()
            with side.ctx():
                _exchange([(blk(X, rank), bgr(rank - o), blk(X, rank + o), bgr(rank + o)) for o in right[:nn]], bg)
                e = side.event()
                mark("side: right blocks by NCCL arrived", side.stream if side.cuda else None)
            return e
        e_n = nccl_right() if (bulk_group is not None and nn) else None
        # publish the local rows for the peers' copy-engine pulls
        pull.xs.copy_(blk(X, rank))
        pull.gs.copy_(blk(G, rank))
    # ---- column mean from slab partials (every rank adds the 64 slabs in the same order).  Issued BEFORE any bulk
    # NCCL exchange on the same communicator.  Its completion also tells this rank that every peer has published
    # its rows ----
    backend.colsum(X)
    _all_gather_in_place(backend.part, backend.part[backend.slab0:backend.slab1], group)
    if pull is not None:
        # ---- bulk movement under the GEMMs by copy engine.  Copies on different streams were measured to share ONE
        # copy engine (no concurrency, 0.22 ms per 82 MB block when all 8 ranks pull), so the pulls go on one stream
        # in priority order ----
        pull.fork()
        if e_n is None and nn:
            e_n = nccl_right()
        ev_a = pull.pull([(blk(X, rank + o), pull.peer_x[(rank + o) % world]) for o in right[nn:]])
        mark("side: right blocks by copy engine arrived", pull.streams[0])
        ev_b = pull.pull([(blk(X, rank + o), pull.peer_x[(rank + o) % world]) for o in range(half + 1, world)] +
                         [(blk(G, rank + o), pull.peer_g[(rank + o) % world]) for o in range(1, world)])
        mark("side: left blocks + G pulled", pull.streams[0])
        wait_a, wait_b = (lambda: (side.join(e_n), joins(ev_a))), (lambda: joins(ev_b))     # join(None) is a no-op
    else:
        # ---- bulk communication over NCCL / gloo on the side stream: right blocks, the rest, G ----
        side.fork()
        with side.ctx():
            ex = lambda offs: _exchange([(blk(X, rank), gr(rank - o), blk(X, rank + o), gr(rank + o)) for o in offs], group)
            ex(right)
            e_a = side.event()
            mark("side: right blocks arrived", side.stream)
            ex(range(half + 1, world))
            _all_gather_in_place(G, blk(G, rank), group)
            e_b = side.event()
            mark("side: left blocks + G arrived", side.stream)
        wait_a, wait_b = (lambda: side.join(e_a)), (lambda: side.join(e_b))
    backend.mean()
    backend.split(X, r0, n)
    mark("mean + own split")
    wait_a()
    mark("waited for the right blocks")
    for o in right:
        backend.split(X, ((rank + o) % world) * n, n)
    backend.gram(0, len(right) + 1)
    mark("split + gram")
    # ---- every unordered block pair was computed once: hand the transposes to their other owner ----
    for s_, o in enumerate(sendt):
        backend.pack_t(((rank + o) % world) * n, s_)
    # (NCCL, not the copy engine: its queue still holds the bulk pulls, and the SMs are free between Gram and select)
    if True:
        _exchange([(backend.sendbuf[s_], gr(rank + o), backend.recvbuf[s_], gr(rank - o)) for s_, o in enumerate(sendt)], group)
    for s_, o in enumerate(sendt):
        backend.unpack(s_, ((rank - o) % world) * n)
    mark("D2 blocks exchanged")

    def wait_rest():
        mark("select + bandwidth")
        wait_b()
        mark("waited for left blocks + G")
    out = _select_and_force(backend, X, G, group, alpha, c_att, median_mode, h_fixed, wait_rest)
    mark("force + loss")
    return out


def stein_force_rows(backend, X, G, group=None, alpha=1.0, c_att=1.0, median_mode=_lib.MEDIAN_JAX_FP32_INDEX,
                     h_fixed=0.0, wait_G=None):
    """Row-sharded Stein force.  X, G: ALL m particles / scores (already all-gathered; G may still
    be in flight -- `wait_G()` is called right before the force stage).  Returns (F_rows, loss, h)."""
    backend.dist(X)
    return _select_and_force(backend, X, G, group, alpha, c_att, median_mode, h_fixed, wait_G)


def _select_and_force(backend, X, G, group, alpha, c_att, median_mode, h_fixed, wait_G):
    m = backend.m
    cnt = mn = None
    if not (h_fixed > 0.0 or m == 1):
        for p in range(3):
            hist = backend.hist(p)
            dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)     # integers: exact, order-free
            backend.scan(p, median_mode)
        cnt, mn = backend.tail()
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
    h = backend.bandwidth(cnt, mn, median_mode, alpha, h_fixed)
    if wait_G is not None:
        wait_G()
    F, sumsq = backend.force(X, G, c_att)
    dist.all_reduce(sumsq, op=dist.ReduceOp.SUM, group=group)
    loss = torch.sqrt(sumsq).to(torch.float32)
    return F, loss, h


class ShardedSteinVI(SteinVI):
    """SteinVI with the particles row-sharded over the ranks of the default process group.
    Same surface as SteinVI; `state.optim_state.x` is the full (m, d) particle matrix (replicated,
    all-gathered every step), the optimizer slots cover the local rows only."""

    def __init__(self, *args, group=None, **kwargs):
        # m <= 256 (every rank evaluates the small / mid Stein path after one all-gather of the scores): the step --
        # two ncclAllGather included -- is captured into a CUDA graph like the single-GPU step; at ~25 launches and
        # ~0.5 ms per step the eager Python / ctypes issue rate was the limit at N = 4 (kernels 542 us, step 623 us).
        # SMI_SHARD_GRAPH=0 keeps eager launches.
        kwargs.setdefault("use_cuda_graph", os.environ.get("SMI_SHARD_GRAPH", "1") != "0")
        super().__init__(*args, **kwargs)
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.row0, self.nrows = partition(self.num_stein_particles, self.world, self.rank)
        self._rows_backend = None
        self._side = None

    def _graphable(self):
        return self.num_stein_particles <= 256

    def init(self, rng_key, *args, **kwargs):
        st = super().init(rng_key, *args, **kwargs)         # every rank draws the same full init
        x = st.optim_state.x
        r0, n = self.row0, self.nrows
        self._state = SteinVIState(OptimState(st.optim_state.step, x, torch.zeros_like(x[r0:r0 + n]),
                                              torch.zeros_like(x[r0:r0 + n])), st.rng_key)
        return self._state

    def _own(self, state):
        if getattr(self, "_state", None) is not None and state.optim_state.x is self._state.optim_state.x:
            return self._state
        raise ValueError("ShardedSteinVI updates the state its own init() returned")

    def _alloc(self, data):
        b = super()._alloc(data)
        if "G_all" not in b:
            m, S, L, d = self.num_stein_particles, self.num_elbo_particles, self.layout.latent_dim, self.layout.particle_dim
            dev, n = self.device, self.nrows
            normal = self.guide.kind == _lib.GUIDE_NORMAL
            b["eps"] = torch.empty(n, S, L, dtype=torch.float32, device=dev) if normal else None
            b["idx"] = torch.empty(n, S, data["B"], dtype=torch.int32, device=dev) if data["idx"] else None
            b["G_all"] = b["G"]
            b["elbo"] = torch.empty(n, dtype=torch.float32, device=dev)
        return b

    def _launch(self, st, b, data, do_update, ring=None):
        m, S, d = self.num_stein_particles, self.num_elbo_particles, self.layout.particle_dim
        r0, n = self.row0, self.nrows
        os_ = st.optim_state
        X = os_.x
        ops.prng_step(self.md, st.rng_key, b["key_next"], m, S, data["n_total"], data["B"] if data["idx"] else 0,
                      r0, n, b["eps"], b["idx"])
        G_rows = b["G_all"][r0:r0 + n]
        ops.elbo_grad(self.md, X, b["eps"], b["idx"], data["x"], data["y"], S, max(data["B"], 1), data["lik_scale"],
                      self._tau(), r0, n, G_rows, b["elbo"], self._ws_elbo)
        if m <= 256:
            # small m: every rank evaluates the (tiny) full force after gathering the scores (one cluster
            # launch up to m = 64 when the set fits, the three mid-path launches up to 256)
            dist.all_gather_into_tensor(b["G_all"], G_rows, group=self.group)
            nbytes = _lib.load().smi_stein_force_workspace_bytes(m, d, _lib.PATH_SMALL)
            w = self._ws_stein.get(nbytes, self.device)
            _lib.call("smi_stein_force_f32", ops._ptr(X), ops._ptr(b["G_all"]), m, d, self.c_att,
                      self.repulsion_temperature, self.kernel_fn.median_mode, self.kernel_fn.h_fixed, _lib.PATH_SMALL,
                      ops._ptr(b["F"]), ops._ptr(b["loss"]), ops._ptr(b["h"]), ops._ptr(w), w.numel(), ops._stream())
            F_rows = b["F"][r0:r0 + n]
        else:
            if self._rows_backend is None:
                self._rows_backend = LibsmiRows(m, d, r0, n, self.device)
                self._side = torch.cuda.Stream(device=self.device)
            cur = torch.cuda.current_stream()
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):         # scores travel while the distances are computed
                dist.all_gather_into_tensor(b["G_all"], G_rows, group=self.group)
            F_rows, loss, h = stein_force_rows(self._rows_backend, X, b["G_all"], self.group,
                                               self.repulsion_temperature, self.c_att, self.kernel_fn.median_mode,
                                               self.kernel_fn.h_fixed, wait_G=lambda: cur.wait_stream(self._side))
            b["loss"].copy_(loss)
            b["h"].copy_(h)
        if do_update:
            rows_state = OptimState(os_.step, X[r0:r0 + n], os_.a, os_.b)
            es = getattr(self, "_es", None)
            skip = es["skip"] if es is not None else None
            self.optim.update_gated(F_rows, rows_state, -1.0, skip)
            if es is None:
                ops.step_commit(b["loss"], os_.step, ring, b["key_next"], st.rng_key)
            else:   # every rank holds the same loss, so every rank takes the same stopping decision
                ops.step_commit(b["loss"], os_.step, ring, b["key_next"], st.rng_key, es["avg"], skip,
                                es["fast_decay"], es["slow_decay"], es["period"])
            dist.all_gather_into_tensor(X, X[r0:r0 + n], group=self.group)      # in place


def bench_sharded_force(m, d, reps, flush):
    """bench.py's Stein-force sweep point at N > 1: rank r owns rows [r0, r0 + n) of X, G; one timed unit =
    all-gather X, all-gather G (overlapped), row-sharded force.  Max over ranks.  Every rank draws the SAME
    full X, G (seed 4, as the N = 1 point does), so h and the loss are those of the N = 1 problem."""
    import numpy as np
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    r0, n = partition(m, world, rank)
    gen = torch.Generator(device=dev).manual_seed(4)
    X = torch.randn(m, d, device=dev, generator=gen)
    G = torch.randn(m, d, device=dev, generator=gen)
    keep = (X[r0:r0 + n].clone(), G[r0:r0 + n].clone())
    be = LibsmiRows(m, d, r0, n, dev)
    side = torch.cuda.Stream(device=dev)
    sym = be.sym_supported() and os.environ.get("SMI_SHARD_SYM", "1") != "0"
    side2 = _SideStream(dev)
    pull = bulk = None
    if sym and world > 3 and os.environ.get("SMI_SHARD_BULK_COMM", "1") != "0":
        bulk = dist.new_group()      # its own NCCL communicator for the right-block exchange
    if sym and os.environ.get("SMI_SHARD_PULL", "1") != "0":
        half = world // 2
        nsend = len([o for o in range(1, half + 1) if not (world % 2 == 0 and o == half)])
        try:
            pull = PeerPull(n, d, nsend, dev)
        except Exception as ex:      # no CUDA IPC on this box: the NCCL exchange remains
            if rank == 0:
                print("PeerPull unavailable (%s): NCCL exchange" % ex, flush=True)

    def scrub():
        # only the local rows are valid before the unit: the gathers must deliver the rest
        X.zero_(), G.zero_()
        X[r0:r0 + n], G[r0:r0 + n] = keep

    def unit():
        if sym:       # block-symmetric ring: the exchange is part of the choreography
            return stein_force_sym(be, X, G, side=side2, pull=pull, bulk_group=bulk)
        cur = torch.cuda.current_stream()
        dist.all_gather_into_tensor(X, X[r0:r0 + n])
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            dist.all_gather_into_tensor(G, G[r0:r0 + n])
        return stein_force_rows(be, X, G, wait_G=lambda: cur.wait_stream(side))

    for _ in range(2):
        scrub()
        unit()
    torch.cuda.synchronize()
    dist.barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        scrub()
        flush()
        dist.barrier()
        a.record()
        F, loss, h = unit()
        b.record()
    torch.cuda.synchronize()
    # per-kernel durations of one more unit on this rank (libsmi's event profiler; untimed)
    lib = _lib.load()
    lib.smi_profile(1)
    unit()
    torch.cuda.synchronize()
    lib.smi_profile(0)
    buf = C.create_string_buffer(1 << 16)
    lib.smi_profile_report(buf, len(buf))
    kern = {}
    for ln in buf.value.decode().splitlines():
        nm, cnt, tms = ln.split()
        kern[nm] = dict(launches=int(cnt), ms_per_step=float(tms))
    timeline = None
    if sym:       # stage timeline of one more unit on this rank (ms since the start of the unit)
        tr = []
        scrub()
        dist.barrier()
        stein_force_sym(be, X, G, side=side2, trace=tr, pull=pull, bulk_group=bulk)
        torch.cuda.synchronize()
        timeline = {nm: round(tr[0][1].elapsed_time(e), 3) for nm, e in tr[1:]}
    ms = torch.tensor([float(np.mean([a.elapsed_time(b) for a, b in evs]))], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return dict(ms=float(ms), loss=float(loss), h=float(h), kernels=kern, X=X, G=G, F_rows=F, row0=r0, timeline=timeline,
                path=("block-symmetric ring, " + ("copy-engine pulls over symmetric memory" if pull is not None else "NCCL exchange"))
                if sym else "row blocks, all columns",
                collectives=("side stream: grouped ncclSend/ncclRecv of the particle blocks right of the rank (before the "
                             "Gram), then of the other blocks, then ncclAllGather(G); main stream: ncclAllGather(64 x d "
                             "column-sum slabs), grouped ncclSend/ncclRecv of the transposed D2 blocks, " if sym else
                             "ncclAllGather(X) blocking, ncclAllGather(G) on a side stream, ") +
                            "3 x ncclAllReduce(SUM, 2048 x int32 histogram), ncclAllReduce(SUM count) + ncclAllReduce(MIN), "
                            "ncclAllReduce(SUM, double sum F^2)")
