"""SteinVI / SVGD / ASVGD: the engine surface the reference drives.

Mirrors numpyro.contrib.einstein.{SteinVI, SVGD, ASVGD} as used at
  experiments/uci_benchmark.py:188-195, experiments/img_classifier.py:160-177,
  experiments/variance_collapse.py:97-123, experiments/synthetic/in_between_uncertainty.py:119-153,
  experiments/synthetic/smi_recovery.py:77-86, src/methods/{svgd,asvgd}.py,
and the calls src/engine_run.py:25-48 and src/model_select/early_stopping.py:16-22,42 make on it:
init / evaluate / update / run / setup_run / get_params.

One `update` is a fixed sequence of libsmi launches on the current CUDA stream
  prng_step -> elbo_grad -> stein_force -> optimizer -> loss_record
with every scalar (loss, bandwidth, step counter, PRNG key) resident on the device, so the
sequence is captured once into a CUDA graph and replayed; nothing synchronises with the host
unless the caller reads the loss (the reference does, src/engine_run.py:56).

State handling differs from JAX in one documented way: buffers are updated IN PLACE (the
equivalent of jit with donated arguments); `update` returns a SteinVIState that aliases the
engine's buffers.
"""

import ctypes as C
import math
from collections import namedtuple
from functools import partial

import numpy as np
import torch

from . import _lib, models, ops
from .kernels import RBFKernel
from .models import AutoDelta, AutoNormal, init_to_uniform
from .optim import OptimState

SteinVIState = namedtuple("SteinVIState", ["optim_state", "rng_key"])
SteinVIRunResult = namedtuple("SteinVIRunResult", ["params", "state", "losses"])


def PRNGKey(seed):
    """jax.random.PRNGKey(seed) for a 32-bit seed: uint32[2] = (0, seed)."""
    return np.array([(int(seed) >> 32) & 0xFFFFFFFF, int(seed) & 0xFFFFFFFF], dtype=np.uint32)


def _key_tensor(key, device):
    if isinstance(key, torch.Tensor):
        t = key.to(device=device)
        return t.view(torch.int32) if t.dtype != torch.int32 else t.clone()
    k = np.asarray(key)
    if k.shape == ():
        k = PRNGKey(int(k))
    k = np.ascontiguousarray(k.astype(np.uint32)).view(np.int32)
    return torch.from_numpy(k.copy()).to(device)


def _radius(fn):
    if fn is None:
        return 2.0
    if isinstance(fn, partial):
        return float(fn.keywords.get("radius", fn.args[0] if fn.args else 2.0))
    if isinstance(fn, init_to_uniform):
        return fn.radius
    if fn is init_to_uniform:
        return 2.0
    raise NotImplementedError("only init_to_uniform(radius) initialisation is on the reference's path")


def _host_array(arr):
    if isinstance(arr, torch.Tensor):
        arr = arr.detach().numpy()
    return arr if (type(arr) is np.ndarray and arr.flags.c_contiguous) else np.ascontiguousarray(arr)


class _Stage:
    """Pinned staging pair (libsmi) + the device tensors a group of host arguments lands in: the
    arrays are packed back to back (256-byte aligned) so one H2D copy moves the whole group."""

    def __init__(self, arrays, dtypes, device):
        self.sig = tuple((a.shape, a.dtype) for a in arrays)
        n = len(arrays)
        offs, total = [], 0
        for a in arrays:
            total = (total + 255) // 256 * 256
            offs.append(total)
            total += int(a.nbytes)
        self.total = max(total, 1)
        self.nbytes = sum(int(a.nbytes) for a in arrays)
        h = C.c_void_p()
        _lib.call("smi_stage_create", self.total, C.byref(h))
        self.handle = h
        self.buf = torch.empty(self.total, dtype=torch.uint8, device=device)
        self.buf_ptr = self.buf.data_ptr()
        self.srcs = (C.c_void_p * n)()
        self.sizes = (C.c_size_t * n)(*[int(a.nbytes) for a in arrays])
        self.n = n
        self.raw, self.dev = [], []
        for a, off, dt in zip(arrays, offs, dtypes):
            src_dtype = torch.from_numpy(np.empty(0, dtype=a.dtype)).dtype
            raw = self.buf[off:off + a.nbytes].view(src_dtype).reshape(a.shape)
            self.raw.append(raw)
            self.dev.append(raw if src_dtype == dt else torch.empty(a.shape, dtype=dt, device=device))

    def __del__(self):
        try:
            _lib.load().smi_stage_destroy(self.handle)
        except Exception:
            pass


class SteinVI:
    """numpyro.contrib.einstein.SteinVI for the reference's models (see models.py)."""

    def __init__(self, model, guide, optim, kernel_fn, num_stein_particles=10, num_elbo_particles=10,
                 loss_temperature=1.0, repulsion_temperature=1.0, non_mixture_guide_params_fn=None,
                 device=None, use_cuda_graph=True, c_att=1.0, **static_kwargs):
        if not isinstance(model, models.Model):
            raise TypeError("model must be one of the descriptors in smi_experiments_b200.models "
                            "(NumPyro callables cannot be traced without NumPyro)")
        if not isinstance(kernel_fn, RBFKernel):
            raise TypeError("kernel_fn must be RBFKernel()")
        self.model = model
        self.guide = guide
        self.optim = optim
        self.kernel_fn = kernel_fn
        self.num_stein_particles = int(num_stein_particles)
        self.num_elbo_particles = int(num_elbo_particles)
        self.loss_temperature = float(loss_temperature)
        self.repulsion_temperature = float(repulsion_temperature)
        self.static_kwargs = static_kwargs
        self.c_att = float(c_att)
        self.device = torch.device(device if device is not None else "cuda")
        self.use_cuda_graph = bool(use_cuda_graph)
        self.md, self.layout = models.make_desc(model, guide.kind)
        self.guide_sites = [n for n in self.layout.names]
        self._ws_elbo, self._ws_stein = ops.Workspace(), ops.Workspace()
        self._bufs = None
        self._graph = None
        self._graph_sig = None
        self.launches_per_update = 0

    # ---- data binding: the reference's call-site kwargs -> device tensors -------------------
    def _to_dev(self, names, arrs, dtypes):
        """Host arrays are staged through pinned memory and copied H2D on the current stream every
        call, the whole group in one copy (smi_stage_h2dv: two pinned buffers used alternately, each
        guarded by the event of its last copy, so a host that runs ahead of the stream cannot
        overwrite a buffer whose copy has not executed); CUDA tensors are used in place."""
        out = [None] * len(arrs)
        host = []
        for i, (arr, dt) in enumerate(zip(arrs, dtypes)):
            if isinstance(arr, torch.Tensor) and arr.is_cuda:
                out[i] = (arr if arr.dtype == dt else arr.to(dt)).contiguous()
            else:
                host.append(i)
        if host:
            hs = [_host_array(arrs[i]) for i in host]
            key = tuple(names[i] for i in host)
            slot = self._stage.get(key)
            if slot is None or slot.sig != tuple((a.shape, a.dtype) for a in hs):
                slot = _Stage(hs, [dtypes[i] for i in host], self.device)
                self._stage[key] = slot
            for j, a in enumerate(hs):
                slot.srcs[j] = a.ctypes.data
            _lib.call("smi_stage_h2dv", slot.handle, slot.srcs, slot.sizes, slot.n, slot.buf_ptr, ops._stream())
            for j, i in enumerate(host):
                if slot.raw[j] is not slot.dev[j]:      # dtype conversion on the device (uint8 pixels -> fp32, int64 labels -> int32)
                    r, dv = slot.raw[j], slot.dev[j]
                    if r.dtype == torch.uint8 and dv.dtype == torch.float32:
                        _lib.call("smi_convert_u8_f32", ops._ptr(r), r.numel(), ops._ptr(dv), ops._stream())
                    elif r.dtype == torch.int64 and dv.dtype == torch.int32:
                        _lib.call("smi_convert_i64_i32", ops._ptr(r), r.numel(), ops._ptr(dv), ops._stream())
                    else:
                        dv.copy_(r)
                out[i] = slot.dev[j]
            self.h2d_bytes += slot.nbytes
        return out

    def _view(self, name, t, fn):
        """fn(t), cached while the same device tensor keeps coming back (the staged one does)."""
        c = self._views.get(name)
        if c is None or c[0] is not t:
            c = (t, fn(t))
            self._views[name] = c
        return c[1]

    def _bind(self, args, kwargs):
        kw = dict(self.static_kwargs)
        kw.update(kwargs)
        m = self.model
        self.h2d_bytes = 0
        if not hasattr(self, "_stage"):
            self._stage, self._views = {}, {}
        if isinstance(m, models.GaussianTarget):
            return dict(x=None, y=None, n_total=0, B=0, lik_scale=1.0, idx=False)
        x = args[0] if len(args) > 0 else kw["x"]
        y = args[1] if len(args) > 1 else kw["y"]
        if isinstance(m, models.BNNRegression):
            xd, yd = self._to_dev(("x", "y"), (x, y), (torch.float32, torch.float32))
            xd = self._view("x", xd, lambda t: t[:, None] if t.dim() == 1 else t)
            yd = self._view("y", yd, lambda t: t.reshape(-1))
            n = xd.shape[0]
            sub = kw.get("subsample", None)
            if sub is None or sub >= n:
                return dict(x=xd, y=yd, n_total=0, B=n, lik_scale=1.0, idx=False)
            return dict(x=xd, y=yd, n_total=n, B=int(sub), lik_scale=float(n) / float(sub), idx=True)
        if isinstance(m, models.BNNClassifier):
            xd, yd = self._to_dev(("x", "y"), (x, y), (torch.float32, torch.int32))
            xd = self._view("x", xd, lambda t: t.reshape(t.shape[0], -1))
            yd = self._view("y", yd, lambda t: t.reshape(-1))
            b = xd.shape[0]
            n = kw.get("n", b)
            bs = kw.get("batch_size", None)
            if not ((bs is None and b == n) or bs == b):
                raise ValueError("batch_size must equal the leading dimension of x (src/models/mlp.py:16)")
            plate = 0 if (bs is None or bs == n) else int(n)
            return dict(x=xd, y=yd, n_total=plate, B=b, lik_scale=float(n) / float(b), idx=False)
        raise TypeError(type(m))

    # ---- buffers ---------------------------------------------------------------------------
    def _alloc(self, data):
        m, S, L, d = self.num_stein_particles, self.num_elbo_particles, self.layout.latent_dim, self.layout.particle_dim
        dev = self.device
        normal = self.guide.kind == _lib.GUIDE_NORMAL
        sig = (data["B"], data["n_total"], data["idx"])
        if self._bufs is not None and self._bufs["sig"] == sig:
            return self._bufs
        b = dict(sig=sig)
        b["eps"] = torch.empty(m, S, L, dtype=torch.float32, device=dev) if normal else None
        b["idx"] = torch.empty(m, S, data["B"], dtype=torch.int32, device=dev) if data["idx"] else None
        b["G"] = torch.empty(m, d, dtype=torch.float32, device=dev)
        b["F"] = torch.empty(m, d, dtype=torch.float32, device=dev)
        b["elbo"] = torch.empty(m, dtype=torch.float32, device=dev)
        b["loss"] = torch.zeros(1, dtype=torch.float32, device=dev)
        b["h"] = torch.zeros(1, dtype=torch.float32, device=dev)
        b["key_next"] = torch.zeros(2, dtype=torch.int32, device=dev)
        self._bufs = b
        self._graph = None
        return b

    def _own(self, state):
        """Make `state` the engine's resident state (copy if it lives elsewhere)."""
        if getattr(self, "_state", None) is not None and state.optim_state.x is self._state.optim_state.x:
            return self._state
        os_ = state.optim_state
        dev = self.device
        self._state = SteinVIState(
            OptimState(os_.step.to(dev).clone(), os_.x.to(dev).clone().contiguous(),
                       os_.a.to(dev).clone(), os_.b.to(dev).clone()),
            _key_tensor(state.rng_key, dev))
        self._graph = None
        self._host_step = int(self._state.optim_state.step.item())
        return self._state

    # ---- the step --------------------------------------------------------------------------
    def _tau(self):
        return self.loss_temperature

    def _launch(self, st, b, data, do_update, ring=None):
        """Enqueue one step on the current stream.  Returns the number of kernel launches."""
        m, S = self.num_stein_particles, self.num_elbo_particles
        os_ = st.optim_state
        ops.prng_step(self.md, st.rng_key, b["key_next"], m, S, data["n_total"], data["B"] if data["idx"] else 0,
                      0, m, b["eps"], b["idx"])
        ops.elbo_grad(self.md, os_.x, b["eps"], b["idx"], data["x"], data["y"], S, max(data["B"], 1),
                      data["lik_scale"], self._tau(), 0, m, b["G"], b["elbo"], self._ws_elbo)
        nbytes = _lib.load().smi_stein_force_workspace_bytes(m, self.layout.particle_dim, _lib.PATH_AUTO)
        w = self._ws_stein.get(nbytes, self.device)
        _lib.call("smi_stein_force_f32", ops._ptr(os_.x), ops._ptr(b["G"]), m, self.layout.particle_dim,
                  self.c_att, self.repulsion_temperature, self.kernel_fn.median_mode, self.kernel_fn.h_fixed,
                  _lib.PATH_AUTO, ops._ptr(b["F"]), ops._ptr(b["loss"]), ops._ptr(b["h"]), ops._ptr(w),
                  w.numel(), ops._stream())
        if do_update:
            self._commit(st, b, ring)

    def _commit(self, st, b, ring):
        """Optimizer + end-of-step bookkeeping, all on the device: gated by the early-stop flag
        when an OnlineEarlyStop run is active (early_stop.py), one commit kernel for the step
        counter, the loss ring, the PRNG key and the stopping rule."""
        es = getattr(self, "_es", None)
        skip = es["skip"] if es is not None else None
        self.optim.update_gated(b["F"], st.optim_state, -1.0, skip)
        if es is None:
            ops.step_commit(b["loss"], st.optim_state.step, ring, b["key_next"], st.rng_key)
        else:
            ops.step_commit(b["loss"], st.optim_state.step, ring, b["key_next"], st.rng_key, es["avg"], skip,
                            es["fast_decay"], es["slow_decay"], es["period"])

    def init(self, rng_key, *args, **kwargs):
        """SteinVI.init (src/engine_run.py:41-47): particles ~ init_to_uniform(radius) per site in
        unconstrained space, AutoNormal scales = init_scale; returns SteinVIState."""
        ops.device_check()
        dev = self.device
        m, L, d = self.num_stein_particles, self.layout.latent_dim, self.layout.particle_dim
        key = _key_tensor(rng_key, dev)
        keys5 = torch.empty(5, 2, dtype=torch.int32, device=dev)
        _lib.call("smi_prng_split", ops._ptr(key), 5, ops._ptr(keys5), ops._stream())
        pkeys = torch.empty(m, 2, dtype=torch.int32, device=dev)
        _lib.call("smi_prng_split", ops._ptr(keys5[4]), m, ops._ptr(pkeys), ops._stream())
        radius = _radius(self.guide.init_loc_fn)
        lat = torch.empty(m, L, dtype=torch.float32, device=dev)
        for i in range(m):
            _lib.call("smi_prng_uniform_f32", ops._ptr(pkeys[i]), L, -radius, radius, ops._ptr(lat[i]),
                      ops._stream())
        psi = torch.empty(m, d, dtype=torch.float32, device=dev)
        for r in self.layout.rows:
            psi[:, r["loc"]:r["loc"] + r["size"]] = lat[:, r["lat"]:r["lat"] + r["size"]]
            if r["rho"] >= 0:
                psi[:, r["rho"]:r["rho"] + r["size"]] = models.softplus_inv(self.guide.init_scale)
        self._state = SteinVIState(self.optim.init(psi), keys5[0].clone())
        self._graph = None
        self._host_step = 0
        return self._state

    def get_params(self, state):
        """{name: constrained value (m, *site_shape)} with NumPyro's AutoGuide names."""
        x = state.optim_state.x
        out = {}
        for name in self.layout.names:
            off, shape, site, which = self.layout.leaves[name]
            v = x[:, off:off + site.size].reshape((x.shape[0],) + tuple(shape))
            if which == "rho":
                v = torch.nn.functional.softplus(v)
            elif self.guide.kind == _lib.GUIDE_DELTA and site.transform == _lib.T_EXP:
                v = torch.exp(v)
            out[name] = v
        return out

    def evaluate(self, state, *args, **kwargs):
        """Loss of `state` without updating it (same key as the next update; src/engine_run.py:48)."""
        st = self._own(state)
        data = self._bind(args, kwargs)
        b = self._alloc(data)
        self._launch(st, b, data, do_update=False)
        return b["loss"].clone().reshape(())

    def update(self, state, *args, **kwargs):
        """One Stein step (src/engine_run.py:25,30).  Returns (state, loss); loss is a device scalar."""
        st = self._own(state)
        data = self._bind(args, kwargs)
        b = self._alloc(data)
        ring = getattr(self, "_ring", None)
        if ring is None:
            # loss history on the device: the commit kernel stores every step's loss in its own slot, so the
            # returned scalar is not overwritten by later steps and costs no extra launch
            if getattr(self, "_hist", None) is None:
                self._hist = torch.zeros(self.LOSS_HISTORY, dtype=torch.float32, device=self.device)
            ring = self._hist
        if not self.use_cuda_graph or not self._graphable():
            self._eager(st, b, data, ring)
            return st, self._loss_out(b, ring)
        # everything the captured launches bake in: buffer addresses AND the host scalars / process-wide PRNG
        # switches (the reference assigns engine.loss_temperature per step, src/engine_run.py:24: a changed
        # scalar re-captures instead of silently replaying the old value)
        sig = (data["x"].data_ptr() if data["x"] is not None else 0,
               data["y"].data_ptr() if data["y"] is not None else 0, data["lik_scale"],
               ring.data_ptr() if ring is not None else 0,
               self._es["avg"].data_ptr() if getattr(self, "_es", None) is not None else 0,
               self.loss_temperature, self.repulsion_temperature, self.c_att, self.kernel_fn.median_mode,
               self.kernel_fn.h_fixed, ops.prng_generation())
        if self._graph is None or self._graph_sig != sig:
            if getattr(self, "_warm_sig", None) != sig:
                # first call with these buffers: run eagerly (sets kernel attributes), capture next time
                self._warm_sig = sig
                self._eager(st, b, data, ring)
                return st, self._loss_out(b, ring)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch(st, b, data, True, ring)
            self._graph, self._graph_sig = g, sig
            # capture does not execute: fall through to the replay for this step
        self._graph.replay()
        return st, self._loss_out(b, ring)

    LOSS_HISTORY = 1 << 20     # slots of the default loss ring (4 MB); a returned loss lives until the ring wraps

    def _loss_out(self, b, ring):
        """The step's loss as its OWN device scalar: `losses.append(loss)` keeps every value.  The commit kernel
        wrote it to ring[t mod len] with t = the step counter before the update, which the host mirrors
        (`_host_step`); with an early-stop flag the device counter may freeze, so that path clones instead."""
        t = self._host_step
        self._host_step = t + 1
        if getattr(self, "_es", None) is not None:
            return b["loss"].clone().reshape(())
        return ring[t % ring.numel()]

    def _graphable(self):
        return True

    def _eager(self, st, b, data, ring):
        n0 = _lib.load().smi_launch_count(0)
        self._launch(st, b, data, True, ring)
        self.launches_per_update = _lib.load().smi_launch_count(0) - n0

    def setup_run(self, rng_key, num_steps, args, init_state, kwargs):
        """(step, diagnostic, collect, extract_state, init_info) as SteinVI.setup_run
        (src/model_select/early_stopping.py:16-22)."""
        state = self.init(rng_key, *args, **kwargs) if init_state is None else self._own(init_state)
        loss0 = self.evaluate(state, *args, **kwargs)

        def step(info):
            st, _ = info
            return self.update(st, *args, **kwargs)

        def collect(info):
            return info[1]

        def extract_state(info):
            return info[0]

        def diagnostic(info):
            return "Stein force {:.2f}.".format(float(info[1]))

        return step, diagnostic, collect, extract_state, (state, loss0)

    def run(self, rng_key, num_steps, *args, progress_bar=True, init_state=None, **kwargs):
        """SteinVI.run: num_steps updates, losses collected ON THE DEVICE (one D2H at the end)."""
        if num_steps < 1:
            raise ValueError("num_steps must be a positive integer.")
        state = self.init(rng_key, *args, **kwargs) if init_state is None else self._own(init_state)
        start = int(state.optim_state.step.item())
        # the step kernel sequence ends with ring[(step - 1) mod len] = loss (smi_loss_record_f32)
        self._ring = torch.zeros(num_steps, dtype=torch.float32, device=self.device)
        try:
            for _ in range(num_steps):
                state, _loss = self.update(state, *args, **kwargs)
            k = start % num_steps
            losses = torch.roll(self._ring, -k) if k else self._ring
        finally:
            self._ring = None
            self._graph = None
        return SteinVIRunResult(self.get_params(state), state, losses)


class SVGD(SteinVI):
    """numpyro.contrib.einstein.SVGD with the reference wrapper's extra `repulsion_temperature`
    (src/methods/svgd.py:4-24).  AutoDelta guide, one ELBO draw, loss_temperature = 1/m.
    compat_b1=True reproduces the wrapper's bug B1 (10 particles and default init regardless of
    the arguments)."""

    def __init__(self, model, optim, kernel_fn, num_stein_particles=10, repulsion_temperature=1.0,
                 guide_kwargs={}, compat_b1=False, **static_kwargs):
        if compat_b1:
            num_stein_particles, guide_kwargs = 10, {}
        super().__init__(model, AutoDelta(model, **guide_kwargs), optim, kernel_fn,
                         num_stein_particles=num_stein_particles, num_elbo_particles=1,
                         loss_temperature=1.0 / float(num_stein_particles),
                         repulsion_temperature=repulsion_temperature, **static_kwargs)


class ASVGD(SVGD):
    """numpyro.contrib.einstein.ASVGD (src/methods/asvgd.py:4-27): SVGD whose loss temperature
    follows a cyclical annealing schedule, tau_t = cyc(t) / m."""

    def __init__(self, model, optim, kernel_fn, num_stein_particles=10, num_cycles=10, trans_speed=10,
                 repulsion_temperature=1.0, guide_kwargs={}, **static_kwargs):
        super().__init__(model, optim, kernel_fn, num_stein_particles, repulsion_temperature, guide_kwargs,
                         **static_kwargs)
        self.num_cycles = int(num_cycles)
        self.trans_speed = int(trans_speed)
        self._sched = None

    @staticmethod
    def _cyclical_annealing(num_steps, num_cycles, trans_speed):
        """Schedule of numpyro ASVGD; mirrored at src/engine_run.py:16-18."""
        norm = float(num_steps + 1) / float(num_cycles)
        cycle_len = num_steps // num_cycles
        last_start = (num_cycles - 1) * cycle_len

        def cycle_fn(t):
            last = min(t // last_start, 1) if last_start > 0 else 1
            return (1 - last) * ((((t - 1) % cycle_len) + 1) / norm) ** trans_speed + last

        return cycle_fn

    # The schedule runs on the device (smi_asvgd_tau_f32 reads the step counter), so the annealed
    # step is CUDA-graph replayable like every other engine: no host scalar changes between steps.
    def _begin_schedule(self, num_steps):
        self._sched = dict(T=int(num_steps), step0=None)
        self._tau_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._graph = self._warm_sig = None

    def _end_schedule(self):
        self._sched = None
        self._graph = self._warm_sig = None

    def _tau(self):
        sched = getattr(self, "_sched", None)
        if sched is None:
            return self.loss_temperature
        step = self._state.optim_state.step
        if sched["step0"] is None:
            sched["step0"] = int(step.item())
        ops.asvgd_tau(step, sched["step0"], sched["T"], self.num_cycles, self.trans_speed,
                      self.num_stein_particles, self._tau_dev)
        return self._tau_dev

    def run(self, rng_key, num_steps, *args, **kwargs):
        self._begin_schedule(num_steps)
        try:
            return super().run(rng_key, num_steps, *args, **kwargs)
        finally:
            self._end_schedule()
