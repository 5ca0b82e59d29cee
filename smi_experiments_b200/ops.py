"""Stage-level operators: torch CUDA tensors in, libsmi (hand-written sm_100a CUDA) underneath.

torch is the container for device memory and streams only; every function here is one
C-ABI call (include/smi.h).  Nothing falls back to torch math or to the CPU.
"""

import ctypes as C

import torch

from . import _lib
from ._lib import (MEDIAN_EXACT_MIDPOINT, MEDIAN_JAX_FP32_INDEX, PATH_AUTO, PATH_SMALL,  # noqa: F401
                   PATH_TC, call, load)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("libsmi operators take CUDA tensors (there is no CPU path)")
        if t is not None and not t.is_contiguous():
            raise ValueError("libsmi operators take contiguous tensors")


class Workspace:
    """Grow-only device scratch buffer handed to the C ABI (the caller owns all memory)."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        nbytes = int(nbytes)
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        return self.buf


_default_ws = {}


def _ws(tag, nbytes, device):
    w = _default_ws.setdefault((tag, device), Workspace())
    return w.get(nbytes, device)


def device_check():
    call("smi_device_check")


def stein_force(X, G, alpha=1.0, c_att=1.0, median_mode=MEDIAN_JAX_FP32_INDEX, h_fixed=0.0,
                path=PATH_AUTO, out=None, ws=None):
    """F, loss, h = Stein force of particles X (m, d) with scores G (m, d).  rows a4-a7."""
    _need_cuda(X, G)
    m, d = X.shape
    F = out if out is not None else torch.empty_like(X)
    loss = torch.empty(1, dtype=torch.float32, device=X.device)
    h = torch.empty(1, dtype=torch.float32, device=X.device)
    nbytes = load().smi_stein_force_workspace_bytes(m, d, path)
    w = ws.get(nbytes, X.device) if ws is not None else _ws("stein", nbytes, X.device)
    call("smi_stein_force_f32", _ptr(X), _ptr(G), m, d, c_att, alpha, median_mode, h_fixed, path,
         _ptr(F), _ptr(loss), _ptr(h), _ptr(w), w.numel(), _stream())
    return F, loss, h


def sqdist(X, path=PATH_AUTO):
    """Gram-form squared distances D2 (m, m).  row a4."""
    _need_cuda(X)
    m, d = X.shape
    D2 = torch.empty(m, m, dtype=torch.float32, device=X.device)
    nbytes = load().smi_sqdist_workspace_bytes(m, d, path)
    w = _ws("stein", nbytes, X.device)
    call("smi_sqdist_f32", _ptr(X), m, d, path, _ptr(D2), _ptr(w), w.numel(), _stream())
    return D2


def median_bandwidth(D2, median_mode=MEDIAN_JAX_FP32_INDEX):
    """h, ranks: exact order-statistic bandwidth of a squared-distance matrix.  row a4 (K2).
    ranks = uint32 bit patterns of the two selected D2 values (as int64 tensor of length 2)."""
    _need_cuda(D2)
    m = D2.shape[0]
    h = torch.empty(1, dtype=torch.float32, device=D2.device)
    ranks = torch.zeros(2, dtype=torch.int32, device=D2.device)
    nbytes = load().smi_median_bandwidth_workspace_bytes(m)
    w = _ws("select", nbytes, D2.device)
    call("smi_median_bandwidth_f32", _ptr(D2), m, median_mode, _ptr(h), _ptr(ranks), _ptr(w), w.numel(),
         _stream())
    return h, ranks


def adam(x, mom, vel, F, step, lr, grad_sign=-1.0, b1=0.9, b2=0.999, eps=1e-8, bump_step=True):
    _need_cuda(x, mom, vel, F, step)
    call("smi_adam_f32", _ptr(x), _ptr(mom), _ptr(vel), _ptr(F), grad_sign, _ptr(step), int(bump_step),
         lr, b1, b2, eps, x.numel(), _stream())


def adagrad(x, acc, mom, F, step, lr, grad_sign=-1.0, momentum=0.9, bump_step=True):
    _need_cuda(x, acc, mom, F, step)
    call("smi_adagrad_f32", _ptr(x), _ptr(acc), _ptr(mom), _ptr(F), grad_sign, _ptr(step), int(bump_step),
         lr, momentum, x.numel(), _stream())


def adam_gated(x, mom, vel, F, step, lr, grad_sign=-1.0, b1=0.9, b2=0.999, eps=1e-8, skip=None):
    """Adam without the step bump, skipped on the device when *skip != 0 (run loops, row n1)."""
    _need_cuda(x, mom, vel, F, step, skip)
    call("smi_adam_gated_f32", _ptr(x), _ptr(mom), _ptr(vel), _ptr(F), grad_sign, _ptr(step), lr, b1, b2, eps,
         x.numel(), _ptr(skip), _stream())


def adagrad_gated(x, acc, mom, F, lr, grad_sign=-1.0, momentum=0.9, skip=None):
    _need_cuda(x, acc, mom, F, skip)
    call("smi_adagrad_gated_f32", _ptr(x), _ptr(acc), _ptr(mom), _ptr(F), grad_sign, lr, momentum, x.numel(),
         _ptr(skip), _stream())


def step_commit(loss, step, ring=None, key_next=None, key=None, es=None, skip=None, fast_decay=0.0,
                slow_decay=0.0, period=1):
    """Close one update on the device: step += 1, ring[t % len] = loss, key <- key_next, and the
    OnlineEarlyStop averages / skip flag (src/model_select/early_stopping.py:24-62).  row n1."""
    _need_cuda(loss, step, ring, key_next, key, es, skip)
    call("smi_step_commit_f32", _ptr(loss), _ptr(step), _ptr(ring), 0 if ring is None else ring.numel(),
         _ptr(key_next), _ptr(key), _ptr(es), _ptr(skip), float(fast_decay), float(slow_decay), int(period),
         _stream())


def elbo_grad(md, psi, eps, idx, x, y, S, B, lik_scale=1.0, tau=1.0, row0=0, nrows=None, G=None,
              elbo=None, ws=None):
    """G (nrows, d), elbo (nrows): per-particle mixture-ELBO gradients.  row a3."""
    _need_cuda(psi, eps, idx, x, y)
    m, d = psi.shape
    nrows = m - row0 if nrows is None else nrows
    if G is None:
        G = torch.empty(nrows, d, dtype=torch.float32, device=psi.device)
    if elbo is None:
        elbo = torch.empty(nrows, dtype=torch.float32, device=psi.device)
    nbytes = load().smi_elbo_grad_workspace_bytes(C.byref(md), m, S, B, nrows)
    w = ws.get(nbytes, psi.device) if ws is not None else _ws("elbo", nbytes, psi.device)
    if isinstance(tau, torch.Tensor):      # device-side temperature (ASVGD schedule inside a CUDA graph)
        _need_cuda(tau)
        call("smi_elbo_grad_tau_f32", C.byref(md), _ptr(psi), _ptr(eps), _ptr(idx), _ptr(x), _ptr(y), m, S, B,
             lik_scale, _ptr(tau), row0, nrows, _ptr(G), _ptr(elbo), _ptr(w), w.numel(), _stream())
    else:
        call("smi_elbo_grad_f32", C.byref(md), _ptr(psi), _ptr(eps), _ptr(idx), _ptr(x), _ptr(y), m, S, B,
             lik_scale, tau, row0, nrows, _ptr(G), _ptr(elbo), _ptr(w), w.numel(), _stream())
    return G, elbo


def asvgd_tau(step, step0, num_steps, num_cycles, trans_speed, m, out):
    """out[0] = cyc(step - step0) / m on the device (ASVGD._cyclical_annealing).  row a10."""
    _need_cuda(step, out)
    call("smi_asvgd_tau_f32", _ptr(step), int(step0), int(num_steps), int(num_cycles), int(trans_speed), int(m),
         _ptr(out), _stream())


PRNG_LAYOUTS = {"partitionable": 1, "legacy": 0}   # SMI_PRNG_PARTITIONABLE / SMI_PRNG_LEGACY


_prng_gen = [0]


def prng_generation():
    """Bumped whenever a process-wide PRNG switch changes: part of the engines' CUDA-graph signature."""
    return _prng_gen[0]


def set_prng_layout(layout):
    """Threefry stream layout of every PRNG launch (jax_threefry_partitionable on / off; JAX >= 0.5 / older).
    Process-wide.  Returns the previous layout's name; ``None`` only queries."""
    code = load().smi_prng_layout(-1 if layout is None else PRNG_LAYOUTS[layout])
    if layout is not None:
        _prng_gen[0] += 1
    return "partitionable" if code == 1 else "legacy"


SUBSAMPLE_RULES = {"mulshift": 0, "jax": 1, "choice": 2}


def set_randint_rule(rule):
    """Rule of the data-plate subsample: partial Fisher-Yates with "mulshift" (one word, default) or "jax"
    (jax.random.randint's two-word construction) indices -- the reference's CPU-backend branch -- or "choice"
    (jax.random.choice without replacement = sort-based permutation, its branch on any other backend).
    Process-wide.  Returns the previous rule's name; ``None`` only queries."""
    code = load().smi_prng_randint(-1 if rule is None else SUBSAMPLE_RULES[rule])
    if rule is not None:
        _prng_gen[0] += 1
    return {v: k for k, v in SUBSAMPLE_RULES.items()}[code]


def prng_step(md, key_in, key_out, m, S, n_total=0, B=0, row0=0, nrows=None, eps=None, idx=None):
    """threefry key chain of one update: fills eps (nrows, S, L) / idx (nrows, S, B), next key.  row a9."""
    _need_cuda(key_in, key_out, eps, idx)
    nrows = m - row0 if nrows is None else nrows
    call("smi_prng_step_f32", C.byref(md), _ptr(key_in), _ptr(key_out), m, S, n_total, B, row0, nrows,
         _ptr(eps), _ptr(idx), _stream())
