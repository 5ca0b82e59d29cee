"""OnlineEarlyStop on the device.

Mirror of the reference's ``OnlineEarlyStop(engine, period)``
(src/model_select/early_stopping.py:10-81, used at experiments/uci_benchmark.py:228-236): a fast
and a slow exponentially-decayed average of the Stein-force loss (decay target 1e-3 over
``period`` and ``10 * period`` steps); the run stops changing state at the first step where
``fast > 2 * slow``.

In the reference the rule runs inside ``fori_collect`` as ``lax.cond(skip, skip_fn, step_fn)``:
all ``num_steps`` iterations execute, the ones after the stop are no-ops.  Here the averages
and the skip flag live in device memory and are updated by the step's commit kernel
(``smi_step_commit_f32``); the optimizer pass is gated by the same flag, so the particles, the
optimizer slots, the step counter and the PRNG key freeze at exactly the reference's step with
no host round trip per step.  The host polls the 4-byte flag once every ``check_every`` steps
and stops launching; the (at most ``check_every - 1``) steps launched after the stop are
no-ops on the state.
"""

import numpy as np
import torch

from .steinvi import SteinVIRunResult

DECAY_TARGET = 0.001  # src/model_select/early_stopping.py:27


def decays(period):
    """(fast_decay, slow_decay) in float32, as jnp.exp(jnp.log(decay_target) / period)."""
    lg = np.log(np.float32(DECAY_TARGET))
    fast = np.exp(lg / np.float32(period), dtype=np.float32)
    slow = np.exp(lg / np.float32(10 * period), dtype=np.float32)
    return float(fast), float(slow)


class OnlineEarlyStop:
    """``OnlineEarlyStop(engine, period).run(key, max_steps, *args, **kwargs)``."""

    def __init__(self, engine, period, check_every=64):
        if int(period) < 1:
            raise ValueError("period must be a positive integer")
        self.engine = engine
        self.period = int(period)
        self.check_every = max(int(check_every), 1)
        self.steps_run = None

    # attributes the reference reads off the wrapped engine (src/metrics.py:217-223)
    @property
    def model(self):
        return self.engine.model

    @property
    def guide(self):
        return self.engine.guide

    @property
    def guide_sites(self):
        return self.engine.guide_sites

    def get_params(self, state):
        return self.engine.get_params(state)

    def _arm(self):
        e = self.engine
        fast, slow = decays(self.period)
        e._es = dict(avg=torch.zeros(3, dtype=torch.float32, device=e.device),
                     skip=torch.zeros(1, dtype=torch.int32, device=e.device),
                     fast_decay=fast, slow_decay=slow, period=self.period)
        e._graph = None
        return e._es

    def _disarm(self):
        self.engine._es = None
        self.engine._graph = None

    def setup_run(self, rng_key, num_steps, args, init_state, kwargs):
        """(step, diagnostic, collect, extract_state, info_init) with the reference's info tuple
        (t, skip, fast_avg, slow_avg, inner_info) held on the device; the tuple handed around on
        the host is (es, inner_info).  Call ``close()`` on the returned step when done."""
        istep, idiag, icol, iext, iinit = self.engine.setup_run(rng_key, num_steps, args, init_state, kwargs)
        es = self._arm()

        def step(info):
            return (es, istep(info[1]))

        step.close = self._disarm

        def diagnostic(info):
            return idiag(info[1])

        def collect(info):
            return icol(info[1])

        def extract_state(info):
            return iext(info[1])

        return step, diagnostic, collect, extract_state, (es, iinit)

    def run(self, rng_key, num_steps, *args, progress_bar=False, init_state=None, **kwargs):
        """Up to ``num_steps`` updates; returns SteinVIRunResult(params, state, losses) with
        ``losses`` of length ``num_steps`` (entries after the stop repeat the last loss, as the
        reference's fori_collect does).  ``self.steps_run`` = updates actually applied."""
        if num_steps < 1:
            raise ValueError("num_steps must be a positive integer.")
        e = self.engine
        state = e.init(rng_key, *args, **kwargs) if init_state is None else e._own(init_state)
        start = int(state.optim_state.step.item())
        es = self._arm()
        e._ring = torch.zeros(num_steps, dtype=torch.float32, device=e.device)
        try:
            launched = 0
            while launched < num_steps:
                n = min(self.check_every, num_steps - launched)
                for _ in range(n):
                    state, _ = e.update(state, *args, **kwargs)
                launched += n
                if int(es["skip"].item()):          # the only host sync: 4 bytes per check_every steps
                    break
            done = int(state.optim_state.step.item()) - start
            ring = e._ring
            k = start % num_steps
            losses = torch.roll(ring, -k) if k else ring.clone()
            if 0 < done < num_steps:
                losses[done:] = losses[done - 1]
            self.steps_run = done
            e._host_step = start + done        # the device counter froze at the stop: re-sync the host mirror
            self.stopped_early = bool(int(es["skip"].item()))
        finally:
            e._ring = None
            self._disarm()
        return SteinVIRunResult(e.get_params(state), state, losses)
