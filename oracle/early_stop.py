"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's OnlineEarlyStop rule.

Follows reference src/model_select/early_stopping.py:24-62 statement by statement in float32
(JAX default dtype; t is the float carried in `info_init = (0.0, False, 0.0, 0.0, iinit)`):

    fast_decay = exp(log(0.001) / period);  slow_decay = exp(log(0.001) / (10 * period))
    per executed step (loss = Stein-force loss of that update):
        tau_f = min(t, period);       foa = loss / (tau_f + 1) + (tau_f / (tau_f + 1)) * fast_decay * foa
        tau_s = min(t, 10 * period);  soa = loss / (tau_s + 1) + (tau_s / (tau_s + 1)) * slow_decay * soa
        skip  = skip or foa > 2 * soa;   t = t + 1
    once skip is set the step is the identity (lax.cond(skip, skip_fn, step_fn)).

Parity unpinned: the reference has no test or fixture for this rule; pinned
here by closed forms (constant loss never stops; first step has foa == soa == loss).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import numpy as np

F = np.float32


def decays(period):
    lg = np.log(F(0.001))
    return np.exp(lg / F(period)), np.exp(lg / F(10 * period))


class EarlyStop:
    def __init__(self, period):
        self.period = int(period)
        self.fast_decay, self.slow_decay = decays(period)
        self.t, self.skip, self.foa, self.soa = F(0.0), False, F(0.0), F(0.0)

    def step(self, loss):
        """Feed the loss of one executed update; returns the skip flag after it."""
        if self.skip:
            return True
        loss = F(loss)
        tf = min(self.t, F(self.period))
        ts = min(self.t, F(10 * self.period))
        one = F(1.0)
        self.foa = F(F(loss / F(tf + one)) + F(F(F(tf / F(tf + one)) * self.fast_decay) * self.foa))
        self.soa = F(F(loss / F(ts + one)) + F(F(F(ts / F(ts + one)) * self.slow_decay) * self.soa))
        self.skip = bool(self.foa > F(F(2.0) * self.soa))
        self.t = F(self.t + one)
        return self.skip


def stop_index(losses, period):
    """Number of updates applied when the rule is run over `losses` (len(losses) if it never fires)."""
    es = EarlyStop(period)
    for i, l in enumerate(losses):
        if es.step(l):
            return i + 1, es
    return len(losses), es
