"""CPU pins of the OnlineEarlyStop restatement (oracle/early_stop.py) and of the host-side
constants the product passes to the device (reference src/model_select/early_stopping.py:24-62)."""
import numpy as np

from oracle import early_stop as oe


def test_decay_constants_match_definition():
    for period in (1, 2, 35, 100):
        fast, slow = oe.decays(period)
        assert fast.dtype == np.float32 and slow.dtype == np.float32
        # decay_target = fast_decay ** period = slow_decay ** (10 period)   (early_stopping.py:26-31)
        assert abs(float(fast) ** period - 1e-3) < 1e-6
        assert abs(float(slow) ** (10 * period) - 1e-3) < 2e-6
    # the product computes the same two floats without importing the oracle
    from smi_experiments_b200 import early_stop as pe
    for period in (1, 7, 35):
        f, s = oe.decays(period)
        assert pe.decays(period) == (float(f), float(s))


def test_first_step_and_constant_loss():
    es = oe.EarlyStop(35)
    assert es.step(3.5) is False
    # t = 0: tau = 0, both averages equal the loss itself
    assert es.foa == np.float32(3.5) and es.soa == np.float32(3.5) and es.t == np.float32(1.0)
    # a constant loss never stops with the reference's period (fast/slow -> 0.1376 / 0.1276 of the loss)
    n, es = oe.stop_index([1.0] * 3000, 35)
    assert n == 3000 and not es.skip
    assert abs(float(es.foa) - 0.13757) < 1e-4 and abs(float(es.soa) - 0.12762) < 1e-4


def test_rule_fires_and_freezes():
    # period 1: fast -> loss / 2, slow -> 0.167 loss: fires once the slow average forgot its start
    n, es = oe.stop_index([2.0] * 50, 1)
    assert 5 <= n <= 10 and es.skip
    t, foa, soa = es.t, es.foa, es.soa
    assert es.step(1e9) is True                      # identity after the stop
    assert (es.t, es.foa, es.soa) == (t, foa, soa)
    # a blow-up of the loss after a long quiet phase stops the run within a few steps
    losses = [1.0] * 2000 + [1.0 * 1.5 ** k for k in range(1, 40)]
    n, es = oe.stop_index(losses, 35)
    assert 2000 < n < 2012 and es.skip


def test_matches_float64_recurrence_loosely():
    rng = np.random.default_rng(0)
    losses = np.abs(rng.standard_normal(400)) + 1.0
    es = oe.EarlyStop(35)
    fd, sd = (float(v) for v in oe.decays(35))
    foa = soa = 0.0
    for t, l in enumerate(losses):
        es.step(l)
        tf, ts = min(t, 35), min(t, 350)
        foa = l / (tf + 1) + tf / (tf + 1) * fd * foa
        soa = l / (ts + 1) + ts / (ts + 1) * sd * soa
    assert abs(float(es.foa) - foa) < 1e-5 * foa and abs(float(es.soa) - soa) < 1e-5 * soa
