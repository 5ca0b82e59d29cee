"""Device-side run loop: step commit kernel, gated optimizers and OnlineEarlyStop
against the CPU restatement of reference src/model_select/early_stopping.py:24-62."""
import numpy as np
import pytest
import torch

from oracle import early_stop as oe

pytestmark = pytest.mark.gpu


def test_step_commit_matches_oracle_bit_exact(cuda):
    from smi_experiments_b200 import early_stop as pe
    from smi_experiments_b200 import ops
    rng = np.random.default_rng(0)
    for period, losses in ((1, np.full(40, 2.0)), (3, np.abs(rng.standard_normal(300)) + 0.5),
                           (35, np.concatenate([np.ones(600), 1.5 ** np.arange(1, 40)]))):
        losses = losses.astype(np.float32)
        n = len(losses)
        fast, slow = pe.decays(period)
        loss = torch.zeros(1, device=cuda)
        step = torch.zeros(1, dtype=torch.int32, device=cuda)
        ring = torch.full((n,), -1.0, device=cuda)
        key = torch.zeros(2, dtype=torch.int32, device=cuda)
        key_next = torch.zeros(2, dtype=torch.int32, device=cuda)
        avg = torch.zeros(3, device=cuda)
        skip = torch.zeros(1, dtype=torch.int32, device=cuda)
        ld = torch.from_numpy(losses).to(cuda)
        for t in range(n):
            loss.copy_(ld[t:t + 1])
            key_next.fill_(t + 1)
            ops.step_commit(loss, step, ring, key_next, key, avg, skip, fast, slow, period)
        n_ref, es = oe.stop_index(losses, period)
        assert int(step.item()) == n_ref
        assert bool(skip.item()) == es.skip
        a = avg.cpu().numpy()
        assert a[0] == es.t and a[1] == es.foa and a[2] == es.soa          # bit-exact float32
        assert key.cpu().tolist() == [n_ref, n_ref]                        # key of the last applied step
        r = ring.cpu().numpy()
        assert np.array_equal(r[:n_ref], losses[:n_ref]) and (r[n_ref:] == -1.0).all()


def test_gated_optimizers_freeze_state(cuda):
    from smi_experiments_b200 import ops
    g = torch.Generator(device=cuda).manual_seed(0)
    x = torch.randn(7, 33, device=cuda, generator=g)
    F = torch.randn(7, 33, device=cuda, generator=g)
    step = torch.zeros(1, dtype=torch.int32, device=cuda)
    skip = torch.zeros(1, dtype=torch.int32, device=cuda)
    for fn in ("adam", "adagrad"):
        xa, a, b = x.clone(), torch.zeros_like(x), torch.zeros_like(x)
        xb, a2, b2 = x.clone(), torch.zeros_like(x), torch.zeros_like(x)
        if fn == "adam":
            ops.adam(xa, a, b, F, step.clone(), 1e-2, bump_step=False)
            ops.adam_gated(xb, a2, b2, F, step, 1e-2, skip=skip)
        else:
            ops.adagrad(xa, a, b, F, step.clone(), 1e-2, bump_step=False)
            ops.adagrad_gated(xb, a2, b2, F, 1e-2, skip=skip)
        assert torch.equal(xa, xb) and torch.equal(a, a2) and torch.equal(b, b2)   # same arithmetic
        skip.fill_(1)
        snap = (xb.clone(), a2.clone(), b2.clone())
        if fn == "adam":
            ops.adam_gated(xb, a2, b2, F, step, 1e-2, skip=skip)
        else:
            ops.adagrad_gated(xb, a2, b2, F, 1e-2, skip=skip)
        assert torch.equal(xb, snap[0]) and torch.equal(a2, snap[1]) and torch.equal(b2, snap[2])
        skip.zero_()
    assert int(step.item()) == 0      # gated forms never bump the counter


def _engine(smi, graph):
    model = smi.uci_bnn(6)
    return smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                       smi.RBFKernel(), num_stein_particles=4, num_elbo_particles=6, use_cuda_graph=graph)


@pytest.mark.parametrize("graph", [False, True])
def test_online_early_stop_run_matches_stepwise_oracle(cuda, graph):
    """The device-side stop freezes the run at the step the oracle's rule picks from the same losses,
    with particles / optimizer slots / key bit-identical to an ungated run cut at that step."""
    import smi_experiments_b200 as smi
    rng = np.random.default_rng(3)
    x = rng.standard_normal((50, 6)).astype(np.float32)
    y = rng.standard_normal(50).astype(np.float32)
    max_steps, period = 40, 1
    # ungated reference trajectory, one D2H per step
    eng = _engine(smi, False)
    st = eng.init(smi.PRNGKey(7), x=x, y=y, subsample=10)
    losses, snaps = [], []
    for _ in range(max_steps):
        st, l = eng.update(st, x=x, y=y, subsample=10)
        losses.append(float(l))
        snaps.append((st.optim_state.x.clone(), st.optim_state.a.clone(), st.rng_key.clone()))
    n_ref, es = oe.stop_index(losses, period)
    assert es.skip and n_ref < max_steps, "the test needs a trajectory on which the rule fires"
    # device-side early stop, host polls every 16 steps (so 16 - n_ref % 16 gated no-op steps run)
    eng2 = _engine(smi, graph)
    stopper = smi.OnlineEarlyStop(eng2, period, check_every=16)
    res = stopper.run(smi.PRNGKey(7), max_steps, x=x, y=y, subsample=10, progress_bar=False)
    assert stopper.steps_run == n_ref and stopper.stopped_early
    assert int(res.state.optim_state.step.item()) == n_ref
    xs, as_, ks = snaps[n_ref - 1]
    assert torch.equal(res.state.optim_state.x, xs) and torch.equal(res.state.optim_state.a, as_)
    assert torch.equal(res.state.rng_key, ks)
    lg = res.losses.cpu().numpy()
    assert lg.shape == (max_steps,)
    assert np.array_equal(lg[:n_ref], np.asarray(losses[:n_ref], np.float32))
    assert (lg[n_ref:] == lg[n_ref - 1]).all()
    assert set(res.params) == set(eng.get_params(st))
    # the engine is left ungated
    assert eng2._es is None


def test_online_early_stop_setup_run_protocol(cuda):
    import smi_experiments_b200 as smi
    model = smi.GaussianTarget(8)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=6)
    stopper = smi.OnlineEarlyStop(eng, 1)
    step, diagnostic, collect, extract_state, info = stopper.setup_run(smi.PRNGKey(0), 30, (), None, {})
    losses = []
    for _ in range(30):
        info = step(info)
        losses.append(float(collect(info)))
    n_ref, es = oe.stop_index(losses, 1)
    assert int(extract_state(info).optim_state.step.item()) == n_ref
    assert isinstance(diagnostic(info), str)
    step.close()
    assert eng._es is None
