"""Full SteinVI.update / run through the engine API against the oracle step (rows a1, a2, a7, a10)."""
import numpy as np
import pytest
import torch

from oracle import models as omodels
from oracle import optim as ooptim
from oracle import prng as oprng
from oracle import step as ostep

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _oracle_steps(om, guide, psi0, key, nsteps, m, S, x, y, N, B, tau, alpha, optimizer, lr, dtype=np.float64):
    psi = psi0.astype(dtype)
    opt = (np.zeros_like(psi), np.zeros_like(psi))
    losses = []
    for it in range(nsteps):
        key, attractive = ostep.draw_keys(key)
        eps, idx = ostep.draw_noise(attractive, m, S, om, guide, N if B else None, B)
        out = ostep.smi_step(om, psi, opt, it, eps, idx, x, y, N, guide, tau, alpha, 1.0, optimizer, lr,
                             dtype=dtype)
        psi, opt = out["psi"], out["opt_state"]
        losses.append(float(out["loss"]))
    return psi, np.array(losses), key


@pytest.mark.parametrize("graph", [False, True])
def test_uci_smi_steps_match_oracle(cuda, graph):
    import smi_experiments_b200 as smi
    I, N, B, m, S = 6, 50, 10, 4, 6
    rng = np.random.default_rng(0)
    x = rng.standard_normal((N, I)).astype(np.float32)
    y = rng.standard_normal(N).astype(np.float32)
    model = smi.uci_bnn(I)
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                      smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S, use_cuda_graph=graph)
    state = eng.init(smi.PRNGKey(1), x=x, y=y, subsample=B)
    psi0 = state.optim_state.x.cpu().numpy().copy()
    key0 = state.rng_key.cpu().numpy().view(np.uint32).copy()
    loss_eval = float(eng.evaluate(state, x=x, y=y, subsample=B))
    nsteps = 4
    losses = []
    for _ in range(nsteps):
        state, loss = eng.update(state, x=x, y=y, subsample=B)
        losses.append(float(loss))
    psi_ref, loss_ref, key_ref = _oracle_steps(omodels.uci_bnn(I), "normal", psi0, key0, nsteps, m, S, x, y, N, B,
                                               1.0, 1.0, "adam", 5e-3)
    assert abs(loss_eval - loss_ref[0]) <= 1e-5 * loss_ref[0]      # evaluate uses the next update's key
    np.testing.assert_allclose(losses, loss_ref, rtol=1e-5)
    assert _rel(state.optim_state.x.cpu().numpy(), psi_ref) <= 1e-5
    assert (state.rng_key.cpu().numpy().view(np.uint32) == key_ref).all()
    assert int(state.optim_state.step.item()) == nsteps
    assert eng.launches_per_update > 0


def test_svgd_gaussian_run_matches_oracle(cuda):
    import smi_experiments_b200 as smi
    d, m, nsteps = 30, 20, 5
    model = smi.GaussianTarget(d)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=m, repulsion_temperature=10.0,
                   guide_kwargs={"init_loc_fn": smi.init_to_uniform(radius=2.0)})
    state = eng.init(smi.PRNGKey(2))
    psi0 = state.optim_state.x.cpu().numpy().copy()
    key0 = state.rng_key.cpu().numpy().view(np.uint32).copy()
    res = eng.run(smi.PRNGKey(2), nsteps, progress_bar=False, init_state=state)
    psi_ref, loss_ref, _ = _oracle_steps(omodels.gaussian(d), "delta", psi0, key0, nsteps, m, 1, None, None, None,
                                         None, 1.0 / m, 10.0, "adam", 0.05)
    assert res.losses.shape == (nsteps,)
    np.testing.assert_allclose(res.losses.cpu().numpy(), loss_ref, rtol=1e-5)
    assert _rel(res.params["x_auto_loc"].cpu().numpy(), psi_ref) <= 1e-5


def test_smi_gaussian_adagrad_single_particle(cuda):
    """m = 1 SMI (variance_collapse.py SMI-1): h = 1, no repulsion; Adagrad path."""
    import smi_experiments_b200 as smi
    d, S, nsteps = 10, 100, 3
    model = smi.GaussianTarget(d)
    eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adagrad(0.05), smi.RBFKernel(), num_stein_particles=1,
                      num_elbo_particles=S)
    state = eng.init(smi.PRNGKey(3))
    psi0 = state.optim_state.x.cpu().numpy().copy()
    key0 = state.rng_key.cpu().numpy().view(np.uint32).copy()
    for _ in range(nsteps):
        state, loss = eng.update(state)
    psi_ref, loss_ref, _ = _oracle_steps(omodels.gaussian(d), "normal", psi0, key0, nsteps, 1, S, None, None, None,
                                         None, 1.0, 1.0, "adagrad", 0.05)
    assert abs(float(loss) - loss_ref[-1]) <= 1e-5 * loss_ref[-1]
    assert _rel(state.optim_state.x.cpu().numpy(), psi_ref) <= 1e-5
    p = eng.get_params(state)
    assert set(p) == {"x_auto_loc", "x_auto_scale"} and (p["x_auto_scale"] > 0).all()


def test_mnist_minibatch_update_host_buffers(cuda):
    """The engine_run.py loop shape: uint8 image batches on the HOST, n / batch_size kwargs."""
    import smi_experiments_b200 as smi
    model = smi.mnist_mlp()
    m, S, B, n = 2, 3, 128, 60000
    eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adagrad(0.8), smi.RBFKernel(), num_stein_particles=m,
                      num_elbo_particles=S)
    rng = np.random.default_rng(5)
    xb = rng.integers(0, 256, (B, 28, 28, 1)).astype(np.uint8)
    yb = rng.integers(0, 10, B).astype(np.int64)
    state = eng.init(smi.PRNGKey(4), xb, yb, n=n, batch_size=B)
    psi0 = state.optim_state.x.cpu().numpy().copy()
    key0 = state.rng_key.cpu().numpy().view(np.uint32).copy()
    state, loss = eng.update(state, xb, yb, n=n, batch_size=B)
    assert eng.h2d_bytes == xb.nbytes + yb.nbytes
    om = omodels.mnist_mlp()
    key, attractive = ostep.draw_keys(key0)
    eps, _ = ostep.draw_noise(attractive, m, S, om, "normal", n, 1)     # plate consumes one subkey
    idx = np.broadcast_to(np.arange(B), (m, S, B))
    out = ostep.smi_step(om, psi0.astype(np.float64), (np.zeros_like(psi0, np.float64),) * 2, 0, eps, idx,
                         xb.reshape(B, -1).astype(np.float32), yb, n, "normal", 1.0, 1.0, 1.0, "adagrad", 0.8,
                         dtype=np.float64)
    assert abs(float(loss) - float(out["loss"])) <= 1e-5 * float(out["loss"])
    assert _rel(state.optim_state.x.cpu().numpy(), out["psi"]) <= 1e-5


def test_asvgd_schedule_and_run(cuda):
    import smi_experiments_b200 as smi
    model = smi.reg_1d_bnn(5)
    eng = smi.ASVGD(model, smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=5, num_cycles=2, trans_speed=3)
    x = np.linspace(-1, 1, 40).astype(np.float32)
    y = np.sin(x).astype(np.float32)
    res = eng.run(smi.PRNGKey(0), 8, x, y, progress_bar=False)
    assert res.losses.shape == (8,) and torch.isfinite(res.losses).all()
    cyc = smi.ASVGD._cyclical_annealing(8, 2, 3)
    assert cyc(4) == 1 and 0 < cyc(1) < 1
