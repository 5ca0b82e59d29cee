"""Full SteinVI.update / run through the engine API against the oracle step (rows a1, a2, a7, a10).

Every step is compared on IDENTICAL inputs: the device state before the update (particles,
optimizer slots, step counter, PRNG key) is handed to the float64 oracle, and the state after the
update is compared with the oracle's.  (Comparing whole trajectories instead would measure how Adam's
normalisation amplifies rounding noise, not the step.)
"""
import numpy as np
import pytest
import torch

from oracle import models as omodels
from oracle import step as ostep

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _snap(state):
    os_ = state.optim_state
    return dict(psi=os_.x.cpu().numpy().astype(np.float64), a=os_.a.cpu().numpy().astype(np.float64),
                b=os_.b.cpu().numpy().astype(np.float64), it=int(os_.step.item()),
                key=state.rng_key.cpu().numpy().view(np.uint32).copy())


def _oracle_step(om, guide, snap, m, S, x, y, N, B, tau, alpha, optimizer, lr, plate_only=False, dtype=np.float64):
    next_key, attractive = ostep.draw_keys(snap["key"])
    if plate_only:      # a subsampling plate that consumes its subkey but whose indices are unused
        eps, _ = ostep.draw_noise(attractive, m, S, om, guide, N, 1)
        idx = np.broadcast_to(np.arange(B), (m, S, B))
    else:
        eps, idx = ostep.draw_noise(attractive, m, S, om, guide, N if (B and B < N) else None,
                                    B if (B and B < N) else None)
    out = ostep.smi_step(om, snap["psi"].astype(dtype), (snap["a"].astype(dtype), snap["b"].astype(dtype)), snap["it"], eps,
                         idx, x, y, N, guide, tau, alpha, 1.0, optimizer, lr, dtype=dtype)
    out["key"] = next_key
    return out


def _check(state, loss, ref, r32=None):
    """r32: the float32 oracle's result of the same step.  Where fp32 itself is ill-conditioned (ReLU kinks at the
    full C2 size: 2 x 10^7 pre-activations per step, a few of them within rounding of zero, each flipping one
    0/1 derivative -- the fp32 oracle differs from float64 by ~1.5e-5 there) the tolerance is
    max(1e-5, 4 x the fp32 oracle's own distance to float64): the device's 3xTF32 pre-activations carry a few
    times the rounding error of a plain fp32 dot product, hence proportionally more flips."""
    def tol(a32, a64, base=RTOL):
        return base if r32 is None else max(base, 4.0 * _rel(a32, a64))
    tl = RTOL if r32 is None else max(RTOL, 4.0 * abs(float(r32["loss"]) - float(ref["loss"])) / float(ref["loss"]))
    assert abs(float(loss) - float(ref["loss"])) <= tl * float(ref["loss"])
    assert _rel(state.optim_state.x.cpu().numpy(), ref["psi"]) <= tol(r32["psi"] if r32 else None, ref["psi"])
    assert _rel(state.optim_state.a.cpu().numpy(), ref["opt_state"][0]) <= tol(r32["opt_state"][0] if r32 else None,
                                                                              ref["opt_state"][0])
    assert _rel(state.optim_state.b.cpu().numpy(), ref["opt_state"][1]) <= tol(r32["opt_state"][1] if r32 else None,
                                                                              ref["opt_state"][1], 2 * RTOL)
    assert (state.rng_key.cpu().numpy().view(np.uint32) == ref["key"]).all()


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
    psi0 = state.optim_state.x.cpu().numpy()
    r1 = eng.layout.rows[1]
    assert np.abs(psi0[:, r1["loc"]:r1["loc"] + r1["size"]]).max() <= 0.1          # init_to_uniform(radius)
    loss_eval = float(eng.evaluate(state, x=x, y=y, subsample=B))
    om = omodels.uci_bnn(I)
    for it in range(5):
        snap = _snap(state)
        state, loss = eng.update(state, x=x, y=y, subsample=B)
        ref = _oracle_step(om, "normal", snap, m, S, x, y, N, B, 1.0, 1.0, "adam", 5e-3)
        _check(state, loss, ref)
        if it == 0:     # evaluate uses the next update's key and leaves the state alone
            assert abs(loss_eval - float(ref["loss"])) <= RTOL * float(ref["loss"])
    assert int(state.optim_state.step.item()) == 5
    assert eng.launches_per_update > 0


def test_svgd_gaussian_run_matches_oracle(cuda):
    import smi_experiments_b200 as smi
    d, m, nsteps = 30, 20, 5
    model = smi.GaussianTarget(d)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=m, repulsion_temperature=10.0,
                   guide_kwargs={"init_loc_fn": smi.init_to_uniform(radius=2.0)})
    state = eng.init(smi.PRNGKey(2))
    om = omodels.gaussian(d)
    losses = []
    for it in range(nsteps):
        snap = _snap(state)
        state, loss = eng.update(state)
        ref = _oracle_step(om, "delta", snap, m, 1, None, None, None, None, 1.0 / m, 10.0, "adam", 0.05)
        _check(state, loss, ref)
        losses.append(float(loss))
    # run(): same trajectory from the same key, losses collected on the device
    res = eng.run(smi.PRNGKey(2), nsteps, progress_bar=False)
    assert res.losses.shape == (nsteps,)
    np.testing.assert_allclose(res.losses.cpu().numpy(), losses, rtol=1e-6)
    assert set(res.params) == {"x_auto_loc"} and res.params["x_auto_loc"].shape == (m, d)
    # resume from a state: the ring offset follows the step counter
    res2 = eng.run(smi.PRNGKey(2), 3, progress_bar=False, init_state=res.state)
    assert res2.losses.shape == (3,) and int(res2.state.optim_state.step.item()) == nsteps + 3
    assert torch.isfinite(res2.losses).all() and float(res2.losses[0]) > 0


def test_smi_gaussian_adagrad_single_particle(cuda):
    """m = 1 SMI (variance_collapse.py SMI-1): h = 1, no repulsion; Adagrad path."""
    import smi_experiments_b200 as smi
    d, S = 10, 100
    model = smi.GaussianTarget(d)
    eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adagrad(0.05), smi.RBFKernel(), num_stein_particles=1,
                      num_elbo_particles=S)
    state = eng.init(smi.PRNGKey(3))
    om = omodels.gaussian(d)
    for _ in range(3):
        snap = _snap(state)
        state, loss = eng.update(state)
        ref = _oracle_step(om, "normal", snap, 1, S, None, None, None, None, 1.0, 1.0, "adagrad", 0.05)
        _check(state, loss, ref)
    p = eng.get_params(state)
    assert set(p) == {"x_auto_loc", "x_auto_scale"} and (p["x_auto_scale"] > 0).all()


def test_smi_collapse_config(cuda):
    """variance_collapse.py shapes: SMI 20 particles x 100 draws, alpha in {0.1, 10}, d = 100."""
    import smi_experiments_b200 as smi
    d, m, S = 100, 20, 100
    model = smi.GaussianTarget(d)
    om = omodels.gaussian(d)
    for alpha in (0.1, 10.0):
        eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=2.0)),
                          smi.Adagrad(0.05), smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S,
                          repulsion_temperature=alpha)
        state = eng.init(smi.PRNGKey(5))
        for _ in range(2):
            snap = _snap(state)
            state, loss = eng.update(state)
            ref = _oracle_step(om, "normal", snap, m, S, None, None, None, None, 1.0, alpha, "adagrad", 0.05)
            _check(state, loss, ref)


def test_mnist_minibatch_update_host_buffers(cuda):
    """The engine_run.py loop shape: uint8 image batches on the HOST, n / batch_size kwargs."""
    import smi_experiments_b200 as smi
    model = smi.mnist_mlp()
    m, S, B, n = 2, 3, 128, 60000
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)), smi.Adagrad(0.8),
                      smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S)
    rng = np.random.default_rng(5)
    xb = (rng.integers(0, 256, (B, 28, 28, 1)) * (rng.random((B, 28, 28, 1)) < 0.2)).astype(np.uint8)
    yb = rng.integers(0, 10, B).astype(np.int64)
    state = eng.init(smi.PRNGKey(4), xb, yb, n=n, batch_size=B)
    snap = _snap(state)
    state, loss = eng.update(state, xb, yb, n=n, batch_size=B)
    assert eng.h2d_bytes == xb.nbytes + yb.nbytes
    om = omodels.mnist_mlp()
    xf = (xb.reshape(B, -1).astype(np.float32))
    ref = _oracle_step(om, "normal", snap, m, S, xf, yb, n, B, 1.0, 1.0, "adagrad", 0.8, plate_only=True)
    # raw-pixel inputs: calibrate with the fp32 oracle's own distance to float64
    next_key, attractive = ostep.draw_keys(snap["key"])
    eps, _ = ostep.draw_noise(attractive, m, S, om, "normal", n, 1)
    idx = np.broadcast_to(np.arange(B), (m, S, B))
    r32 = ostep.smi_step(om, snap["psi"].astype(np.float32), (snap["a"].astype(np.float32), snap["b"].astype(np.float32)),
                         snap["it"], eps, idx, xf, yb, n, "normal", 1.0, 1.0, 1.0, "adagrad", 0.8, dtype=np.float32)
    tol = max(RTOL, 2.0 * abs(float(r32["loss"]) - float(ref["loss"])) / float(ref["loss"]))
    assert abs(float(loss) - float(ref["loss"])) <= tol * float(ref["loss"])
    assert (state.rng_key.cpu().numpy().view(np.uint32) == ref["key"]).all()
    tolx = max(RTOL, 2.0 * _rel(r32["psi"], ref["psi"]))
    assert _rel(state.optim_state.x.cpu().numpy(), ref["psi"]) <= tolx


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


def test_setup_run_tuple(cuda):
    """The 5-tuple src/model_select/early_stopping.py:16-22 unpacks."""
    import smi_experiments_b200 as smi
    model = smi.GaussianTarget(8)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=6)
    step, diagnostic, collect, extract_state, info = eng.setup_run(smi.PRNGKey(0), 10, (), None, {})
    l0 = float(collect(info))
    info = step(info)
    assert float(collect(info)) > 0 and l0 > 0
    assert int(extract_state(info).optim_state.step.item()) == 1
    assert isinstance(diagnostic(info), str)


def test_asvgd_device_schedule_matches_host_formula(cuda):
    """smi_asvgd_tau_f32 against ASVGD._cyclical_annealing (numpyro; mirrored at src/engine_run.py:16-24)."""
    import smi_experiments_b200 as smi
    from smi_experiments_b200 import ops
    step = torch.zeros(1, dtype=torch.int32, device=cuda)
    tau = torch.zeros(1, device=cuda)
    for T, Cc, p, m, step0 in ((8, 2, 3, 5, 0), (100, 10, 10, 20, 7), (15000, 10, 10, 5, 0), (9, 1, 2, 3, 2)):
        cyc = smi.ASVGD._cyclical_annealing(T, Cc, p)
        for t in sorted(set([0, 1, 2, T // 2, T - 1, T, T + 3] + list(range(0, T + 1, max(T // 17, 1))))):
            step.fill_(step0 + t)
            ops.asvgd_tau(step, step0, T, Cc, p, m, tau)
            want = np.float32(cyc(t) / float(m))
            got = float(tau.item())
            assert abs(got - want) <= 1e-6 * max(abs(want), 1e-30), (T, Cc, p, t, got, want)


def test_asvgd_graph_run_matches_host_scheduled_eager_loop(cuda):
    """ASVGD.run with the schedule on the device (CUDA-graph replay) == an eager loop that sets
    loss_temperature = cyc(t) / m on the host before every update (the reference's body_fn)."""
    import smi_experiments_b200 as smi
    x = np.linspace(-1, 1, 40).astype(np.float32)
    y = np.sin(3 * x).astype(np.float32)
    T = 12

    def mk(graph):
        model = smi.reg_1d_bnn(5)
        return smi.ASVGD(model, smi.Adam(1e-2), smi.RBFKernel(), num_stein_particles=5, num_cycles=3, trans_speed=2,
                         use_cuda_graph=graph)
    res = mk(True).run(smi.PRNGKey(0), T, x, y, progress_bar=False)
    eng = mk(False)
    cyc = smi.ASVGD._cyclical_annealing(T, 3, 2)
    st = eng.init(smi.PRNGKey(0), x, y)
    losses = []
    for t in range(T):
        eng.loss_temperature = cyc(t) / 5.0
        st, l = eng.update(st, x, y)
        losses.append(float(l))
    np.testing.assert_allclose(res.losses.cpu().numpy(), np.asarray(losses, np.float32), rtol=2e-6)
    np.testing.assert_allclose(res.state.optim_state.x.cpu().numpy(), st.optim_state.x.cpu().numpy(), rtol=1e-5, atol=1e-7)


# ---- the shapes bench.py times (BASELINE.json configs[1] and configs[3]), one full update each ----------------
def test_c2_timed_shape_full_update(cuda):
    """C2 exactly as bench.py runs it: 455 x 13 train rows, BNN 13-50-50-1, 20 particles, 100 draws, minibatch 100,
    Adam 5e-3, CUDA-graph replay from the third call on (experiments/uci_benchmark.py:188-195)."""
    import smi_experiments_b200 as smi
    import bench
    xt, yt = bench.make_uci()
    c = bench.C2
    m, S, B, N = c["m"], c["S"], c["B"], xt.shape[0]
    model = smi.uci_bnn(c["in_dim"])
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=c["radius"])),
                      smi.Adam(c["lr"]), smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S)
    state = eng.init(smi.PRNGKey(1), x=xt, y=yt, subsample=B)
    om = omodels.uci_bnn(c["in_dim"])
    for it in range(3):          # eager, eager (warm), graph replay
        snap = _snap(state)
        state, loss = eng.update(state, x=xt, y=yt, subsample=B)
        ref = _oracle_step(om, "normal", snap, m, S, xt, yt, N, B, 1.0, 1.0, "adam", c["lr"])
        r32 = _oracle_step(om, "normal", snap, m, S, xt, yt, N, B, 1.0, 1.0, "adam", c["lr"], dtype=np.float32)
        _check(state, loss, ref, r32)
    assert eng._graph is not None


def test_c4_timed_shape_full_update(cuda):
    """C4 as bench.py runs it: MLP 784-100-100-10 tanh, 10 particles, 55 draws, batch 128 of n = 60000, Adagrad 0.8
    (experiments/img_classifier.py:66-73,170-177); pixel values scaled to [0, 1] as in bench.py."""
    import smi_experiments_b200 as smi
    model = smi.mnist_mlp()
    m, S, B, n = 10, 55, 128, 60000
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)), smi.Adagrad(0.8),
                      smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S)
    rng = np.random.default_rng(0)
    xb = (rng.integers(0, 256, (B, 784)).astype(np.float32) / 255.0)
    yb = rng.integers(0, 10, B).astype(np.int32)
    state = eng.init(smi.PRNGKey(4), xb, yb, n=n, batch_size=B)
    snap = _snap(state)
    state, loss = eng.update(state, xb, yb, n=n, batch_size=B)
    om = omodels.mnist_mlp()
    ref = _oracle_step(om, "normal", snap, m, S, xb, yb, n, B, 1.0, 1.0, "adagrad", 0.8, plate_only=True)
    assert abs(float(loss) - float(ref["loss"])) <= RTOL * float(ref["loss"])
    assert (state.rng_key.cpu().numpy().view(np.uint32) == ref["key"]).all()
    # Adagrad's FIRST step is lr * 0.1 * g / sqrt(g^2) = +-0.08 per coordinate: a sign function of the force, which no
    # finite precision can reproduce where the force is ~0 (one flipped coordinate of the 1.79 M is a 7.6e-5 error of
    # the whole particle matrix).  Compare where the step is well conditioned (|F| above 1e-3 of its rms: > 99 % of the
    # coordinates) at the strict tolerance, and the accumulator g^2 -- which is smooth -- everywhere.
    F = ref["F"]
    ok = np.abs(F) > 1e-3 * np.sqrt(np.mean(F * F))
    assert ok.mean() > 0.99
    assert _rel(state.optim_state.x.cpu().numpy()[ok], ref["psi"][ok]) <= RTOL
    assert _rel(state.optim_state.a.cpu().numpy(), ref["opt_state"][0]) <= 2 * RTOL
    assert _rel(state.optim_state.b.cpu().numpy()[ok], ref["opt_state"][1][ok]) <= RTOL


def test_returned_losses_do_not_alias(cuda):
    """`losses.append(loss)` without reading (the pattern NumPyro's fori_collect / a caller that stacks at the end
    uses): every returned scalar keeps its own value, on the eager, warm-up and graph-replay paths."""
    import smi_experiments_b200 as smi
    model = smi.GaussianTarget(12)
    eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=8)
    state = eng.init(smi.PRNGKey(7))
    kept, read = [], []
    for _ in range(6):
        state, loss = eng.update(state)
        kept.append(loss)
        read.append(float(loss))
    assert len(set(read)) == 6
    assert [float(v) for v in torch.stack(kept).cpu()] == read


def test_many_particle_engine_graph_replay_matches_eager(cuda):
    """100 particles on one GPU (the shapes of the 8-GPU weak-scaling runs): the tcgen05 mixture filter with its internal
    side-stream fork, the sparse gradient kernel and the mid-size Stein path captured into ONE CUDA graph; the replayed
    steps must equal the eagerly launched ones bit for bit, and the first of them the float64 oracle."""
    import smi_experiments_b200 as smi
    I, N, B, m, S = 13, 80, 20, 100, 6
    rng = np.random.default_rng(3)
    x = rng.standard_normal((N, I)).astype(np.float32)
    y = rng.standard_normal(N).astype(np.float32)
    model = smi.uci_bnn(I)

    def mk(graph):
        return smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                           smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S, use_cuda_graph=graph)
    eg, ee = mk(True), mk(False)
    sg, se = eg.init(smi.PRNGKey(9), x=x, y=y, subsample=B), ee.init(smi.PRNGKey(9), x=x, y=y, subsample=B)
    snap = _snap(se)
    om = omodels.uci_bnn(I)
    for it in range(5):
        sg, lg = eg.update(sg, x=x, y=y, subsample=B)
        se, le = ee.update(se, x=x, y=y, subsample=B)
        assert float(lg) == float(le)
        assert torch.equal(sg.optim_state.x, se.optim_state.x) and torch.equal(sg.rng_key, se.rng_key)
        if it == 0:
            ref = _oracle_step(om, "normal", snap, m, S, x, y, N, B, 1.0, 1.0, "adam", 5e-3)
            _check(se, le, ref)
    assert eg._graph is not None
