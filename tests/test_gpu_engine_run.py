"""Epoch/minibatch run driver (reference src/engine_run.py:11-70) over the device-side loss ring."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _engine(smi, graph=True):
    model = smi.BNNClassifier(in_dim=36, hidden=16, out_dim=4)
    return smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)), smi.Adagrad(0.05),
                       smi.RBFKernel(), num_stein_particles=3, num_elbo_particles=4, use_cuda_graph=graph)


def test_setup_run_matches_manual_loop(cuda):
    import smi_experiments_b200 as smi
    from smi_experiments_b200 import data, engine_run
    x, y = data.synthetic_images(n=200, shape=(6, 6, 1), classes=4, seed=1)
    B, epochs = 32, 2
    res = engine_run.setup_run(_engine(smi), (x, y), seed=9)(smi.PRNGKey(3), epochs, B, None)
    # manual: the reference's loop shape, one device_get per step (src/engine_run.py:36-56)
    eng = _engine(smi, graph=False)
    state, losses = None, []
    for e in range(epochs):
        batches, n = data.train_batches(B, (x, y), 9 + e)
        for xb, yb in batches:
            if state is None:
                state = eng.init(smi.PRNGKey(3), xb, yb, n=n, batch_size=B)
                losses.append(float(eng.evaluate(state, xb, yb, n=n, batch_size=B)))
            else:
                state, l = eng.update(state, xb, yb, n=n, batch_size=B)
                losses.append(float(l))
    assert res.losses.shape == (epochs * (200 // B),)
    assert np.array_equal(res.losses.cpu().numpy(), np.asarray(losses, np.float32))
    assert torch.equal(res.state.optim_state.x, state.optim_state.x)
    assert set(res.params) == set(eng.get_params(state))
    with pytest.raises(ValueError):
        engine_run.setup_run(eng, (x, y))(smi.PRNGKey(0), 0, B, None)
