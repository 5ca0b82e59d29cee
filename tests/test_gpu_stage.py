"""Host-buffer staging (smi_stage_*): what lands on the device is what the host held at call time,
also when the host runs ahead of the stream and reuses its array."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_stage_h2d_roundtrip_and_reuse():
    from smi_experiments_b200 import _lib, ops
    n = 1 << 16
    h = C.c_void_p()
    _lib.call("smi_stage_create", n * 4, C.byref(h))
    try:
        dev = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(8)]
        src = np.empty(n, dtype=np.float32)
        for i in range(8):            # the same host array is overwritten right after each call
            src[:] = float(i) + np.arange(n, dtype=np.float32) / n
            _lib.call("smi_stage_h2d", h, src.ctypes.data, n * 4, dev[i].data_ptr(), ops._stream())
        torch.cuda.synchronize()
        for i in range(8):
            want = float(i) + np.arange(n, dtype=np.float32) / n
            assert np.array_equal(dev[i].cpu().numpy(), want)
        with pytest.raises(_lib.SmiError):
            _lib.call("smi_stage_h2d", h, src.ctypes.data, n * 8, dev[0].data_ptr(), ops._stream())
    finally:
        _lib.call("smi_stage_destroy", h)


def test_update_from_host_arrays_matches_device_inputs():
    import smi_experiments_b200 as smi
    rng = np.random.default_rng(0)
    x = rng.normal(size=(64, 5)).astype(np.float32)
    y = rng.normal(size=64).astype(np.float32)
    losses = []
    for host in (True, False):
        model = smi.uci_bnn(5)
        eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adam(1e-2), smi.RBFKernel(), num_stein_particles=4,
                          num_elbo_particles=6)
        xs, ys = (x, y) if host else (torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
        st = eng.init(smi.PRNGKey(3), x=xs, y=ys, subsample=16)
        out = []
        for _ in range(12):           # no host sync between steps: the staging must keep up
            st, loss = eng.update(st, x=xs, y=ys, subsample=16)
            out.append(loss.clone())
        losses.append(torch.stack(out).cpu().numpy())
    assert np.array_equal(losses[0], losses[1])
