"""The no-allocation entry point smi_median_bandwidth_noalloc_f32 (include/smi.h; no counterpart in the reference)
against the existing smi_median_bandwidth_f32, through the C ABI.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(cuda, D2, mode):
    from smi_experiments_b200 import _lib, ops
    m = D2.shape[0]
    h0, r0 = ops.median_bandwidth(D2, mode)                      # existing entry point, caller-owned workspace
    h1 = torch.full((1,), -1.0, dtype=torch.float32, device=cuda)
    r1 = torch.zeros(2, dtype=torch.int32, device=cuda)
    _lib.call("smi_median_bandwidth_noalloc_f32", ops._ptr(D2), m, mode, ops._ptr(h1), ops._ptr(r1), ops._stream())
    torch.cuda.synchronize()
    return (h0.cpu().numpy(), r0.cpu().numpy()), (h1.cpu().numpy(), r1.cpu().numpy())


def test_noalloc_matches_workspace_entry_two_calls_in_a_row(cuda, lib):
    from smi_experiments_b200 import _lib, ops
    assert lib.smi_median_bandwidth_noalloc_max_m() == 32768
    rng = np.random.default_rng(123)
    # two different problems back to back: the shared scratch array keeps the first call's contents
    for m, d, mode in ((300, 257, _lib.MEDIAN_JAX_FP32_INDEX), (1024, 64, _lib.MEDIAN_EXACT_MIDPOINT),
                       (300, 257, _lib.MEDIAN_JAX_FP32_INDEX)):
        X = torch.from_numpy(rng.standard_normal((m, d)).astype(np.float32) * (1.0 + 0.01 * m)).to(cuda)
        D2 = ops.sqdist(X, path=ops.PATH_TC)
        (h0, r0), (h1, r1) = _both(cuda, D2, mode)
        assert (r0 == r1).all()                      # the two selected order statistics: bit-exact (integer select)
        assert h0.tobytes() == h1.tobytes()          # and therefore the same bandwidth bits
        assert np.isfinite(h1).all() and h1[0] > 0
    # the same problem twice in a row gives the same bits
    (h0, r0), (h1, r1) = _both(cuda, D2, mode)
    (_, _), (h2, r2) = _both(cuda, D2, mode)
    assert h1.tobytes() == h2.tobytes() and (r1 == r2).all()
    # m == 1: h = 1 by definition, as the existing entry point
    one = torch.zeros(1, 1, dtype=torch.float32, device=cuda)
    (h0, _), (h1, _) = _both(cuda, one, _lib.MEDIAN_JAX_FP32_INDEX)
    assert float(h0[0]) == float(h1[0]) == 1.0


def test_noalloc_rejects_sizes_beyond_its_maximum(cuda, lib):
    from smi_experiments_b200 import _lib, ops
    D2 = torch.zeros(4, 4, dtype=torch.float32, device=cuda)      # never read: the size check comes first
    h = torch.zeros(1, dtype=torch.float32, device=cuda)
    with pytest.raises(_lib.SmiError) as ei:
        _lib.call("smi_median_bandwidth_noalloc_f32", ops._ptr(D2), 32769, 1, ops._ptr(h), C.c_void_p(0), ops._stream())
    assert ei.value.code == -5 and "32768" in str(ei.value)
    with pytest.raises(_lib.SmiError):
        _lib.call("smi_median_bandwidth_noalloc_f32", C.c_void_p(0), 4, 1, ops._ptr(h), C.c_void_p(0), ops._stream())
