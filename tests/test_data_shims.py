"""Host-side data shims against the formats of the reference's loaders
(datasets/uci_reg/load_uci.py:11-76, src/preprocess.py:4-11, load_ood_detect.py:33-46)."""
import os

import numpy as np
import pytest

from smi_experiments_b200 import data

REF_UCI = "/root/reference/datasets/uci_reg"


def _make_uci(tmp_path, n=40, f=3, splits=2):
    rng = np.random.default_rng(0)
    d = tmp_path / "toy" / "data"
    d.mkdir(parents=True)
    arr = rng.standard_normal((n, f + 1))
    arr[:, 1] = 2.5                                     # a constant feature: std 0 -> 1
    np.savetxt(d / "data.txt", arr)
    np.savetxt(d / "n_splits.txt", np.asarray([splits]))
    for i in range(splits):
        perm = rng.permutation(n)
        np.savetxt(d / ("index_train_%d.txt" % i), perm[:30])
        np.savetxt(d / ("index_test_%d.txt" % i), perm[30:])
    (tmp_path / "not_a_dataset").mkdir()
    return arr


def test_uci_loader_and_splits(tmp_path):
    arr = _make_uci(tmp_path)
    ds = data.load_datasets(str(tmp_path))
    assert list(ds) == ["toy"]
    x, y = ds["toy"]
    assert x.shape == (40, 3) and y.shape == (40,) and x.dtype == np.float32
    assert np.allclose(x, arr[:, :-1].astype(np.float32)) and np.allclose(y, arr[:, -1].astype(np.float32))
    sp = data.load_standard_splits(str(tmp_path))["toy"]
    assert len(sp) == 2 and sp[0]["val"] is None
    assert sorted(np.concatenate([sp[0]["tr"], sp[0]["te"]]).tolist()) == list(range(40))
    spv = data.load_standard_splits(str(tmp_path), use_validation=True)["toy"]
    assert len(spv[0]["val"]) == 3 and len(spv[0]["tr"]) == 27          # ceil(0.1 * 30) validation rows
    assert not set(spv[0]["val"]) & set(spv[0]["tr"])
    gap = data.load_modified_gap_splits(str(tmp_path))["toy"]
    assert len(gap) == 3
    for dim, s in enumerate(gap):
        assert len(s["te"]) == 26 - 13 and len(s["tr"]) == 40 - 13      # floor(40/3) .. floor(80/3)
        if dim != 1:
            assert x[s["te"], dim].min() >= np.sort(x[:, dim])[12] and x[s["te"], dim].max() <= np.sort(x[:, dim])[26]


def test_normalize_matches_reference_definition():
    rng = np.random.default_rng(1)
    x = rng.standard_normal((50, 4)).astype(np.float32) * 3 + 1
    x[:, 2] = 7.0
    xn, mean, std = data.normalize(x)
    assert np.allclose(xn[:, [0, 1, 3]].mean(0), 0, atol=1e-5) and np.allclose(xn[:, [0, 1, 3]].std(0), 1, atol=1e-5)
    assert std[0, 2] == 1.0 and np.all(xn[:, 2] == 0)
    xt, m2, s2 = data.normalize(x[:5] + 1, mean, std)                   # test rows use the train statistics
    assert m2 is mean and s2 is std and np.allclose(xt, (x[:5] + 1 - mean) / std)


def test_train_batches_shuffle_and_drop_remainder():
    x, y = data.synthetic_images(n=1000, seed=0)
    assert x.dtype == np.uint8 and x.shape == (1000, 28, 28, 1) and y.max() <= 9
    it, n = data.train_batches(128, (x, y), seed=5)
    batches = list(it)
    assert n == 1000 and len(batches) == 7 and batches[0][0].shape == (128, 28, 28, 1)
    it2, _ = data.train_batches(128, (x, y), seed=5)
    assert np.array_equal(next(it2)[1], batches[0][1])
    it3, _ = data.train_batches(128, (x, y), seed=6)
    assert not np.array_equal(next(it3)[1], batches[0][1])


@pytest.mark.skipif(not os.path.isdir(REF_UCI), reason="reference datasets not mounted (GPU box)")
def test_reads_the_reference_uci_tree():
    ds = data.load_datasets(REF_UCI)
    assert "boston-housing" in ds
    x, y = ds["boston-housing"]
    assert x.shape == (506, 13) and y.shape == (506,)                    # SURVEY.md 8 config C2 shape
    sp = data.load_standard_splits(REF_UCI)["boston-housing"]
    assert len(sp) == 20 and len(sp[0]["tr"]) + len(sp[0]["te"]) == 506
