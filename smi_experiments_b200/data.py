"""Data shims either side of the path: the formats the reference's run
drivers hand to ``engine.update``.

  normalize                  src/preprocess.py:4-11 (train-set z-score, zero std -> 1)
  load_datasets / splits     datasets/uci_reg/load_uci.py:11-76 (``<name>/data/data.txt``, last column
                             is the target; ``index_train_<i>.txt`` / ``index_test_<i>.txt`` /
                             ``n_splits.txt``; gap splits by sorting one feature)
  train_batches              datasets/ood_detection/load_ood_detect.py:33-46 (shuffled, drop_remainder
                             minibatches of uint8 images + integer labels), over in-memory arrays
                             instead of tensorflow-datasets (absent from this image)
Host-side numpy: none of this is on the device path; batches are staged through pinned memory by
``SteinVI._to_dev``.
"""
import math
import os
from collections import defaultdict

import numpy as np


def normalize(val, mean=None, std=None):
    """(val - mean) / std with train-set statistics; returns (normalized, mean, std)."""
    val = np.asarray(val, np.float32)
    if mean is None and std is None:
        std = val.std(0, keepdims=True)
        std = np.where(std == 0, np.float32(1.0), std).astype(np.float32)
        mean = val.mean(0, keepdims=True).astype(np.float32)
    return (val - mean) / std, mean, std


def _dataset_dirs(root):
    for name in sorted(os.listdir(root)):
        d = os.path.join(root, name)
        if os.path.isdir(d) and os.path.exists(os.path.join(d, "data", "data.txt")):
            yield name, d


def load_datasets(root):
    """{name: (x (n, f) float32, y (n,) float32)} for every ``<root>/<name>/data/data.txt``."""
    out = {}
    for name, d in _dataset_dirs(root):
        data = np.loadtxt(os.path.join(d, "data", "data.txt"))
        out[name] = (data[:, :-1].astype(np.float32), data[:, -1].astype(np.float32))
    return out


def _train_val(tr, use_validation, rng):
    if not use_validation:
        return tr.astype(int), None
    # sklearn.model_selection.train_test_split(tr, test_size=0.1): shuffled, ceil(0.1 n) validation rows
    perm = rng.permutation(len(tr))
    n_val = int(math.ceil(0.1 * len(tr)))
    return tr[perm[n_val:]].astype(int), tr[perm[:n_val]].astype(int)


def load_standard_splits(root, use_validation=False, seed=0):
    """{name: [{"tr", "val", "te"}]} from the index files of https://arxiv.org/abs/1811.09385."""
    rng = np.random.default_rng(seed)
    splits = defaultdict(list)
    for name, d in _dataset_dirs(root):
        n = int(np.loadtxt(os.path.join(d, "data", "n_splits.txt")).item())
        for i in range(n):
            tr = np.loadtxt(os.path.join(d, "data", "index_train_%d.txt" % i)).astype(int)
            te = np.loadtxt(os.path.join(d, "data", "index_test_%d.txt" % i)).astype(int)
            tr, val = _train_val(tr, use_validation, rng)
            splits[name].append({"tr": tr, "val": val, "te": te})
    return splits


def load_modified_gap_splits(root, use_validation=False, test_size=1.0 / 3.0, seed=0):
    """One split per feature: the middle ``test_size`` of the rows sorted by that feature is held out."""
    assert 0.0 <= test_size <= 1.0
    rng = np.random.default_rng(seed)
    splits = defaultdict(list)
    for name, (x, _) in load_datasets(root).items():
        n, f = x.shape
        tail = (1 - test_size) / 2
        before, after = math.floor(n * tail), math.floor(n * (tail + test_size))
        for dim in range(f):
            order = np.argsort(x[:, dim], kind="stable")
            tr = np.concatenate((order[:before], order[after:])).astype(int)
            te = order[before:after].astype(int)
            tr, val = _train_val(tr, use_validation, rng)
            splits[name].append({"tr": tr, "val": val, "te": te})
    return splits


def train_batches(batch_size, dataset, seed=None):
    """(iterator of (x_batch, y_batch), n): one shuffled epoch, remainder dropped.
    ``dataset`` = (images (n, ...) uint8 or float, labels (n,))."""
    x, y = dataset
    n = len(x)
    perm = np.random.default_rng(seed).permutation(n)

    def it():
        for b in range(n // batch_size):
            sel = perm[b * batch_size:(b + 1) * batch_size]
            yield x[sel], y[sel]

    return it(), n


def synthetic_images(n=60000, shape=(28, 28, 1), classes=10, seed=3):
    """MNIST-shape stand-in: uint8 pixels U{0..255}, labels U{0..classes-1}."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (n,) + tuple(shape), dtype=np.uint8), rng.integers(0, classes, n).astype(np.int32)
