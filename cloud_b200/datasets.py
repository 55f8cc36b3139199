"""``tensorflow_datasets``-shaped catalogue (``import cloud_b200.datasets as tfds``).

The reference's notebooks and examples pull their data with ``tfds.load`` -
``tfds.load(name="mnist", with_info=True, as_supervised=True)``
(core/tests/examples/call_run_within_nb_on_colab.ipynb, testdata/mnist_example_using_fit.py:41-47)
and ``tfds.load("stanford_dogs", split=["train", "test"], ...)``
(core/tests/examples/dogs_classification.ipynb, call_run_within_script_with_keras_fit.py:29-37).
This box has no network, so every catalogue entry is SYNTHETIC data of the named shape, dtype and
class count (class-dependent templates plus noise, so a model can learn on it).  The number of
examples defaults to the real dataset's, capped by ``TFC_SYNTHETIC_SAMPLES``.
"""
from __future__ import annotations

import os

import numpy as np

from .data import Dataset

# name -> (image shape, classes, train examples, test examples, seed)
_CATALOGUE = {
    "mnist": ((28, 28, 1), 10, 60000, 10000, 11),
    "fashion_mnist": ((28, 28, 1), 10, 60000, 10000, 12),
    "cifar10": ((32, 32, 3), 10, 50000, 10000, 13),
    "cifar100": ((32, 32, 3), 100, 50000, 10000, 14),
    # the real images vary in size; the notebooks resize to 224 x 224 right after loading
    "stanford_dogs": ((64, 64, 3), 120, 12000, 8580, 15),
}


class _ClassLabel:
    def __init__(self, num_classes):
        self.num_classes = num_classes
        self.names = [str(i) for i in range(num_classes)]


class _Image:
    def __init__(self, shape):
        self.shape = shape
        self.dtype = np.uint8


class _Split:
    def __init__(self, num_examples):
        self.num_examples = num_examples


class DatasetInfo:
    def __init__(self, name, shape, classes, splits):
        self.name = name
        self.features = {"image": _Image(shape), "label": _ClassLabel(classes)}
        self.splits = {k: _Split(n) for k, n in splits.items()}
        self.description = "synthetic stand-in for {} (no network on this box)".format(name)

    def __repr__(self):
        return "DatasetInfo(name={!r}, synthetic=True, splits={})".format(
            self.name, {k: s.num_examples for k, s in self.splits.items()})


def _cap(n):
    limit = os.environ.get("TFC_SYNTHETIC_SAMPLES")
    return min(n, int(limit)) if limit else n


def _make(shape, classes, n, seed, split_seed):
    templates = np.random.default_rng(seed).integers(0, 256, size=(classes,) + shape).astype(np.float32)
    r = np.random.default_rng(seed * 1000 + split_seed)
    y = r.integers(0, classes, size=n).astype(np.int64)
    x = np.empty((n,) + shape, dtype=np.uint8)
    for i in range(0, n, 4096):
        yy = y[i:i + 4096]
        noise = r.normal(0.0, 48.0, size=(len(yy),) + shape).astype(np.float32)
        x[i:i + 4096] = np.clip(0.6 * templates[yy] + noise + 40.0, 0, 255).astype(np.uint8)
    return x, y


def _one_split(name, split, as_supervised):
    shape, classes, n_train, n_test, seed = _CATALOGUE[name]
    base, _, sl = split.partition("[")
    if base not in ("train", "test"):
        raise ValueError("Unknown split {!r}: {} has 'train' and 'test'".format(split, name))
    n = _cap(n_train if base == "train" else n_test)
    x, y = _make(shape, classes, n, seed, 1 if base == "train" else 2)
    if sl:                                          # "train[:80%]" / "train[1000:]"
        lo, _, hi = sl.rstrip("]").partition(":")

        def pos(text, default):
            if not text:
                return default
            return int(n * float(text[:-1]) / 100.0) if text.endswith("%") else int(text)
        x, y = x[pos(lo, 0):pos(hi, n)], y[pos(lo, 0):pos(hi, n)]
    if as_supervised:
        return Dataset.from_tensor_slices((x, y))
    return Dataset.from_tensor_slices({"image": x, "label": y})


def load(name, split=None, data_dir=None, batch_size=None, shuffle_files=False, download=True,
         as_supervised=False, with_info=False, **kwargs):
    """``tfds.load``: a Dataset (one split), a list of them (list of splits) or a dict
    ``{"train", "test"}`` (``split=None``); with ``with_info`` a ``(datasets, info)`` pair."""
    if name not in _CATALOGUE:
        raise ValueError("Dataset {!r} is not in the local catalogue {}".format(name, sorted(_CATALOGUE)))
    if split is None:
        out = {s: _one_split(name, s, as_supervised) for s in ("train", "test")}
    elif isinstance(split, (list, tuple)):
        out = [_one_split(name, s, as_supervised) for s in split]
    else:
        out = _one_split(name, split, as_supervised)
    if batch_size:
        batch = (lambda d: d.batch(batch_size))
        out = {k: batch(v) for k, v in out.items()} if isinstance(out, dict) else \
            [batch(v) for v in out] if isinstance(out, list) else batch(out)
    if not with_info:
        return out
    shape, classes, n_train, n_test, _seed = _CATALOGUE[name]
    return out, DatasetInfo(name, shape, classes, {"train": _cap(n_train), "test": _cap(n_test)})


def list_builders():
    return sorted(_CATALOGUE)


def disable_progress_bar():
    """Nothing is downloaded, so there is no progress bar to disable."""


def as_numpy(dataset):
    return dataset.as_numpy_iterator()
