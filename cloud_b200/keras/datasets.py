"""``keras.datasets`` stand-ins.  There is no network on this box, so ``load_data()`` returns
deterministic SYNTHETIC data of the named shape and dtype (class-dependent templates plus noise, so
models can actually learn on it).  Reference usage: ``tf.keras.datasets.mnist.load_data()``
README.md:66-70, tuner/tests/integration/tuner_integration_test.py:51-57."""
from __future__ import annotations

import numpy as np


def _synthetic(n_train, n_test, shape, classes, seed):
    rng = np.random.default_rng(seed)
    templates = rng.integers(0, 256, size=(classes,) + shape).astype(np.float32)

    def make(n, s):
        r = np.random.default_rng(s)
        y = r.integers(0, classes, size=n).astype(np.uint8)
        x = np.empty((n,) + shape, dtype=np.uint8)
        step = 8192
        for i in range(0, n, step):
            yy = y[i:i + step]
            noise = r.normal(0.0, 48.0, size=(len(yy),) + shape).astype(np.float32)
            x[i:i + step] = np.clip(0.6 * templates[yy] + noise + 40.0, 0, 255).astype(np.uint8)
        return x, y
    return make(n_train, seed + 1), make(n_test, seed + 2)


class _Loader:
    def __init__(self, shape, classes, n_train, n_test, seed, label_2d=False):
        self.shape, self.classes, self.n_train, self.n_test = shape, classes, n_train, n_test
        self.seed, self.label_2d = seed, label_2d
        self._cache = None

    def load_data(self, path=None):
        if self._cache is None:
            (xt, yt), (xv, yv) = _synthetic(self.n_train, self.n_test, self.shape, self.classes,
                                            self.seed)
            if self.label_2d:
                yt, yv = yt.reshape(-1, 1), yv.reshape(-1, 1)
            self._cache = ((xt, yt), (xv, yv))
        return self._cache


mnist = _Loader((28, 28), 10, 60000, 10000, seed=1)
fashion_mnist = _Loader((28, 28), 10, 60000, 10000, seed=2)
cifar10 = _Loader((32, 32, 3), 10, 50000, 10000, seed=3, label_2d=True)
cifar100 = _Loader((32, 32, 3), 100, 50000, 10000, seed=4, label_2d=True)
