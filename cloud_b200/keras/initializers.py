"""``keras.initializers``.  Reference usage: ``tf.keras.initializers.Constant([[0.5]])`` in the
cloud_fit test model (experimental/cloud_fit/tests/unit/client_test.py:69-90); the other classes are
the ones Keras layers name as defaults.  An initializer fills a float32 tensor of the weight's
Keras shape (Dense ``[in, out]``, Conv2D ``[h, w, in, out]``) given its fans."""
from __future__ import annotations

import math

import numpy as np
import torch


class Initializer:
    name = None

    def fill(self, t: torch.Tensor, fan_in: int, fan_out: int) -> torch.Tensor:
        raise NotImplementedError

    def __call_tensor__(self, t, fan_in=1, fan_out=1):
        return self.fill(t, fan_in, fan_out)

    def get_config(self):
        return {}


class Constant(Initializer):
    name = "constant"

    def __init__(self, value=0.0):
        self.value = value

    def fill(self, t, fan_in, fan_out):
        v = torch.as_tensor(np.asarray(self.value, dtype=np.float32))
        if v.numel() == 1:
            return t.fill_(float(v.reshape(())))
        return t.copy_(v.reshape(t.shape))

    def get_config(self):
        return {"value": np.asarray(self.value).tolist()}


class Zeros(Constant):
    name = "zeros"

    def __init__(self):
        super().__init__(0.0)

    def get_config(self):
        return {}


class Ones(Constant):
    name = "ones"

    def __init__(self):
        super().__init__(1.0)

    def get_config(self):
        return {}


def _generator(seed):
    if seed is None:
        return None
    g = torch.Generator()
    g.manual_seed(int(seed))
    return g


class RandomNormal(Initializer):
    name = "random_normal"

    def __init__(self, mean=0.0, stddev=0.05, seed=None):
        self.mean, self.stddev, self.seed = mean, stddev, seed

    def fill(self, t, fan_in, fan_out):
        return t.normal_(self.mean, self.stddev, generator=_generator(self.seed))

    def get_config(self):
        return {"mean": self.mean, "stddev": self.stddev, "seed": self.seed}


class TruncatedNormal(RandomNormal):
    name = "truncated_normal"

    def fill(self, t, fan_in, fan_out):
        return torch.nn.init.trunc_normal_(t, self.mean, self.stddev, self.mean - 2 * self.stddev,
                                           self.mean + 2 * self.stddev, generator=_generator(self.seed))


class RandomUniform(Initializer):
    name = "random_uniform"

    def __init__(self, minval=-0.05, maxval=0.05, seed=None):
        self.minval, self.maxval, self.seed = minval, maxval, seed

    def fill(self, t, fan_in, fan_out):
        return t.uniform_(self.minval, self.maxval, generator=_generator(self.seed))

    def get_config(self):
        return {"minval": self.minval, "maxval": self.maxval, "seed": self.seed}


class VarianceScaling(Initializer):
    name = "variance_scaling"

    def __init__(self, scale=1.0, mode="fan_in", distribution="truncated_normal", seed=None):
        if mode not in ("fan_in", "fan_out", "fan_avg"):
            raise ValueError("Invalid `mode` argument: {}".format(mode))
        self.scale, self.mode, self.distribution, self.seed = scale, mode, distribution, seed

    def fill(self, t, fan_in, fan_out):
        n = {"fan_in": fan_in, "fan_out": fan_out, "fan_avg": (fan_in + fan_out) / 2.0}[self.mode]
        var = self.scale / max(1.0, n)
        g = _generator(self.seed)
        if self.distribution == "uniform":
            limit = math.sqrt(3.0 * var)
            return t.uniform_(-limit, limit, generator=g)
        std = math.sqrt(var)
        if self.distribution in ("truncated_normal", "normal"):
            std = std / 0.87962566103423978     # keeps the variance after clipping at two sigma
            return torch.nn.init.trunc_normal_(t, 0.0, std, -2 * std, 2 * std, generator=g)
        return t.normal_(0.0, std, generator=g)

    def get_config(self):
        return {"scale": self.scale, "mode": self.mode, "distribution": self.distribution, "seed": self.seed}


class GlorotUniform(VarianceScaling):
    name = "glorot_uniform"

    def __init__(self, seed=None):
        super().__init__(1.0, "fan_avg", "uniform", seed)

    def get_config(self):
        return {"seed": self.seed}


class GlorotNormal(VarianceScaling):
    name = "glorot_normal"

    def __init__(self, seed=None):
        super().__init__(1.0, "fan_avg", "truncated_normal", seed)

    def get_config(self):
        return {"seed": self.seed}


class HeNormal(VarianceScaling):
    name = "he_normal"

    def __init__(self, seed=None):
        super().__init__(2.0, "fan_in", "truncated_normal", seed)

    def get_config(self):
        return {"seed": self.seed}


class HeUniform(VarianceScaling):
    name = "he_uniform"

    def __init__(self, seed=None):
        super().__init__(2.0, "fan_in", "uniform", seed)

    def get_config(self):
        return {"seed": self.seed}


class LecunNormal(VarianceScaling):
    name = "lecun_normal"

    def __init__(self, seed=None):
        super().__init__(1.0, "fan_in", "truncated_normal", seed)

    def get_config(self):
        return {"seed": self.seed}


_BY_NAME = {c.name: c for c in (Constant, Zeros, Ones, RandomNormal, TruncatedNormal, RandomUniform,
                                VarianceScaling, GlorotUniform, GlorotNormal, HeNormal, HeUniform, LecunNormal)}
_BY_CLASS = {c.__name__: c for c in _BY_NAME.values()}


def get(identifier):
    """str / dict / instance -> Initializer."""
    if identifier is None:
        return GlorotUniform()
    if isinstance(identifier, Initializer):
        return identifier
    if isinstance(identifier, str):
        cls = _BY_NAME.get(identifier) or _BY_CLASS.get(identifier)
        if cls is None:
            raise ValueError("Unknown initializer: {}".format(identifier))
        return cls()
    if isinstance(identifier, dict):
        cls = _BY_CLASS.get(identifier.get("class_name")) or _BY_NAME.get(identifier.get("class_name"))
        if cls is None:
            raise ValueError("Unknown initializer: {}".format(identifier))
        return cls(**identifier.get("config", {}))
    raise ValueError("Could not interpret initializer identifier: {}".format(identifier))


def serialize(init):
    init = get(init)
    return {"class_name": type(init).__name__, "config": init.get_config()}
