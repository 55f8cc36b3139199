"""HyperParameters container and hyperparameter types (the KerasTuner 1.0.x surface the reference
builds on: tuner/tuner.py:21-25, tuner/utils.py:220-282, and SURVEY.md appendix B2).

``Int(name, min, max, step=1, sampling=None, default)``, ``Float(name, min, max, step=None,
sampling=None, default)``, ``Choice(name, values, ordered, default)``, ``Boolean(name, default)``,
``Fixed(name, value)``; the container offers the same registering calls plus ``get``, ``values``,
``space``, ``copy`` and ``get_config`` / ``from_config``.
"""
from __future__ import annotations

import contextlib
import copy
import math
import random


class HyperParameter:
    def __init__(self, name, default=None):
        self.name = name
        self._default = default

    @property
    def default(self):
        return self._default

    def random_sample(self, seed=None):
        raise NotImplementedError

    def get_config(self):
        return {"name": self.name, "default": self.default}

    @classmethod
    def from_config(cls, config):
        return cls(**config)

    def __repr__(self):
        return "{}({})".format(type(self).__name__, ", ".join(
            f"{k}={v!r}" for k, v in self.get_config().items()))


def _check_sampling(sampling, step, min_value, max_value, hp_type="int"):
    if sampling is None:
        return None
    sampling = sampling.lower()
    if sampling not in ("linear", "log", "reverse_log"):
        raise ValueError('`sampling` must be one of "linear", "log", "reverse_log"')
    if sampling in ("log", "reverse_log") and min_value <= 0:
        raise ValueError('`sampling="{}"` is not supported for negative values, found '
                         "`min_value`: {}".format(sampling, min_value))
    return sampling


def _sample_range(rng, lo, hi, sampling):
    u = rng.random()
    if sampling == "log":
        return math.exp(math.log(lo) + u * (math.log(hi) - math.log(lo)))
    if sampling == "reverse_log":
        return hi + lo - math.exp(math.log(lo) + u * (math.log(hi) - math.log(lo)))
    return lo + u * (hi - lo)


class Choice(HyperParameter):
    def __init__(self, name, values, ordered=None, default=None):
        if not values:
            raise ValueError("`values` must be provided.")
        types = {type(v) for v in values}
        if len(types) > 1 and types != {int, float}:
            raise TypeError("A `Choice` can contain only one type of value, found values: "
                            + str(values) + " with types " + str(types))
        if not isinstance(values[0], (str, int, float, bool)):
            raise TypeError("A `Choice` can contain only `int`, `float`, `str`, or `bool`, found "
                            "values: " + str(values))
        if default is not None and default not in values:
            raise ValueError("The default value should be one of the choices. You passed: "
                             "values=%s, default=%s" % (values, default))
        super().__init__(name, default)
        self.values = list(values)
        is_numeric = isinstance(values[0], (int, float)) and not isinstance(values[0], bool)
        if ordered and not is_numeric:
            raise ValueError("`ordered` must be `False` for non-numeric types.")
        self.ordered = is_numeric if ordered is None else ordered

    @property
    def default(self):
        return self.values[0] if self._default is None else self._default

    def random_sample(self, seed=None):
        return random.Random(seed).choice(self.values)

    def get_config(self):
        return {"name": self.name, "default": self._default, "values": self.values,
                "ordered": self.ordered}


class Int(HyperParameter):
    def __init__(self, name, min_value, max_value, step=1, sampling=None, default=None):
        super().__init__(name, default)
        self.min_value, self.max_value = int(min_value), int(max_value)
        if self.min_value > self.max_value:
            raise ValueError("`min_value` cannot exceed `max_value`.")
        self.step = int(step) if step is not None else None
        self.sampling = _check_sampling(sampling, step, self.min_value, self.max_value)

    @property
    def default(self):
        return self.min_value if self._default is None else self._default

    def random_sample(self, seed=None):
        rng = random.Random(seed)
        v = _sample_range(rng, self.min_value, self.max_value + 1, self.sampling)
        v = int(min(max(math.floor(v), self.min_value), self.max_value))
        if self.step and self.step > 1:
            v = self.min_value + ((v - self.min_value) // self.step) * self.step
        return v

    def get_config(self):
        return {"name": self.name, "default": self._default, "min_value": self.min_value,
                "max_value": self.max_value, "step": self.step, "sampling": self.sampling}


class Float(HyperParameter):
    def __init__(self, name, min_value, max_value, step=None, sampling=None, default=None):
        super().__init__(name, default)
        self.min_value, self.max_value = float(min_value), float(max_value)
        if self.min_value > self.max_value:
            raise ValueError("`min_value` cannot exceed `max_value`.")
        self.step = float(step) if step is not None else None
        self.sampling = _check_sampling(sampling, step, self.min_value, self.max_value, "float")

    @property
    def default(self):
        return self.min_value if self._default is None else self._default

    def random_sample(self, seed=None):
        rng = random.Random(seed)
        v = _sample_range(rng, self.min_value, self.max_value, self.sampling)
        if self.step:
            v = self.min_value + round((v - self.min_value) / self.step) * self.step
        return min(max(v, self.min_value), self.max_value)

    def get_config(self):
        return {"name": self.name, "default": self._default, "min_value": self.min_value,
                "max_value": self.max_value, "step": self.step, "sampling": self.sampling}


class Boolean(HyperParameter):
    def __init__(self, name, default=False):
        if default not in (True, False):
            raise ValueError("`default` must be a Python boolean. You passed: default=%s" % default)
        super().__init__(name, default)

    def random_sample(self, seed=None):
        return random.Random(seed).choice((True, False))


class Fixed(HyperParameter):
    def __init__(self, name, value):
        if not isinstance(value, (int, float, str, bool)):
            raise ValueError("`Fixed` value must be an `int`, `float`, `str`, or `bool`, found {}"
                             .format(value))
        super().__init__(name, value)
        self.value = value

    def random_sample(self, seed=None):
        return self.value

    def get_config(self):
        return {"name": self.name, "value": self.value}


_TYPES = {c.__name__: c for c in (Choice, Int, Float, Boolean, Fixed)}


class HyperParameters:
    """Container for both the search space (``space``) and current values (``values``)."""

    def __init__(self):
        self._hps = {}          # name -> HyperParameter
        self._space = []
        self.values = {}
        self._scopes = []

    # -- registration / retrieval -----------------------------------------------------------------
    def _name(self, name):
        return "/".join(self._scopes + [name]) if self._scopes else name

    @contextlib.contextmanager
    def name_scope(self, name):
        self._scopes.append(name)
        try:
            yield
        finally:
            self._scopes.pop()

    @contextlib.contextmanager
    def conditional_scope(self, parent_name, parent_values):
        """Conditional scopes are accepted; the local oracle treats children as always active."""
        yield

    def _retrieve(self, hp):
        name = self._name(hp.name)
        hp.name = name
        if name in self._hps:
            if self.values is not None and name in self.values:
                return self.values[name]
            return self._hps[name].default
        return self.register(hp)

    def register(self, hp):
        self._hps[hp.name] = hp
        self._space.append(hp)
        if self.values is None:
            self.values = {}
        if hp.name not in self.values:
            self.values[hp.name] = hp.default
        return self.values[hp.name]

    def Choice(self, name, values, ordered=None, default=None, **kw):
        return self._retrieve(Choice(name, values, ordered, default))

    def Int(self, name, min_value, max_value, step=1, sampling=None, default=None, **kw):
        return self._retrieve(Int(name, min_value, max_value, step, sampling, default))

    def Float(self, name, min_value, max_value, step=None, sampling=None, default=None, **kw):
        return self._retrieve(Float(name, min_value, max_value, step, sampling, default))

    def Boolean(self, name, default=False, **kw):
        return self._retrieve(Boolean(name, default))

    def Fixed(self, name, value, **kw):
        return self._retrieve(Fixed(name, value))

    def get(self, name):
        name = self._name(name)
        if self.values is not None and name in self.values:
            return self.values[name]
        if name in self._hps:
            return self._hps[name].default
        raise KeyError("{} does not exist.".format(name))

    def __getitem__(self, name):
        return self.get(name)

    def __contains__(self, name):
        return self._name(name) in self._hps

    @property
    def space(self):
        return list(self._space)

    # -- (de)serialisation ------------------------------------------------------------------------
    def copy(self):
        return HyperParameters.from_config(self.get_config())

    def get_config(self):
        return {"space": [{"class_name": type(p).__name__, "config": p.get_config()}
                          for p in self._space],
                "values": dict(self.values) if self.values is not None else None}

    @classmethod
    def from_config(cls, config):
        hps = cls()
        for p in config["space"]:
            hp = _TYPES[p["class_name"]].from_config(copy.deepcopy(p["config"]))
            hps._hps[hp.name] = hp
            hps._space.append(hp)
        hps.values = dict(config["values"]) if config.get("values") is not None else {}
        return hps

    def merge(self, hps, overwrite=True):
        space = hps.space if isinstance(hps, HyperParameters) else hps
        for hp in space:
            if hp.name not in self._hps or overwrite:
                if hp.name in self._hps:
                    self._space = [p for p in self._space if p.name != hp.name]
                self._hps[hp.name] = hp
                self._space.append(hp)
                if self.values is not None:
                    self.values.setdefault(hp.name, hp.default)
