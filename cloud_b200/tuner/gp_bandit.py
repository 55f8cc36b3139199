"""Gaussian-process bandit for the local optimizer service.

The CAIP Optimizer's ``GAUSSIAN_PROCESS_BANDIT`` - also what ``ALGORITHM_UNSPECIFIED`` means to the
service, and ``make_study_config`` in the reference always asks for that (tuner/utils.py:58-68) -
runs inside Google's service, so the reference has no code for it; the wire contract only names it
(tuner/api/ml_public_google_rest_v1.json, ``StudyConfig.algorithm``).  This module is the local
stand-in: a plain numpy GP regression over the unit-cube encoding of the study's parameters with
an expected-improvement acquisition.

* encoding: DOUBLE / INTEGER -> one coordinate in [0, 1] through the parameter's ``scale_type``
  (linear, log, reverse log); DISCRETE -> its position between the smallest and the largest
  feasible value; CATEGORICAL -> one-hot.
* model: zero-mean GP on standardised objectives, Matern-5/2 kernel, one length scale picked from
  a small grid by log marginal likelihood, a noise term that is part of the same grid search.
* trials that are still running are "kriging believers": they enter the model with the posterior
  mean at their point, so eight workers asking at once are spread out instead of being handed
  eight neighbours of the same optimum.  Infeasible trials count as the worst value seen.
* acquisition: expected improvement over candidates drawn by the caller (uniform samples of the
  real search space plus perturbations of the best trials), so every suggestion is a feasible
  parameter assignment by construction.
"""
from __future__ import annotations

import math

import numpy as np

_LENGTH_SCALES = (0.08, 0.15, 0.3, 0.6, 1.2, 2.4)
_NOISES = (1e-6, 1e-3, 1e-2, 1e-1)
MAX_MODEL_POINTS = 256


def _value_spec(spec, snake, camel):
    return spec.get(snake) or spec.get(camel) or {}


def _bounds(vs):
    return float(vs.get("min_value", vs.get("minValue"))), float(vs.get("max_value", vs.get("maxValue")))


def param_value(entry):
    """The python value of one ``{"parameter", floatValue|intValue|stringValue}`` entry."""
    if "floatValue" in entry:
        return float(entry["floatValue"])
    if "intValue" in entry:
        return int(entry["intValue"])
    return entry.get("stringValue")


class SpaceEncoder:
    """Maps a trial's parameter list to a point of the unit cube."""

    def __init__(self, specs):
        self.specs = list(specs)
        self.columns = []           # (spec, first column, width)
        col = 0
        for spec in self.specs:
            width = 1
            if spec["type"] == "CATEGORICAL":
                vs = _value_spec(spec, "categorical_value_spec", "categoricalValueSpec")
                width = max(1, len(vs.get("values", [])))
            self.columns.append((spec, col, width))
            col += width
        self.dims = col

    def _unit(self, spec, v):
        kind = spec["type"]
        scale = spec.get("scale_type") or spec.get("scaleType")
        if kind in ("DOUBLE", "INTEGER"):
            vs = _value_spec(spec, "double_value_spec", "doubleValueSpec") or \
                _value_spec(spec, "integer_value_spec", "integerValueSpec")
            lo, hi = _bounds(vs)
            v = min(max(float(v), lo), hi)
            if hi <= lo:
                return 0.5
            if scale == "UNIT_LOG_SCALE" and lo > 0:
                return (math.log(v) - math.log(lo)) / (math.log(hi) - math.log(lo))
            if scale == "UNIT_REVERSE_LOG_SCALE" and lo > 0:
                return 1.0 - (math.log(hi + lo - v) - math.log(lo)) / (math.log(hi) - math.log(lo))
            return (v - lo) / (hi - lo)
        vs = _value_spec(spec, "discrete_value_spec", "discreteValueSpec")
        values = sorted(float(x) for x in vs.get("values", []))
        if len(values) < 2 or values[-1] == values[0]:
            return 0.5
        return (float(v) - values[0]) / (values[-1] - values[0])

    def encode(self, parameters) -> np.ndarray:
        by_name = {p["parameter"]: param_value(p) for p in parameters}
        x = np.zeros(self.dims)
        for spec, col, width in self.columns:
            v = by_name.get(spec["parameter"])
            if spec["type"] == "CATEGORICAL":
                values = [str(s) for s in
                          _value_spec(spec, "categorical_value_spec", "categoricalValueSpec").get("values", [])]
                if v is not None and str(v) in values:
                    x[col + values.index(str(v))] = 1.0
                continue
            x[col] = 0.5 if v is None else self._unit(spec, v)
        return x


def _matern52(a: np.ndarray, b: np.ndarray, length: float) -> np.ndarray:
    d2 = np.maximum(((a[:, None, :] - b[None, :, :]) ** 2).sum(-1), 0.0)
    r = np.sqrt(5.0 * d2) / length
    return (1.0 + r + r * r / 3.0) * np.exp(-r)


class GaussianProcess:
    """Zero-mean GP with a Matern-5/2 kernel on already standardised targets."""

    def __init__(self, x: np.ndarray, y: np.ndarray):
        self.x = x
        best = None
        for length in _LENGTH_SCALES:
            k0 = _matern52(x, x, length * math.sqrt(max(1, x.shape[1])))
            for noise in _NOISES:
                try:
                    chol = np.linalg.cholesky(k0 + noise * np.eye(len(x)))
                except np.linalg.LinAlgError:
                    continue
                alpha = np.linalg.solve(chol.T, np.linalg.solve(chol, y))
                lml = -0.5 * float(y @ alpha) - float(np.log(np.diag(chol)).sum())
                if best is None or lml > best[0]:
                    best = (lml, length, noise, chol, alpha)
        if best is None:
            raise np.linalg.LinAlgError("no positive-definite kernel matrix")
        self.log_marginal_likelihood, self.length, self.noise, self._chol, self._alpha = best

    def predict(self, q: np.ndarray):
        ks = _matern52(q, self.x, self.length * math.sqrt(max(1, self.x.shape[1])))
        mean = ks @ self._alpha
        v = np.linalg.solve(self._chol, ks.T)
        var = np.maximum(1.0 - (v * v).sum(0), 1e-12)
        return mean, np.sqrt(var)


def _norm_pdf(z):
    return np.exp(-0.5 * z * z) / math.sqrt(2.0 * math.pi)


try:
    from scipy.special import erf as _erf
except ImportError:                     # scipy is optional
    _erf = np.vectorize(math.erf)


def _norm_cdf(z):
    return 0.5 * (1.0 + _erf(z / math.sqrt(2.0)))


def expected_improvement(mean, std, incumbent, xi=0.01):
    """EI for maximisation of a standardised objective."""
    imp = mean - incumbent - xi
    z = imp / std
    return imp * _norm_cdf(z) + std * _norm_pdf(z)


def min_initial_trials(dims: int) -> int:
    """Completed trials wanted before the model replaces random exploration."""
    return min(max(3, dims + 1), 10)


def choose(specs, observed, pending, candidates, maximize: bool):
    """Index of the candidate with the largest expected improvement.

    ``observed``: list of (parameters, objective or None for infeasible) of finished trials;
    ``pending``: parameter lists of trials still running; ``candidates``: parameter lists.
    Returns ``None`` when there is not enough signal for a model (caller falls back to random).
    """
    enc = SpaceEncoder(specs)
    feasible = [v for _p, v in observed if v is not None and math.isfinite(v)]
    if enc.dims == 0 or len(feasible) < min_initial_trials(enc.dims) or not candidates:
        return None
    sign = 1.0 if maximize else -1.0
    worst = min(sign * v for v in feasible)
    xs, ys = [], []
    for params, value in observed:
        xs.append(enc.encode(params))
        ys.append(sign * value if value is not None and math.isfinite(value) else worst)
    if len(xs) > MAX_MODEL_POINTS:      # keep the best half and the most recent half
        order = np.argsort(ys)[::-1][:MAX_MODEL_POINTS // 2].tolist()
        recent = [i for i in range(len(xs) - 1, -1, -1) if i not in set(order)][:MAX_MODEL_POINTS // 2]
        keep = sorted(order + recent)
        xs, ys = [xs[i] for i in keep], [ys[i] for i in keep]
    x = np.stack(xs)
    y = np.asarray(ys, dtype=np.float64)
    mu, sd = float(y.mean()), float(y.std())
    if sd < 1e-12:
        return None                     # flat objective so far: nothing to model
    y = (y - mu) / sd
    try:
        gp = GaussianProcess(x, y)
        if pending:
            xp = np.stack([enc.encode(p) for p in pending])
            believed, _ = gp.predict(xp)
            gp = GaussianProcess(np.concatenate([x, xp]), np.concatenate([y, believed]))
        q = np.stack([enc.encode(c) for c in candidates])
        mean, std = gp.predict(q)
    except np.linalg.LinAlgError:
        return None
    ei = expected_improvement(mean, std, float(y.max()))
    return int(np.argmax(ei))
