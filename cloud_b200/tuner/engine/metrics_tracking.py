"""Per-trial metric histories and metric-direction inference (KerasTuner ``metrics_tracking``;
``infer_metric_direction`` is relied on by tuner/utils.py:285-323 and its test
tuner/tests/unit/utils_test.py:249-262)."""
from __future__ import annotations


class MetricObservation:
    def __init__(self, value, step):
        self.value = float(value)
        self.step = int(step)

    def get_config(self):
        return {"value": self.value, "step": self.step}


class MetricHistory:
    def __init__(self, direction="min"):
        if direction not in ("min", "max"):
            raise ValueError("`direction` should be one of {'min', 'max'}, but got: %s" % direction)
        self.direction = direction
        self._observations = {}

    def update(self, value, step):
        self._observations[int(step)] = MetricObservation(value, step)

    def get_history(self):
        return [self._observations[k] for k in sorted(self._observations)]

    def get_best_value(self):
        vals = [o.value for o in self._observations.values()]
        if not vals:
            return None
        return min(vals) if self.direction == "min" else max(vals)

    def get_best_step(self):
        if not self._observations:
            return None
        pick = min if self.direction == "min" else max
        return pick(self._observations.values(), key=lambda o: o.value).step

    def get_last_value(self):
        hist = self.get_history()
        return hist[-1].value if hist else None

    def get_config(self):
        return {"direction": self.direction, "observations": [o.get_config() for o in self.get_history()]}

    @classmethod
    def from_config(cls, config):
        h = cls(config["direction"])
        for o in config["observations"]:
            h.update(o["value"], o["step"])
        return h


class MetricsTracker:
    def __init__(self, metrics=None):
        self.metrics = {}
        for m in metrics or []:
            name = m if isinstance(m, str) else getattr(m, "name", str(m))
            self.register(name)

    def exists(self, name):
        return name in self.metrics

    def register(self, name, direction=None):
        if direction is None:
            direction = infer_metric_direction(name) or "min"
        self.metrics[name] = MetricHistory(direction)

    def update(self, name, value, step=0):
        if not self.exists(name):
            self.register(name)
        prev_best = self.metrics[name].get_best_value()
        self.metrics[name].update(value, step)
        new_best = self.metrics[name].get_best_value()
        return prev_best is None or new_best != prev_best

    def get_history(self, name):
        return self.metrics[name].get_history()

    def get_best_value(self, name):
        return self.metrics[name].get_best_value()

    def get_best_step(self, name):
        return self.metrics[name].get_best_step()

    def get_last_value(self, name):
        return self.metrics[name].get_last_value()

    def get_config(self):
        return {"metrics": {k: v.get_config() for k, v in self.metrics.items()}}

    @classmethod
    def from_config(cls, config):
        t = cls()
        t.metrics = {k: MetricHistory.from_config(v) for k, v in config["metrics"].items()}
        return t


_MAX_NAMES = ("acc", "accuracy", "auc", "precision", "recall", "f1", "iou", "cosine_similarity",
              "cosine_proximity", "true_positives", "true_negatives")
_MIN_NAMES = ("loss", "error", "mse", "mae", "mape", "msle", "rmse", "hinge", "logcosh",
              "kullback_leibler_divergence", "kld", "poisson", "crossentropy", "false_positives",
              "false_negatives")


def infer_metric_direction(metric):
    """'max' for accuracy-like names, 'min' for loss/error-like names, else None."""
    name = metric if isinstance(metric, str) else getattr(metric, "name", str(metric))
    name = name.lower()
    if name.startswith("val_"):
        name = name[4:]
    if name in ("acc", "accuracy") or name.endswith("accuracy") or name.endswith("_acc"):
        return "max"
    if name == "loss" or name.endswith("loss"):
        return "min"
    for m in _MAX_NAMES:
        if m in name:
            return "max"
    for m in _MIN_NAMES:
        if m in name:
            return "min"
    return None
