"""``Trial`` and ``TrialStatus`` (KerasTuner engine; used by tuner/tuner.py:121-322)."""
from __future__ import annotations

import hashlib
import json
import os
import random
import time

from . import hyperparameters as hp_module
from . import metrics_tracking


class TrialStatus:
    RUNNING = "RUNNING"
    IDLE = "IDLE"
    INVALID = "INVALID"
    STOPPED = "STOPPED"
    COMPLETED = "COMPLETED"


def generate_trial_id():
    s = str(time.time()) + str(random.randint(1, int(1e7)))
    return hashlib.sha256(s.encode("utf-8")).hexdigest()[:32]


class Trial:
    def __init__(self, hyperparameters, trial_id=None, status=TrialStatus.RUNNING):
        self.hyperparameters = hyperparameters
        self.trial_id = generate_trial_id() if trial_id is None else trial_id
        self.metrics = metrics_tracking.MetricsTracker()
        self.score = None
        self.best_step = None
        self.status = status

    def summary(self):
        print("Trial summary")
        print("|-Trial ID: {}".format(self.trial_id))
        print("|-Score: {}".format(self.score))
        print("|-Best step: {}".format(self.best_step))
        print("Hyperparameters:")
        values = getattr(self.hyperparameters, "values", None) or {}
        for name, value in values.items():
            print("|-{}: {}".format(name, value))

    def get_state(self):
        hps = self.hyperparameters
        return {"trial_id": self.trial_id,
                "hyperparameters": hps.get_config() if hasattr(hps, "get_config") else None,
                "metrics": self.metrics.get_config(), "score": self.score,
                "best_step": self.best_step, "status": self.status}

    def set_state(self, state):
        self.trial_id = state["trial_id"]
        if state.get("hyperparameters"):
            self.hyperparameters = hp_module.HyperParameters.from_config(state["hyperparameters"])
        self.metrics = metrics_tracking.MetricsTracker.from_config(state["metrics"])
        self.score = state["score"]
        self.best_step = state["best_step"]
        self.status = state["status"]

    @classmethod
    def from_state(cls, state):
        t = cls(hyperparameters=None)
        t.set_state(state)
        return t

    def save(self, fname):
        os.makedirs(os.path.dirname(fname), exist_ok=True)
        tmp = fname + ".tmp"
        with open(tmp, "w") as f:
            json.dump(self.get_state(), f)
        os.replace(tmp, fname)

    @classmethod
    def load(cls, fname):
        with open(fname) as f:
            return cls.from_state(json.load(f))
