"""A compact re-creation of the KerasTuner 1.0.x engine the reference builds on (SURVEY.md appendix
B2): ``hyperparameters``, ``trial``, ``metrics_tracking``, ``oracle``, ``hypermodel``, ``tuner``."""
from . import hyperparameters, trial, metrics_tracking, oracle, hypermodel, tuner  # noqa: F401
from .hyperparameters import HyperParameters  # noqa: F401
from .oracle import Objective, Oracle  # noqa: F401
from .trial import Trial, TrialStatus  # noqa: F401
from .hypermodel import HyperModel, DefaultHyperModel  # noqa: F401
from .tuner import Tuner  # noqa: F401
