"""Hyper-parameter tuning: ``CloudOracle`` / ``CloudTuner`` / ``DistributingCloudTuner`` backed by
a local Vizier-compatible optimizer, plus a multi-GPU trial scheduler
(reference: src/python/tensorflow_cloud/tuner)."""
from .tuner import CloudOracle, CloudTuner, DistributingCloudTuner  # noqa: F401
from .scheduler import TrialScheduler, WorkerPool  # noqa: F401
from .engine import HyperParameters, Objective  # noqa: F401
from .engine.tuner import BayesianOptimization, RandomSearch  # noqa: F401
from .engine.hyperband import Hyperband  # noqa: F401
