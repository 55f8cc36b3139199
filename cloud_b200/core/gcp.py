"""Alias kept for code that imports ``<pkg>.core.gcp`` like the reference does
(src/python/tensorflow_cloud/core/gcp.py); the implementation lives in ``platform.py``."""
from .platform import *  # noqa: F401,F403
from .platform import _get_valid_machine_configurations  # noqa: F401
