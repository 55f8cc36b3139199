"""Shared bits of cloud_fit (reference: experimental/cloud_fit/utils.py:17-39)."""
from ... import distribute

MULTI_WORKER_MIRRORED_STRATEGY_NAME = "MultiWorkerMirroredStrategy"
MIRRORED_STRATEGY_NAME = "MirroredStrategy"

SUPPORTED_DISTRIBUTION_STRATEGIES = {
    MULTI_WORKER_MIRRORED_STRATEGY_NAME: distribute.MultiWorkerMirroredStrategy,
    MIRRORED_STRATEGY_NAME: distribute.MirroredStrategy,
}


def is_tf_v1():
    """The reference refuses TF1 (client.py:95-98); there is no TF here at all."""
    return False


def enable_eager_for_tf_1():
    """No-op kept for API parity (torch is always eager)."""
