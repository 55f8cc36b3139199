"""Version helpers (reference: src/python/tensorflow_cloud/utils/tf_utils.py:16-20 returns
``tf.__version__``; the runtime here is cloud_b200 on torch)."""
from ..version import __version__


def get_version():
    return __version__


def get_torch_version():
    import torch
    return torch.__version__
