"""cloud_b200 - a B200-native framework with the capabilities of TensorFlow Cloud.

``import cloud_b200 as tfc`` keeps the reference's spelling
(src/python/tensorflow_cloud/__init__.py:16-27): ``tfc.run``, ``tfc.remote``,
``tfc.MachineConfig``, ``tfc.AcceleratorType``, ``tfc.COMMON_MACHINE_CONFIGS``,
``tfc.CloudOracle``, ``tfc.CloudTuner`` (+ ``tfc.DistributingCloudTuner``), and adds the runtime
the reference outsources to TensorFlow: ``tfc.distribute`` (strategies + data-parallel engine on
hand-written sm_100a kernels), ``tfc.keras`` (Keras-shaped front end), ``tfc.data`` (tf.data-shaped
pipeline), ``tfc.datasets`` (tfds-shaped synthetic catalogue), ``tfc.image``.
"""
from .version import __version__  # noqa: F401

# Sub-packages load on first use (``tfc.keras``, ``from cloud_b200 import keras`` and
# ``import cloud_b200.keras`` all work): the launcher's supervisor, the job CLI and the client side of
# ``tfc.run`` never pay for importing torch.
_SUBMODULES = ("data", "datasets", "image", "distribute", "keras", "ops", "nn", "models", "monitoring",
               "tuner", "automl", "launch", "core", "experimental", "io", "parallel", "utils", "build")

_LAZY = {
    "GradientTape": ("cloud_b200.keras.tape", "GradientTape"),
    "run": ("cloud_b200.core.run", "run"),
    "remote": ("cloud_b200.core.run", "remote"),
    "MachineConfig": ("cloud_b200.core.machine_config", "MachineConfig"),
    "AcceleratorType": ("cloud_b200.core.machine_config", "AcceleratorType"),
    "COMMON_MACHINE_CONFIGS": ("cloud_b200.core.machine_config", "COMMON_MACHINE_CONFIGS"),
    "CloudOracle": ("cloud_b200.tuner.tuner", "CloudOracle"),
    "CloudTuner": ("cloud_b200.tuner.tuner", "CloudTuner"),
    "DistributingCloudTuner": ("cloud_b200.tuner.tuner", "DistributingCloudTuner"),
}


def __dir__():
    return sorted(list(globals()) + list(_LAZY) + list(_SUBMODULES))


def __getattr__(name):
    if name in _LAZY:
        import importlib
        mod, attr = _LAZY[name]
        value = getattr(importlib.import_module(mod), attr)
        globals()[name] = value
        return value
    if name in _SUBMODULES:
        import importlib
        return importlib.import_module("cloud_b200." + name)
    raise AttributeError(f"module 'cloud_b200' has no attribute {name!r}")
