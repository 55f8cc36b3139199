"""``cloud_b200.keras`` - the Keras-shaped front end (``import cloud_b200.keras as keras``).

The reference's user scripts are written against ``tf.keras`` (README.md:66-81 and every file under
core/tests/testdata); this package provides that surface on torch + the B200 kernels so those
scripts run after changing only their import line.
"""
from . import layers, losses, metrics, optimizers, callbacks, applications, datasets, saving  # noqa: F401
from . import initializers, regularizers  # noqa: F401
from .layers import Input  # noqa: F401
from .tape import GradientTape  # noqa: F401
from .models import Model, Sequential, load_model, mixed_precision, model_from_json  # noqa: F401
from . import models  # noqa: F401


class backend:
    @staticmethod
    def clear_session():
        layers.reset_name_counts()

    @staticmethod
    def backend():
        return "cloud_b200"


class utils:
    from .batch_sources import Sequence  # noqa: F401

    @staticmethod
    def to_categorical(y, num_classes=None, dtype="float32"):
        import numpy as np
        y = np.asarray(y, dtype="int64").reshape(-1)
        n = num_classes or int(y.max()) + 1
        out = np.zeros((len(y), n), dtype=dtype)
        out[np.arange(len(y)), y] = 1
        return out

    @staticmethod
    def normalize(x, axis=-1, order=2):
        import numpy as np
        n = np.linalg.norm(x, order, axis, keepdims=True)
        n[n == 0] = 1
        return x / n
