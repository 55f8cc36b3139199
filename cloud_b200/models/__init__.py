"""Model zoo: the networks the reference's documentation, tests and examples train, as builders.

Every function returns an uncompiled ``cloud_b200.keras`` model with the layer stack (and therefore
the parameter count) of the reference program named in its docstring; ``compiled(name, ...)``
also applies that program's optimizer / loss / metrics.  ``bench.py``, ``bench/mlp.py`` and
``bench/tuner8.py`` time these architectures.
"""
from __future__ import annotations

from .. import keras

layers = keras.layers


def mnist_mlp(hidden=512, dropout=0.2, classes=10, input_shape=(784,)):
    """README MNIST MLP: 784-512-10, dropout 0.2, softmax (README.md:71-75): 407 050 parameters."""
    return keras.Sequential([
        layers.Dense(hidden, activation="relu", input_shape=tuple(input_shape)),
        layers.Dropout(dropout),
        layers.Dense(classes, activation="softmax"),
    ], name="mnist_mlp")


def mnist_cnn(classes=10, input_shape=(28, 28, 1)):
    """core/tests/testdata/mnist_example_using_fit.py:54-63: conv32 - pool - dense64 - softmax:
    347 146 parameters."""
    return keras.Sequential([
        layers.Conv2D(32, 3, activation="relu", input_shape=tuple(input_shape)),
        layers.MaxPooling2D(),
        layers.Flatten(),
        layers.Dense(64, activation="relu"),
        layers.Dense(classes, activation="softmax"),
    ], name="mnist_cnn")


def mnist_ctl_cnn(classes=10, input_shape=(28, 28, 1)):
    """core/tests/testdata/mnist_example_using_ctl.py:72-86: two conv + pool stages, dense64:
    121 930 parameters."""
    return keras.Sequential([
        layers.Conv2D(32, 3, activation="relu", input_shape=tuple(input_shape)),
        layers.MaxPooling2D(),
        layers.Conv2D(64, 3, activation="relu"),
        layers.MaxPooling2D(),
        layers.Flatten(),
        layers.Dense(64, activation="relu"),
        layers.Dense(classes, activation="softmax"),
    ], name="mnist_ctl_cnn")


def tuner_mlp(num_layers=2, units=64, classes=10, input_shape=(28, 28)):
    """The hypermodel of the tuner tests (tuner/tests/integration/tuner_integration_test.py:60-79):
    Flatten, ``num_layers`` x Dense(64, relu), softmax."""
    model = keras.Sequential(name="tuner_mlp")
    model.add(layers.Flatten(input_shape=tuple(input_shape)))
    for _ in range(int(num_layers)):
        model.add(layers.Dense(int(units), activation="relu"))
    model.add(layers.Dense(classes, activation="softmax"))
    return model


def linear_probe():
    """``y = w x + 1`` with ``w`` trainable and an L2 penalty: the model of the cloud_fit tests
    (experimental/cloud_fit/tests/unit/client_test.py:69-90)."""
    inp = layers.Input(shape=(1,), dtype="float32")
    times_w = layers.Dense(1, kernel_initializer=keras.initializers.Constant([[0.5]]),
                           kernel_regularizer=keras.regularizers.l2(0.01), use_bias=False)
    plus_1 = layers.Dense(1, kernel_initializer=keras.initializers.Constant([[1.0]]),
                          bias_initializer=keras.initializers.Constant([1.0]), trainable=False)
    return keras.Model(inp, plus_1(times_w(inp)), name="linear_probe")


def _classifier(base_fn, name, classes, input_shape, normalize, freeze_base, dropout, logits):
    inp = keras.Input(shape=tuple(input_shape))
    # uint8 NHWC images are normalised on the device, straight into the stem's operand layout
    x = layers.ImageNormalization(name="normalize")(inp) if normalize else inp
    base = base_fn(include_top=False, weights=None, input_tensor=x)
    if freeze_base:
        base.trainable = False
    y = layers.GlobalAveragePooling2D(name="avg_pool")(base.output)
    if dropout:
        y = layers.Dropout(dropout)(y)
    out = layers.Dense(classes, activation=None if logits else "softmax", name="predictions")(y)
    return keras.Model(inputs=inp, outputs=out, name=name)


def resnet50_classifier(classes=1000, input_shape=(224, 224, 3), normalize=True, freeze_base=False,
                        dropout=0.0, logits=False):
    """ResNet-50 + global average pooling + Dense head: the headline benchmark model, and with
    ``freeze_base=True, dropout=0.5, classes=120`` the transfer-learning model of
    core/tests/examples/call_run_within_script_with_keras_fit.py:79-87."""
    return _classifier(keras.applications.ResNet50, "resnet50_classifier", classes, input_shape,
                       normalize, freeze_base, dropout, logits)


def resnet18_classifier(classes=10, input_shape=(32, 32, 3), normalize=True, freeze_base=False,
                        dropout=0.0, logits=False):
    """ResNet-18 on small images: the second search space of the 8-trial tuner benchmark
    (BASELINE.json config 4)."""
    return _classifier(keras.applications.ResNet18, "resnet18_classifier", classes, input_shape,
                       normalize, freeze_base, dropout, logits)


_ZOO = {
    "mnist_mlp": (mnist_mlp, dict(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])),
    "mnist_cnn": (mnist_cnn, dict(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])),
    "mnist_ctl_cnn": (mnist_ctl_cnn, dict(optimizer="adam", loss="sparse_categorical_crossentropy",
                                          metrics=["accuracy"])),
    "tuner_mlp": (tuner_mlp, dict(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["acc"])),
    "linear_probe": (linear_probe, dict(optimizer=("sgd", 0.002), loss="mean_squared_error")),
    "resnet50": (resnet50_classifier, dict(optimizer=("adam", 1e-3), loss="sparse_categorical_crossentropy",
                                           metrics=["accuracy"])),
    "resnet18": (resnet18_classifier, dict(optimizer=("adam", 1e-3), loss="sparse_categorical_crossentropy",
                                           metrics=["accuracy"])),
}


def list_models():
    return sorted(_ZOO)


def get(name, **kwargs):
    """An uncompiled model of the zoo by name."""
    if name not in _ZOO:
        raise ValueError("Unknown model {!r}; available: {}".format(name, ", ".join(list_models())))
    return _ZOO[name][0](**kwargs)


def compiled(name, **kwargs):
    """The model compiled the way its reference program compiles it."""
    model = get(name, **kwargs)
    spec = dict(_ZOO[name][1])
    opt = spec.pop("optimizer")
    if isinstance(opt, tuple):
        opt = {"adam": keras.optimizers.Adam, "sgd": keras.optimizers.SGD}[opt[0]](opt[1])
    model.compile(optimizer=opt, **spec)
    return model
