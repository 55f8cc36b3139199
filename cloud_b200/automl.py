"""``ImageClassifier`` / ``StructuredDataClassifier``: model search with one call, in the shape of
the AutoKeras API the reference demonstrates on top of ``tfc.run``
(core/tests/examples/call_run_within_script_with_autokeras.py:40-48)::

    clf = automl.ImageClassifier(max_trials=2)
    clf.fit(x_train, y_train, epochs=10)
    clf.evaluate(x_test, y_test)
    clf.export_model().save("model.h5")

AutoKeras itself is a separate library that is not on this box; this is a compact stand-in built on
the framework's own tuner engine: a fixed family of models (conv blocks -> global pooling -> dense
head for images, an MLP for feature vectors) whose depth, width, dropout and learning rate are the
search space, searched with ``RandomSearch`` (or any oracle-backed tuner passed in as ``tuner``),
each trial trained with the Keras front end on the B200 kernels.  Inputs are normalised inside the
model, so the exported model takes raw pixels / raw features like an AutoKeras export does.
"""
from __future__ import annotations

import tempfile

import numpy as np

from . import keras
from .tuner.engine.tuner import RandomSearch


class _AutoModel:
    _task = "classification"

    def __init__(self, num_classes=None, max_trials=100, directory=None, project_name="auto_model",
                 objective="val_loss", overwrite=False, seed=None, tuner=None, **kwargs):
        self.num_classes = num_classes
        self.max_trials = int(max_trials)
        self.directory = directory
        self.project_name = project_name
        self.objective = objective
        self.overwrite = overwrite
        self.seed = seed
        self._tuner_factory = tuner
        self.tuner = None
        self._best = None
        self._input_shape = None
        self._scale = 1.0

    # -- the search space -----------------------------------------------------------------------
    def _body(self, hp, x):
        raise NotImplementedError

    def _build(self, hp):
        inputs = keras.Input(shape=self._input_shape)
        x = keras.layers.Rescaling(self._scale)(inputs) if self._scale != 1.0 else inputs
        x = self._body(hp, x)
        drop = hp.Choice("dropout", [0.0, 0.25, 0.5], default=0.25)
        if drop:
            x = keras.layers.Dropout(drop)(x)
        outputs = keras.layers.Dense(self.num_classes, activation="softmax")(x)
        model = keras.Model(inputs, outputs)
        opt_name = hp.Choice("optimizer", ["adam", "sgd"], default="adam")
        lr = hp.Choice("learning_rate", [1e-1, 1e-2, 1e-3, 1e-4], default=1e-3)
        opt = keras.optimizers.Adam(lr) if opt_name == "adam" else keras.optimizers.SGD(lr, momentum=0.9)
        model.compile(optimizer=opt, loss="sparse_categorical_crossentropy", metrics=["accuracy"])
        return model

    # -- AutoKeras-facing -------------------------------------------------------------------------
    def _prepare(self, x, y):
        x = np.asarray(x)
        y = np.asarray(y).reshape(len(x), -1)
        if y.shape[1] > 1:                              # one-hot labels
            y = y.argmax(1)
        y = y.reshape(-1).astype(np.int64)
        return x, y

    def fit(self, x=None, y=None, epochs=None, callbacks=None, validation_split=0.2,
            validation_data=None, batch_size=32, verbose=1, **kwargs):
        x, y = self._prepare(x, y)
        self._infer_input(x)
        if self.num_classes is None:
            self.num_classes = int(y.max()) + 1
        if validation_data is None and validation_split:
            n_val = max(1, int(len(x) * validation_split))
            validation_data = (x[-n_val:], y[-n_val:])
            x, y = x[:-n_val], y[:-n_val]
        elif validation_data is not None:
            validation_data = self._prepare(*validation_data)
        directory = self.directory or tempfile.mkdtemp(prefix="automl_")
        if self._tuner_factory is not None:
            self.tuner = self._tuner_factory(self._build)
        else:
            self.tuner = RandomSearch(self._build, objective=self.objective, max_trials=self.max_trials,
                                      seed=self.seed, directory=directory, project_name=self.project_name,
                                      overwrite=self.overwrite)
        epochs = 10 if epochs is None else epochs       # AutoKeras trains up to 1000 with early stopping
        cbs = list(callbacks or [])
        if not any(isinstance(c, keras.callbacks.EarlyStopping) for c in cbs):
            cbs.append(keras.callbacks.EarlyStopping(monitor="val_loss", patience=max(2, epochs // 10)))
        self.tuner.search(x=x, y=y, epochs=epochs, validation_data=validation_data, batch_size=batch_size,
                          callbacks=cbs, verbose=2 if verbose else 0, **kwargs)
        self._best = self.tuner.get_best_models(1)[0]
        return self

    def _require_fit(self):
        if self._best is None:
            raise RuntimeError("Call fit() before {}.".format("predict/evaluate/export_model"))
        return self._best

    def predict(self, x, batch_size=32, **kwargs):
        probs = self._require_fit().predict(np.asarray(x), batch_size=batch_size)
        return probs.argmax(1).reshape(-1, 1)

    def evaluate(self, x, y=None, batch_size=32, verbose=0, **kwargs):
        x, y = self._prepare(x, y)
        return self._require_fit().evaluate(x, y, batch_size=batch_size, verbose=verbose)

    def export_model(self):
        """The best model found, as a plain Keras-shaped model (``.save()``, ``.predict()``, ...)."""
        return self._require_fit()


class ImageClassifier(_AutoModel):
    """Searches conv nets for images given as ``[N, H, W]`` or ``[N, H, W, C]`` arrays."""

    def _infer_input(self, x):
        self._expand = x.ndim == 3
        shape = x.shape[1:] + ((1,) if self._expand else ())
        self._input_shape = tuple(shape) if not self._expand else tuple(x.shape[1:])
        self._scale = 1.0 / 255.0 if x.dtype == np.uint8 or x.max() > 1.5 else 1.0

    def _body(self, hp, x):
        if self._expand:
            x = keras.layers.Reshape(self._input_shape + (1,))(x)
        blocks = hp.Int("conv_blocks", 1, 3, default=2)
        for b in range(blocks):
            filters = hp.Choice("filters_%d" % b, [16, 32, 64, 128], default=32)
            x = keras.layers.Conv2D(filters, 3, padding="same", activation="relu")(x)
            if hp.Boolean("second_conv_%d" % b, default=False):
                x = keras.layers.Conv2D(filters, 3, padding="same", activation="relu")(x)
            x = keras.layers.MaxPooling2D()(x)
        pooling = hp.Choice("spatial_reduction", ["flatten", "global_avg", "global_max"], default="flatten")
        if pooling == "flatten":
            x = keras.layers.Flatten()(x)
        elif pooling == "global_max":
            x = keras.layers.GlobalMaxPooling2D()(x)
        else:
            x = keras.layers.GlobalAveragePooling2D()(x)
        units = hp.Choice("dense_units", [32, 64, 128, 256], default=64)
        return keras.layers.Dense(units, activation="relu")(x)


class StructuredDataClassifier(_AutoModel):
    """Searches MLPs for feature vectors ``[N, F]``."""

    def _infer_input(self, x):
        self._input_shape = tuple(x.shape[1:])
        self._scale = 1.0

    def _body(self, hp, x):
        if len(self._input_shape) > 1:
            x = keras.layers.Flatten()(x)
        for i in range(hp.Int("num_layers", 1, 4, default=2)):
            x = keras.layers.Dense(hp.Choice("units_%d" % i, [16, 32, 64, 128, 256], default=64),
                                   activation="relu")(x)
        return x
