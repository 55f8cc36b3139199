"""The pieces around the Keras front end that the reference's examples import from other libraries:
``tensorflow_datasets`` (tfds.load in testdata/mnist_example_using_fit.py:41-47 and the notebooks),
``tf.image.resize`` (call_run_within_script_with_keras_fit.py:60-64), AutoKeras' ``ImageClassifier``
(call_run_within_script_with_autokeras.py:40-48), and a kernel to run notebooks in."""
import json
import os

import numpy as np
import pytest
import torch

import cloud_b200 as tfc
from cloud_b200 import automl
from cloud_b200.utils import notebook

tfds = tfc.datasets


@pytest.fixture(autouse=True)
def _cpu(monkeypatch):
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    monkeypatch.setenv("TFC_SYNTHETIC_SAMPLES", "300")
    torch.manual_seed(0)


def test_load_returns_the_tfds_shapes():
    datasets, info = tfds.load(name="mnist", with_info=True, as_supervised=True)
    assert set(datasets) == {"train", "test"}
    assert info.splits["train"].num_examples == 300 and info.features["label"].num_classes == 10
    image, label = next(iter(datasets["train"]))
    assert image.shape == (28, 28, 1) and image.dtype == np.uint8 and 0 <= int(label) < 10
    (train, test), meta = tfds.load("stanford_dogs", split=["train", "test"], shuffle_files=True,
                                    with_info=True, as_supervised=True)
    assert meta.features["label"].num_classes == 120
    assert tfc.data.experimental.cardinality(train) == 300 == len(test)
    features = next(iter(tfds.load("cifar10", split="test")))
    assert set(features) == {"image", "label"} and features["image"].shape == (32, 32, 3)
    assert len(tfds.load("mnist", split="train[:50%]")) == 150
    assert len(tfds.load("mnist", split="train[100:]")) == 200
    batch = next(iter(tfds.load("mnist", split="train", batch_size=32, as_supervised=True)))
    assert batch[0].shape == (32, 28, 28, 1)
    with pytest.raises(ValueError, match="not in the local catalogue"):
        tfds.load("imagenet2012")
    with pytest.raises(ValueError, match="Unknown split"):
        tfds.load("mnist", split="validation")
    assert "mnist" in tfds.list_builders()
    # deterministic: the same split is the same data in every process
    again = next(iter(tfds.load("mnist", split="train", as_supervised=True)))
    assert np.array_equal(again[0], image)


def test_synthetic_data_is_learnable():
    train = tfds.load("mnist", split="train", as_supervised=True)
    train = train.map(lambda im, lb: (im.astype("float32") / 255.0, lb)).batch(50)
    model = tfc.keras.Sequential([tfc.keras.layers.Flatten(input_shape=(28, 28, 1)),
                                  tfc.keras.layers.Dense(10, activation="softmax")])
    model.compile(tfc.keras.optimizers.Adam(1e-2), "sparse_categorical_crossentropy", metrics=["accuracy"])
    hist = model.fit(train, epochs=4, verbose=0)
    assert hist.history["accuracy"][-1] > 0.9


def test_resize_matches_torch_interpolate():
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, size=(2, 9, 7, 3)).astype(np.uint8)
    for size in ((18, 14), (5, 4), (9, 7), (224, 224)):
        got = tfc.image.resize(img, size)
        ref = torch.nn.functional.interpolate(torch.tensor(img).permute(0, 3, 1, 2).float(), size=size,
                                              mode="bilinear", align_corners=False, antialias=False)
        np.testing.assert_allclose(got, ref.permute(0, 2, 3, 1).numpy(), atol=1e-3)
        assert got.dtype == np.float32
    single = tfc.image.resize(img[0], (4, 4), method="nearest")
    assert single.shape == (4, 4, 3) and set(np.unique(single)) <= set(np.unique(img[0]).astype(np.float32))
    with pytest.raises(ValueError):
        tfc.image.resize(np.zeros((4, 4)), (2, 2))
    assert tfc.image.convert_image_dtype(img, "float32").max() <= 1.0


def test_image_classifier_searches_trains_and_exports(tmp_path):
    (x, y), (xt, yt) = tfc.keras.datasets.mnist.load_data()
    x, y, xt, yt = x[:400], y[:400], xt[:100], yt[:100]
    clf = automl.ImageClassifier(max_trials=2, seed=0, directory=str(tmp_path / "search"))
    with pytest.raises(RuntimeError, match="fit"):
        clf.export_model()
    clf.fit(x, y, epochs=3, verbose=0)
    assert len(clf.tuner.oracle.trials) == 2
    loss, acc = clf.evaluate(xt, yt)
    assert acc > 0.2 and np.isfinite(loss)              # chance is 0.1
    pred = clf.predict(xt[:7])
    assert pred.shape == (7, 1) and pred.dtype.kind == "i"
    # the export takes raw uint8 pixels (the scaling is part of the model) and reloads from disk
    path = str(tmp_path / "model.h5")
    clf.export_model().save(path)
    again = tfc.keras.models.load_model(path)
    np.testing.assert_allclose(again.predict(xt[:7]), clf.export_model().predict(xt[:7]), atol=1e-5)


def test_structured_data_classifier_on_feature_vectors(tmp_path):
    rng = np.random.RandomState(0)
    centers = rng.randn(3, 12) * 3
    y = rng.randint(0, 3, 300)
    x = (centers[y] + rng.randn(300, 12)).astype("float32")
    onehot = tfc.keras.utils.to_categorical(y, 3)                       # one-hot labels are accepted
    clf = automl.StructuredDataClassifier(max_trials=2, seed=1, directory=str(tmp_path / "s"))
    clf.fit(x, onehot, epochs=5, verbose=0, validation_data=(x[:60], y[:60]))
    assert clf.evaluate(x, y)[1] > 0.8


def _write_notebook(path, cells):
    nb = {"cells": [{"cell_type": kind, "metadata": {}, "source": src.splitlines(keepends=True)}
                    for kind, src in cells], "metadata": {}, "nbformat": 4, "nbformat_minor": 4}
    path.write_text(json.dumps(nb))


def test_notebook_runner_is_a_run_all(tmp_path, monkeypatch):
    nb = tmp_path / "demo.ipynb"
    _write_notebook(nb, [
        ("markdown", "# not code"),
        ("code", "!pip install something\n%matplotlib inline\nimport os\nvalues = [1, 2]\n"),
        ("code", "if values:\n    !echo shell lines inside a block keep the block valid\nvalues.append(3)\n"),
        ("code", "from cloud_b200.core import run as run_module\n"
                 "in_notebook = run_module._called_from_notebook()\n"
                 "nb_path = os.environ['TFC_NOTEBOOK_PATH']\n"),
    ])
    monkeypatch.delenv("TFC_NOTEBOOK_PATH", raising=False)
    ns = notebook.run_notebook(str(nb))
    assert ns["values"] == [1, 2, 3] and ns["in_notebook"] is True and ns["nb_path"] == str(nb)
    assert "TFC_NOTEBOOK_PATH" not in os.environ                    # restored afterwards
    assert len(notebook.code_cells(str(nb))) == 3
    bad = tmp_path / "bad.ipynb"
    _write_notebook(bad, [("code", "x = 1\n"), ("code", "raise ValueError('cell two')\n")])
    with pytest.raises(ValueError, match="cell two"):
        notebook.run_notebook(str(bad))
