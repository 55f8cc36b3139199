"""The model zoo reproduces the reference programs' architectures: parameter counts are the ones
Keras reports for them (MLP 407 050, CNN 347 146, CTL-CNN 121 930,
ResNet-50 about 25.56 M)."""
import numpy as np
import pytest
import torch

import cloud_b200 as tfc
from cloud_b200 import models


@pytest.fixture(autouse=True)
def _cpu():
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    torch.manual_seed(0)


@pytest.mark.parametrize("name,kwargs,shape,params", [
    ("mnist_mlp", {}, (2, 784), 407050),
    ("mnist_cnn", {}, (2, 28, 28, 1), 347146),
    ("mnist_ctl_cnn", {}, (2, 28, 28, 1), 121930),
    ("tuner_mlp", {"num_layers": 3}, (2, 28, 28), 784 * 64 + 64 + 2 * (64 * 64 + 64) + 650),
    ("linear_probe", {}, (2, 1), 3),
])
def test_small_models(name, kwargs, shape, params):
    model = models.compiled(name, **kwargs)
    out = model.predict(np.random.rand(*shape).astype("float32"))
    assert out.shape[0] == 2 and model.count_params() == params


def test_resnet_classifiers():
    # Keras: ResNet50(weights=None) has 25 636 712 parameters with its 1000-class head, of which
    # 53 120 are batch-norm moving statistics
    m = models.get("resnet50", classes=1000, input_shape=(32, 32, 3))
    m.predict(np.random.randint(0, 256, (1, 32, 32, 3)).astype("uint8"))
    assert m.count_params() == 25636712
    assert sum(p.numel() for p in m.trainable_weights) == 25636712 - 53120
    frozen = models.resnet50_classifier(classes=120, input_shape=(32, 32, 3), freeze_base=True, dropout=0.5)
    frozen.predict(np.random.randint(0, 256, (1, 32, 32, 3)).astype("uint8"))
    assert sum(p.numel() for p in frozen.trainable_weights) == 2048 * 120 + 120
    r18 = models.get("resnet18")
    r18.predict(np.random.randint(0, 256, (1, 32, 32, 3)).astype("uint8"))
    assert 11.1e6 < r18.count_params() < 11.3e6


def test_registry():
    assert {"mnist_mlp", "resnet50", "resnet18"} <= set(models.list_models())
    with pytest.raises(ValueError, match="Unknown model"):
        models.get("alexnet")
    m = models.compiled("linear_probe")
    x = np.array([[9.], [10.], [11.]], dtype="float32")
    assert m.evaluate(x, x / 2 + 6, verbose=0) == pytest.approx(25.0025, rel=1e-5)


def test_more_applications_have_the_parameter_counts_keras_reports():
    apps = tfc.keras.applications
    x = np.random.rand(1, 32, 32, 3).astype("float32")
    for build, params in ((apps.ResNet101, 44707176), (apps.ResNet152, 60419944)):
        m = build(weights=None, input_shape=(32, 32, 3))
        assert m.predict(x).shape == (1, 1000) and m.count_params() == params
    vgg = apps.VGG16(weights=None, include_top=False, pooling="avg", input_shape=(32, 32, 3))
    assert vgg.predict(x).shape == (1, 512) and vgg.count_params() == 14714688
    mobile = apps.MobileNet(weights=None, input_shape=(32, 32, 3), classes=10, alpha=0.5)
    assert mobile.predict(x).shape == (1, 10)
    full = apps.MobileNet(weights=None, input_shape=(32, 32, 3))
    full.predict(x)
    assert 4.2e6 < full.count_params() < 4.3e6
    with pytest.raises(ValueError, match="no network"):
        apps.ResNet101(weights="imagenet")
