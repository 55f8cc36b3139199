"""``keras.applications``-shaped model builders.

``ResNet50`` follows the Keras v1 graph (bias-carrying convolutions, stride on the first 1x1 of a
stage, BN epsilon 1.001e-5, layer names ``conv2_block1_1_conv`` ...) so weight files line up with
Keras'; it is the model of the reference's flagship example
(core/tests/examples/call_run_within_script_with_keras_fit.py:79-87) and of BASELINE.json's
headline config.  BN+ReLU and BN+Add+ReLU are single fused layers (one memory pass each on the
hand-written NHWC batch-norm kernels).  ``ResNet18`` is the basic-block variant used by the
tuner benchmark (BASELINE.json config 4).  ``ResNet101`` / ``ResNet152`` share ResNet50's blocks,
``VGG16`` and ``MobileNet`` (depthwise-separable blocks) round off the zoo; parameter counts equal Keras'
(tests/test_models_zoo.py).  Weights are always random-init (no network here).
"""
from __future__ import annotations

from . import layers as L
from .models import Model


def _bottleneck(x, filters, stride, conv_shortcut, name):
    if conv_shortcut:
        sc = L.Conv2D(4 * filters, 1, strides=stride, name=name + "_0_conv")(x)
        sc = L.BatchNormalization(epsilon=1.001e-5, name=name + "_0_bn")(sc)
    else:
        sc = x
    y = L.Conv2D(filters, 1, strides=stride, name=name + "_1_conv")(x)
    y = L.BatchNormalization(epsilon=1.001e-5, activation="relu", name=name + "_1_bn")(y)
    y = L.Conv2D(filters, 3, padding="same", name=name + "_2_conv")(y)
    y = L.BatchNormalization(epsilon=1.001e-5, activation="relu", name=name + "_2_bn")(y)
    y = L.Conv2D(4 * filters, 1, name=name + "_3_conv")(y)
    return L.BatchNormalization(epsilon=1.001e-5, activation="relu", name=name + "_3_bn")([y, sc])


def _basic(x, filters, stride, conv_shortcut, name):
    if conv_shortcut:
        sc = L.Conv2D(filters, 1, strides=stride, use_bias=False, name=name + "_0_conv")(x)
        sc = L.BatchNormalization(epsilon=1e-5, name=name + "_0_bn")(sc)
    else:
        sc = x
    y = x
    if stride > 1:
        y = L.Conv2D(filters, 3, strides=stride, padding=1, use_bias=False, name=name + "_1_conv")(y)
    else:
        y = L.Conv2D(filters, 3, padding="same", use_bias=False, name=name + "_1_conv")(y)
    y = L.BatchNormalization(epsilon=1e-5, activation="relu", name=name + "_1_bn")(y)
    y = L.Conv2D(filters, 3, padding="same", use_bias=False, name=name + "_2_conv")(y)
    return L.BatchNormalization(epsilon=1e-5, activation="relu", name=name + "_2_bn")([y, sc])


def _stem(inp, use_bias, eps):
    # Keras: conv1_pad (ZeroPadding2D 3) + 'valid' conv, pool1_pad (ZeroPadding2D 1) + 'valid' pool.
    # The paddings are folded into the conv / pool (same maths, two full activation copies saved).
    x = L.Conv2D(64, 7, strides=2, padding=3, use_bias=use_bias, name="conv1_conv")(inp)
    x = L.BatchNormalization(epsilon=eps, activation="relu", name="conv1_bn")(x)
    return L.MaxPooling2D(3, strides=2, padding=1, name="pool1_pool")(x)


def _head(x, include_top, pooling, classes, classifier_activation):
    if include_top:
        x = L.GlobalAveragePooling2D(name="avg_pool")(x)
        x = L.Dense(classes, activation=classifier_activation, name="predictions")(x)
    elif pooling == "avg":
        x = L.GlobalAveragePooling2D(name="avg_pool")(x)
    elif pooling == "max":
        x = L.GlobalMaxPooling2D(name="max_pool")(x)
    return x


def _check_weights(weights):
    if weights not in (None,):
        raise ValueError("Pretrained weights cannot be downloaded on this box (no network); "
                         "pass weights=None and load a local file with load_weights().")


def _resnet_v1(name, stack_blocks, include_top, weights, input_tensor, input_shape, pooling, classes,
               classifier_activation):
    _check_weights(weights)
    inp = input_tensor if input_tensor is not None else L.Input(shape=input_shape or (224, 224, 3))
    x = _stem(inp, True, 1.001e-5)
    for stage, (filters, blocks, stride) in enumerate(zip((64, 128, 256, 512), stack_blocks, (1, 2, 2, 2)), 2):
        for b in range(blocks):
            x = _bottleneck(x, filters, stride if b == 0 else 1, b == 0, f"conv{stage}_block{b + 1}")
    x = _head(x, include_top, pooling, classes, classifier_activation)
    return Model(inp, x, name=name)


def ResNet50(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
             classes=1000, classifier_activation="softmax", **kwargs):
    return _resnet_v1("resnet50", (3, 4, 6, 3), include_top, weights, input_tensor, input_shape, pooling,
                      classes, classifier_activation)


def ResNet101(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
              classes=1000, classifier_activation="softmax", **kwargs):
    """Keras ``ResNet101``: 44 707 176 parameters with the 1000-class head."""
    return _resnet_v1("resnet101", (3, 4, 23, 3), include_top, weights, input_tensor, input_shape, pooling,
                      classes, classifier_activation)


def ResNet152(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
              classes=1000, classifier_activation="softmax", **kwargs):
    """Keras ``ResNet152``: 60 419 944 parameters with the 1000-class head."""
    return _resnet_v1("resnet152", (3, 8, 36, 3), include_top, weights, input_tensor, input_shape, pooling,
                      classes, classifier_activation)


def VGG16(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
          classes=1000, classifier_activation="softmax", **kwargs):
    """Keras ``VGG16`` (layer names ``block1_conv1`` ... ``fc1``, ``fc2``, ``predictions``):
    138 357 544 parameters at 224 x 224 with the top."""
    _check_weights(weights)
    inp = input_tensor if input_tensor is not None else L.Input(shape=input_shape or (224, 224, 3))
    x = inp
    for block, (filters, convs) in enumerate(((64, 2), (128, 2), (256, 3), (512, 3), (512, 3)), 1):
        for c in range(convs):
            x = L.Conv2D(filters, 3, padding="same", activation="relu", name=f"block{block}_conv{c + 1}")(x)
        x = L.MaxPooling2D(2, strides=2, name=f"block{block}_pool")(x)
    if include_top:
        x = L.Flatten(name="flatten")(x)
        x = L.Dense(4096, activation="relu", name="fc1")(x)
        x = L.Dense(4096, activation="relu", name="fc2")(x)
        x = L.Dense(classes, activation=classifier_activation, name="predictions")(x)
    elif pooling == "avg":
        x = L.GlobalAveragePooling2D()(x)
    elif pooling == "max":
        x = L.GlobalMaxPooling2D()(x)
    return Model(inp, x, name="vgg16")


def MobileNet(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
              classes=1000, alpha=1.0, classifier_activation="softmax", **kwargs):
    """MobileNet v1 graph: a strided 3x3 stem, then thirteen depthwise-separable blocks
    (depthwise 3x3 + BN + ReLU, pointwise 1x1 + BN + ReLU), global average pooling, classifier.
    ``alpha`` scales every width.  4.2 M parameters at alpha = 1 with 1000 classes."""
    _check_weights(weights)
    inp = input_tensor if input_tensor is not None else L.Input(shape=input_shape or (224, 224, 3))

    def width(c):
        return max(8, int(c * alpha))
    x = L.Conv2D(width(32), 3, strides=2, padding="same", use_bias=False, name="conv1")(inp)
    x = L.BatchNormalization(activation="relu", name="conv1_bn")(x)
    plan = [(64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2)] + [(512, 1)] * 5 + [(1024, 2), (1024, 1)]
    for i, (filters, stride) in enumerate(plan, 1):
        x = L.DepthwiseConv2D(3, strides=stride, padding="same", use_bias=False, name=f"conv_dw_{i}")(x)
        x = L.BatchNormalization(activation="relu", name=f"conv_dw_{i}_bn")(x)
        x = L.Conv2D(width(filters), 1, use_bias=False, name=f"conv_pw_{i}")(x)
        x = L.BatchNormalization(activation="relu", name=f"conv_pw_{i}_bn")(x)
    x = _head(x, include_top, pooling, classes, classifier_activation)
    return Model(inp, x, name="mobilenet")


def ResNet18(include_top=True, weights=None, input_tensor=None, input_shape=None, pooling=None,
             classes=1000, classifier_activation="softmax", **kwargs):
    if weights not in (None,):
        raise ValueError("Pretrained weights cannot be downloaded on this box (no network).")
    inp = input_tensor if input_tensor is not None else L.Input(shape=input_shape or (224, 224, 3))
    x = _stem(inp, False, 1e-5)
    for stage, (filters, blocks, stride) in enumerate([(64, 2, 1), (128, 2, 2), (256, 2, 2), (512, 2, 2)], 2):
        for b in range(blocks):
            first = b == 0
            x = _basic(x, filters, stride if first else 1, first and (stride > 1),
                       f"conv{stage}_block{b + 1}")
    x = _head(x, include_top, pooling, classes, classifier_activation)
    return Model(inp, x, name="resnet18")


class resnet:  # tf.keras.applications.resnet.*
    ResNet50 = staticmethod(ResNet50)
    ResNet101 = staticmethod(ResNet101)
    ResNet152 = staticmethod(ResNet152)

    @staticmethod
    def preprocess_input(x, data_format=None):
        """caffe-style: RGB->BGR, subtract the ImageNet channel means (no scaling)."""
        import numpy as np
        x = np.asarray(x, dtype=np.float32)[..., ::-1]
        return x - np.array([103.939, 116.779, 123.68], dtype=np.float32)


class resnet50:
    ResNet50 = staticmethod(ResNet50)
    preprocess_input = resnet.preprocess_input
