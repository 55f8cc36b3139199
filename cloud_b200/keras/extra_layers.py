"""Layers beyond the ones the reference's own examples use (those are in ``layers.py``): what a
script written against ``tf.keras`` commonly contains besides them.  TensorFlow Cloud runs any Keras
model (``README.md:55-108`` - "your ``tf.keras`` model"), so a user switching over finds these too.
They run on PyTorch ops (cuDNN / cuBLAS on the GPU) - library rows, not hand-written kernels - and
keep the Keras weight names and layouts, so HDF5 weight files stay interchangeable:

    Embedding ``embeddings[vocab, dim]``; LayerNormalization ``gamma, beta``;
    Conv1D ``kernel[width, in, out]``; DepthwiseConv2D ``depthwise_kernel[h, w, in, mult]``;
    SeparableConv2D ``depthwise_kernel``, ``pointwise_kernel[1, 1, in * mult, out]``, ``bias``.

Tensor layout inside a model: 4-D activations are NCHW-logical / channels_last (see ``layers.py``),
3-D ones stay ``[batch, steps, channels]`` as in Keras.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .layers import Layer, _init_tensor, _pair, _same_pad, get_activation


# -- activations as layers ---------------------------------------------------------------------------
class LeakyReLU(Layer):
    def __init__(self, alpha=0.3, **kwargs):
        super().__init__(**kwargs)
        self.alpha = float(alpha)

    def call(self, x, training=False):
        return F.leaky_relu(x, self.alpha)

    def get_config(self):
        return {**super().get_config(), "alpha": self.alpha}


class ELU(Layer):
    def __init__(self, alpha=1.0, **kwargs):
        super().__init__(**kwargs)
        self.alpha = float(alpha)

    def call(self, x, training=False):
        return F.elu(x, self.alpha)

    def get_config(self):
        return {**super().get_config(), "alpha": self.alpha}


class PReLU(Layer):
    """``max(x, 0) + alpha * min(x, 0)`` with one learned ``alpha`` per channel (``shared_axes``
    = the spatial axes, the usual setting for convolutional nets, is what is implemented)."""

    def build(self, input_shape):
        self.alpha = nn.Parameter(torch.zeros(int(input_shape[-1])))
        self.add_weight_spec("alpha", "alpha")

    def call(self, x, training=False):
        a = self.alpha.to(x.dtype)
        if x.dim() == 4:
            a = a.view(1, -1, 1, 1)
        return torch.relu(x) + a * torch.clamp(x, max=0)


# -- merges ------------------------------------------------------------------------------------------
class _Merge(Layer):
    def compute_output_shape(self, input_shape):
        return tuple(input_shape[0])


class Subtract(_Merge):
    def call(self, xs, training=False):
        return xs[0] - xs[1]


class Multiply(_Merge):
    def call(self, xs, training=False):
        y = xs[0]
        for t in xs[1:]:
            y = y * t
        return y


class Average(_Merge):
    def call(self, xs, training=False):
        return sum(xs[1:], xs[0]) / float(len(xs))


class Maximum(_Merge):
    def call(self, xs, training=False):
        y = xs[0]
        for t in xs[1:]:
            y = torch.maximum(y, t)
        return y


# -- normalisation / embedding -------------------------------------------------------------------------
class LayerNormalization(Layer):
    """Normalises over the last (channel) axis."""

    def __init__(self, axis=-1, epsilon=1e-3, center=True, scale=True, **kwargs):
        super().__init__(**kwargs)
        if axis not in (-1,):
            raise NotImplementedError("LayerNormalization over the last axis only")
        self.epsilon, self.center, self.scale = float(epsilon), center, scale

    def build(self, input_shape):
        c = int(input_shape[-1])
        self.gamma = nn.Parameter(torch.ones(c), requires_grad=self.scale)
        self.beta = nn.Parameter(torch.zeros(c), requires_grad=self.center)
        self.add_weight_spec("gamma", "gamma")
        self.add_weight_spec("beta", "beta")

    def call(self, x, training=False):
        if x.dim() == 4:                       # channels are axis 1 internally
            y = F.layer_norm(x.permute(0, 2, 3, 1).float(), (x.shape[1],), self.gamma, self.beta, self.epsilon)
            return y.permute(0, 3, 1, 2).to(x.dtype)
        return F.layer_norm(x.float(), (x.shape[-1],), self.gamma, self.beta, self.epsilon).to(x.dtype)

    def get_config(self):
        return {**super().get_config(), "epsilon": self.epsilon, "center": self.center, "scale": self.scale}


class Embedding(Layer):
    def __init__(self, input_dim, output_dim, embeddings_initializer="random_uniform", mask_zero=False,
                 input_length=None, **kwargs):
        if input_length is not None and "input_shape" not in kwargs:
            kwargs["input_shape"] = (input_length,)
        super().__init__(**kwargs)
        self.input_dim, self.output_dim = int(input_dim), int(output_dim)
        self.embeddings_initializer = embeddings_initializer

    def build(self, input_shape):
        w = _init_tensor(self.embeddings_initializer, (self.input_dim, self.output_dim), self.input_dim,
                         self.output_dim)
        self.embeddings = nn.Parameter(w)
        self.add_weight_spec("embeddings", "embeddings")

    def compute_output_shape(self, input_shape):
        return tuple(input_shape) + (self.output_dim,)

    def call(self, x, training=False):
        return F.embedding(x.long(), self.embeddings)

    def get_config(self):
        return {**super().get_config(), "input_dim": self.input_dim, "output_dim": self.output_dim}


# -- 1-D convolution / pooling -------------------------------------------------------------------------
class Conv1D(Layer):
    def __init__(self, filters, kernel_size, strides=1, padding="valid", activation=None, use_bias=True,
                 dilation_rate=1, kernel_initializer="glorot_uniform", bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.filters, self.kernel_size, self.strides = int(filters), int(kernel_size), int(strides)
        self.padding = padding.lower()
        if self.padding not in ("valid", "same", "causal"):
            raise ValueError("padding must be 'valid', 'same' or 'causal'")
        self.dilation = int(dilation_rate)
        self.activation_name = activation if isinstance(activation, (str, type(None))) else None
        self._activation_fn = get_activation(activation)
        self.use_bias = use_bias
        self.kernel_initializer, self.bias_initializer = kernel_initializer, bias_initializer

    def build(self, input_shape):
        cin = int(input_shape[-1])
        fan_in, fan_out = cin * self.kernel_size, self.filters * self.kernel_size
        w = _init_tensor(self.kernel_initializer, (self.kernel_size, cin, self.filters), fan_in, fan_out)
        self.kernel = nn.Parameter(w.permute(2, 1, 0).contiguous())          # torch: [out, in, width]
        self.add_weight_spec("kernel", "kernel", lambda t: t.permute(2, 1, 0), lambda t: t.permute(2, 1, 0))
        if self.use_bias:
            self.bias = nn.Parameter(_init_tensor(self.bias_initializer, (self.filters,), fan_in, fan_out))
            self.add_weight_spec("bias", "bias")
        else:
            self.bias = None

    def _pads(self, steps):
        eff = (self.kernel_size - 1) * self.dilation
        if self.padding == "causal":
            return eff, 0
        if self.padding == "same":
            return _same_pad(steps, self.kernel_size, self.strides, self.dilation)
        return 0, 0

    def compute_output_shape(self, input_shape):
        b, steps, _c = input_shape
        if steps is not None:
            lo, hi = self._pads(steps)
            steps = (steps + lo + hi - ((self.kernel_size - 1) * self.dilation + 1)) // self.strides + 1
        return (b, steps, self.filters)

    def call(self, x, training=False):
        x = x.transpose(1, 2)                                                # [B, C, T]
        lo, hi = self._pads(x.shape[-1])
        if lo or hi:
            x = F.pad(x, (lo, hi))
        w = self.kernel
        if w.dtype != x.dtype:
            x = x.to(w.dtype)
        y = F.conv1d(x, w, self.bias, stride=self.strides, dilation=self.dilation).transpose(1, 2)
        return y if self._activation_fn is None else self._activation_fn(y)

    def get_config(self):
        return {**super().get_config(), "filters": self.filters, "kernel_size": self.kernel_size,
                "strides": self.strides, "padding": self.padding, "activation": self.activation_name,
                "use_bias": self.use_bias, "dilation_rate": self.dilation}


class _Pool1D(Layer):
    def __init__(self, pool_size=2, strides=None, padding="valid", **kwargs):
        super().__init__(**kwargs)
        self.pool_size = int(pool_size)
        self.strides = int(strides) if strides is not None else self.pool_size
        self.padding = padding.lower()

    def compute_output_shape(self, input_shape):
        b, steps, c = input_shape
        if steps is not None:
            steps = -(-steps // self.strides) if self.padding == "same" else (steps - self.pool_size) // self.strides + 1
        return (b, steps, c)

    def _prepare(self, x, fill):
        x = x.transpose(1, 2)
        if self.padding == "same":
            lo, hi = _same_pad(x.shape[-1], self.pool_size, self.strides)
            if lo or hi:
                x = F.pad(x, (lo, hi), value=fill)
        return x

    def get_config(self):
        return {**super().get_config(), "pool_size": self.pool_size, "strides": self.strides,
                "padding": self.padding}


class MaxPooling1D(_Pool1D):
    def call(self, x, training=False):
        return F.max_pool1d(self._prepare(x, float("-inf")), self.pool_size, self.strides).transpose(1, 2)


class AveragePooling1D(_Pool1D):
    def call(self, x, training=False):
        if self.padding == "same":
            raise NotImplementedError("AveragePooling1D(padding='same')")
        return F.avg_pool1d(self._prepare(x, 0.0), self.pool_size, self.strides).transpose(1, 2)


class GlobalAveragePooling1D(Layer):
    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        return x.mean(dim=1)


class GlobalMaxPooling1D(GlobalAveragePooling1D):
    def call(self, x, training=False):
        return x.amax(dim=1)


# -- depthwise / separable convolutions, resampling ------------------------------------------------------
class DepthwiseConv2D(Layer):
    def __init__(self, kernel_size, strides=(1, 1), padding="valid", depth_multiplier=1, activation=None,
                 use_bias=True, depthwise_initializer="glorot_uniform", bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.kernel_size, self.strides = _pair(kernel_size), _pair(strides)
        self.padding = padding.lower()
        self.depth_multiplier = int(depth_multiplier)
        self.activation_name = activation if isinstance(activation, (str, type(None))) else None
        self._activation_fn = get_activation(activation)
        self.use_bias = use_bias
        self.depthwise_initializer, self.bias_initializer = depthwise_initializer, bias_initializer

    def _build_depthwise(self, cin):
        kh, kw = self.kernel_size
        m = self.depth_multiplier
        w = _init_tensor(self.depthwise_initializer, (kh, kw, cin, m), kh * kw * cin, kh * kw * m)
        # Keras [h, w, in, mult] <-> torch grouped conv [in * mult, 1, h, w] (output channel = in * mult + k)
        self.depthwise_kernel = nn.Parameter(w.permute(2, 3, 0, 1).reshape(cin * m, 1, kh, kw).contiguous())
        self.add_weight_spec("depthwise_kernel", "depthwise_kernel",
                             lambda t: t.reshape(cin, m, kh, kw).permute(2, 3, 0, 1),
                             lambda t: t.permute(2, 3, 0, 1).reshape(cin * m, 1, kh, kw))
        self._cin = cin

    def build(self, input_shape):
        cin = int(input_shape[-1])
        self._build_depthwise(cin)
        if self.use_bias:
            self.bias = nn.Parameter(_init_tensor(self.bias_initializer, (cin * self.depth_multiplier,), cin, cin))
            self.add_weight_spec("bias", "bias")
        else:
            self.bias = None

    def _out_hw(self, h, w):
        def one(n, k, s):
            if n is None:
                return None
            return -(-n // s) if self.padding == "same" else (n - k) // s + 1
        return one(h, self.kernel_size[0], self.strides[0]), one(w, self.kernel_size[1], self.strides[1])

    def compute_output_shape(self, input_shape):
        b, h, w, c = input_shape
        oh, ow = self._out_hw(h, w)
        return (b, oh, ow, c * self.depth_multiplier)

    def _depthwise(self, x):
        if self.padding == "same":
            ph = _same_pad(x.shape[2], self.kernel_size[0], self.strides[0])
            pw = _same_pad(x.shape[3], self.kernel_size[1], self.strides[1])
            if any(ph) or any(pw):
                x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
        w = self.depthwise_kernel
        if w.dtype != x.dtype:
            x = x.to(w.dtype)
        return F.conv2d(x, w, None, stride=self.strides, groups=self._cin)

    def call(self, x, training=False):
        y = self._depthwise(x)
        if self.bias is not None:
            y = y + self.bias.to(y.dtype).view(1, -1, 1, 1)
        return y if self._activation_fn is None else self._activation_fn(y)

    def get_config(self):
        return {**super().get_config(), "kernel_size": self.kernel_size, "strides": self.strides,
                "padding": self.padding, "depth_multiplier": self.depth_multiplier,
                "activation": self.activation_name, "use_bias": self.use_bias}


class SeparableConv2D(DepthwiseConv2D):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", depth_multiplier=1,
                 activation=None, use_bias=True, pointwise_initializer="glorot_uniform", **kwargs):
        super().__init__(kernel_size, strides=strides, padding=padding, depth_multiplier=depth_multiplier,
                         activation=activation, use_bias=use_bias, **kwargs)
        self.filters = int(filters)
        self.pointwise_initializer = pointwise_initializer

    def build(self, input_shape):
        cin = int(input_shape[-1])
        self._build_depthwise(cin)
        mid = cin * self.depth_multiplier
        w = _init_tensor(self.pointwise_initializer, (1, 1, mid, self.filters), mid, self.filters)
        self.pointwise_kernel = nn.Parameter(w.permute(3, 2, 0, 1).contiguous(memory_format=torch.channels_last))
        self.add_weight_spec("pointwise_kernel", "pointwise_kernel", lambda t: t.permute(2, 3, 1, 0),
                             lambda t: t.permute(3, 2, 0, 1))
        if self.use_bias:
            self.bias = nn.Parameter(_init_tensor(self.bias_initializer, (self.filters,), mid, self.filters))
            self.add_weight_spec("bias", "bias")
        else:
            self.bias = None

    def compute_output_shape(self, input_shape):
        b, h, w, _c = input_shape
        oh, ow = self._out_hw(h, w)
        return (b, oh, ow, self.filters)

    def call(self, x, training=False):
        y = self._depthwise(x)
        pw = self.pointwise_kernel
        y = F.conv2d(y, pw.to(y.dtype), None if self.bias is None else self.bias.to(y.dtype))
        return y if self._activation_fn is None else self._activation_fn(y)

    def get_config(self):
        return {**super().get_config(), "filters": self.filters}


class UpSampling2D(Layer):
    def __init__(self, size=(2, 2), interpolation="nearest", **kwargs):
        super().__init__(**kwargs)
        self.size, self.interpolation = _pair(size), interpolation

    def compute_output_shape(self, input_shape):
        b, h, w, c = input_shape
        return (b, None if h is None else h * self.size[0], None if w is None else w * self.size[1], c)

    def call(self, x, training=False):
        if self.interpolation == "nearest":
            return F.interpolate(x, scale_factor=self.size, mode="nearest")
        return F.interpolate(x, scale_factor=self.size, mode="bilinear", align_corners=False)

    def get_config(self):
        return {**super().get_config(), "size": self.size, "interpolation": self.interpolation}


class Cropping2D(Layer):
    def __init__(self, cropping=((0, 0), (0, 0)), **kwargs):
        super().__init__(**kwargs)
        if isinstance(cropping, int):
            cropping = ((cropping, cropping), (cropping, cropping))
        self.cropping = tuple(_pair(c) for c in cropping)

    def compute_output_shape(self, input_shape):
        b, h, w, c = input_shape
        (t, bo), (le, r) = self.cropping
        return (b, None if h is None else h - t - bo, None if w is None else w - le - r, c)

    def call(self, x, training=False):
        (t, bo), (le, r) = self.cropping
        return x[:, :, t:x.shape[2] - bo, le:x.shape[3] - r]

    def get_config(self):
        return {**super().get_config(), "cropping": self.cropping}


EXTRA_LAYERS = (LeakyReLU, ELU, PReLU, Subtract, Multiply, Average, Maximum, LayerNormalization, Embedding,
                Conv1D, MaxPooling1D, AveragePooling1D, GlobalAveragePooling1D, GlobalMaxPooling1D,
                DepthwiseConv2D, SeparableConv2D, UpSampling2D, Cropping2D)
