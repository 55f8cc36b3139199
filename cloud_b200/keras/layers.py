"""Keras-shaped layers on torch, tuned for B200.

User-visible conventions follow Keras (what the reference's example scripts are written against,
e.g. README.md:71-75, core/tests/testdata/mnist_example_using_fit.py:54-63): NHWC image tensors,
``Dense`` kernels ``[in, out]``, ``Conv2D`` kernels ``HWIO``, auto names ``dense``, ``dense_1``...
Internally 4-D activations are NCHW-logical tensors in ``channels_last`` memory (zero-copy view of
the NHWC batch, and the layout cuDNN's tensor-core kernels want); Dense layers run on the
hand-written tcgen05 GEMM (``cloud_b200.ops.dense``).
"""
from __future__ import annotations

import os
import re

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops

# 1x1 / stride-1 convolutions go through ops/conv.py, which puts their weight gradient (default) or
# all three GEMMs (TFC_CONV1X1_GEMM=all) on the tcgen05 kernel; TFC_CONV1X1_GEMM=0: plain cuDNN
_CONV1X1_GEMM = ops.conv1x1_mode() != "0"
# k x k weight gradients on the tcgen05 shifted-box kernel: correct but slower than cuDNN's sm_100
# kernels on the 3x3 shapes of ResNet-50 (256 vs 110 us at 56x56x64, batch 256), so opt-in
_CONV_KXK = os.environ.get("TFC_CONV_KXK_WGRAD", "0") == "1"
# strided 1x1 convolutions as a subsample + a plain 1x1 convolution (ops.subsample); 0: leave the stride to cuDNN
_STRIDED_1X1 = os.environ.get("TFC_STRIDED_1X1", "1") != "0"

_NAME_COUNTS: dict[str, int] = {}


def reset_name_counts() -> None:
    _NAME_COUNTS.clear()


def _snake(name: str) -> str:
    s = re.sub(r"(.)([A-Z][a-z0-9]+)", r"\1_\2", name)
    s = re.sub(r"([a-z])([A-Z])", r"\1_\2", s).lower()
    return s.replace("2_d", "2d").replace("1_d", "1d").replace("3_d", "3d")


def _auto_name(cls_name: str) -> str:
    base = _snake(cls_name)
    n = _NAME_COUNTS.get(base, 0)
    _NAME_COUNTS[base] = n + 1
    if base == "input_layer":
        return f"input_{n + 1}"
    return base if n == 0 else f"{base}_{n}"


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


# ----------------------------------------------------------------------------------------------
# Symbolic tensors for the functional API
# ----------------------------------------------------------------------------------------------
class KerasTensor:
    """Placeholder produced by ``Input`` / by calling a layer on placeholders."""

    def __init__(self, shape, layer=None, inputs=(), dtype="float32"):
        self.shape = tuple(shape)
        self.layer = layer
        self.inputs = tuple(inputs)
        self.dtype = dtype

    def __repr__(self):
        return f"<KerasTensor shape={self.shape} from {getattr(self.layer, 'name', None)}>"


_ACTIVATIONS = {
    None: None, "linear": None,
    "relu": torch.relu, "elu": F.elu, "selu": F.selu, "gelu": F.gelu, "sigmoid": torch.sigmoid,
    "tanh": torch.tanh, "softplus": F.softplus, "swish": F.silu, "silu": F.silu,
    "softmax": lambda x: torch.softmax(x.float(), dim=-1).to(x.dtype),
}


def get_activation(name):
    if callable(name):
        return name
    if name not in _ACTIVATIONS:
        raise ValueError(f"Unknown activation function: {name}")
    return _ACTIVATIONS[name]


def _init_tensor(initializer, shape, fan_in, fan_out):
    """A float32 tensor of ``shape`` filled by a Keras initializer (name, config dict or object)."""
    from . import initializers
    return initializers.get(initializer).fill(torch.empty(shape, dtype=torch.float32), fan_in, fan_out)


class Layer(nn.Module):
    """Base layer: lazy ``build`` on the first call, Keras naming and weight (de)serialisation."""

    def __init__(self, name=None, trainable=True, dtype=None, input_shape=None,
                 batch_input_shape=None, **kwargs):
        super().__init__()
        if kwargs:
            unknown = set(kwargs) - {"input_dim", "activity_regularizer", "kernel_regularizer",
                                     "bias_regularizer", "kernel_constraint", "bias_constraint"}
            if unknown:
                raise TypeError(f"Unknown keyword arguments for {type(self).__name__}: {sorted(unknown)}")
        self.name = name or _auto_name(type(self).__name__)
        self._trainable = trainable
        self.built = False
        from . import regularizers
        self.kernel_regularizer = regularizers.get(kwargs.get("kernel_regularizer"))
        self.bias_regularizer = regularizers.get(kwargs.get("bias_regularizer"))
        if input_shape is None and "input_dim" in kwargs:
            input_shape = (kwargs["input_dim"],)
        if batch_input_shape is not None:
            self._declared_input_shape = tuple(batch_input_shape)
        elif input_shape is not None:
            self._declared_input_shape = (None,) + tuple(input_shape)
        else:
            self._declared_input_shape = None
        self._keras_weight_specs: list = []   # (keras_suffix, attr, to_keras, from_keras)
        self.input_shape_ = None
        self.output_shape_ = None

    # -- Keras-facing -------------------------------------------------------------------------
    @property
    def trainable(self):
        return self._trainable

    @trainable.setter
    def trainable(self, v):
        self._trainable = bool(v)
        for p in self.parameters():
            p.requires_grad_(bool(v))
        self._on_trainable_changed()

    def _on_trainable_changed(self):
        pass

    def build(self, input_shape):
        pass

    def call(self, x, training=False):
        return x

    def compute_output_shape(self, input_shape):
        return input_shape

    def _maybe_build(self, input_shape):
        if not self.built:
            self.build(tuple(input_shape))
            self.built = True
            self.input_shape_ = tuple(input_shape)
            self.output_shape_ = self.compute_output_shape(tuple(input_shape))
            if not self._trainable:
                for p in self.parameters(recurse=False):
                    p.requires_grad_(False)

    def __call__(self, inputs, *args, **kwargs):
        if isinstance(inputs, KerasTensor) or (
                isinstance(inputs, (list, tuple)) and inputs and isinstance(inputs[0], KerasTensor)):
            return self._symbolic_call(inputs)
        return super().__call__(inputs, *args, **kwargs)

    def _symbolic_call(self, inputs):
        ins = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        shape = [t.shape for t in ins] if len(ins) > 1 else ins[0].shape
        self._maybe_build(shape)
        return KerasTensor(self.compute_output_shape(shape), layer=self, inputs=ins)

    def forward(self, x, training=None):
        if training is None:
            training = self.training
        if not self.built:
            self._maybe_build(_keras_shape(x))
            dev = _first_device(x)
            if dev is not None:
                self.to(dev)
        return self.call(x, training=training)

    # -- weights ------------------------------------------------------------------------------
    def add_weight_spec(self, suffix, attr, to_keras=None, from_keras=None):
        self._keras_weight_specs.append((suffix, attr, to_keras, from_keras))

    @property
    def weight_names(self):
        return [f"{self.name}/{s}:0" for s, *_ in self._keras_weight_specs]

    def get_weights(self):
        out = []
        for _s, attr, to_keras, _f in self._keras_weight_specs:
            src = getattr(self, attr)
            acc = getattr(src, "_cb2_fp32", None)
            # a parameter owned by a training engine answers with its fp32 master, not with the bf16
            # rounding the kernels compute on: what is saved is what training continues from
            t = acc.get() if acc is not None else None
            if t is None:
                t = src.detach().float().cpu()
            if to_keras is not None:
                t = to_keras(t)
            out.append(np.ascontiguousarray(t.numpy()))
        return out

    def set_weights(self, weights):
        if len(weights) != len(self._keras_weight_specs):
            raise ValueError(f'You called `set_weights(weights)` on layer "{self.name}" with a weight '
                             f"list of length {len(weights)}, but the layer was expecting "
                             f"{len(self._keras_weight_specs)} weights.")
        with torch.no_grad():
            for w, (_s, attr, _t, from_keras) in zip(weights, self._keras_weight_specs):
                t = torch.as_tensor(np.asarray(w), dtype=torch.float32)
                if from_keras is not None:
                    t = from_keras(t)
                dst = getattr(self, attr)
                if tuple(t.shape) != tuple(dst.shape):
                    raise ValueError(f"Layer {self.name}: weight shape {tuple(t.shape)} is not "
                                     f"compatible with {tuple(dst.shape)}")
                acc = getattr(dst, "_cb2_fp32", None)
                if acc is not None:
                    acc.set(t)                 # fp32 into the master, its rounding into the arena
                else:
                    dst.copy_(t.to(dst.device, dst.dtype))

    def count_params(self):
        return sum(int(getattr(self, attr).numel()) for _s, attr, *_ in self._keras_weight_specs)

    @property
    def losses(self):
        """Weight penalties of this layer (``kernel_regularizer`` / ``bias_regularizer``)."""
        out = []
        for reg, attr in ((self.kernel_regularizer, "kernel"), (self.bias_regularizer, "bias")):
            w = getattr(self, attr, None) if reg is not None else None
            if w is not None and (w.requires_grad or not torch.is_grad_enabled()):
                out.append(reg(w))
        return out

    def get_config(self):
        from . import regularizers
        cfg = {"name": self.name, "trainable": self._trainable}
        for key in ("kernel_regularizer", "bias_regularizer"):
            if getattr(self, key) is not None:
                cfg[key] = regularizers.serialize(getattr(self, key))
        return cfg

    @classmethod
    def from_config(cls, config):
        return cls(**config)


def _first_device(x):
    if isinstance(x, torch.Tensor):
        return x.device
    if isinstance(x, (list, tuple)) and x and isinstance(x[0], torch.Tensor):
        return x[0].device
    return None


def _keras_shape(x):
    """Keras-convention shape (None batch, NHWC) of an internal tensor (NCHW for 4-D)."""
    if isinstance(x, (list, tuple)):
        return [_keras_shape(t) for t in x]
    s = tuple(x.shape)
    if len(s) == 4:
        return (None, s[2], s[3], s[1])
    return (None,) + s[1:]


# ----------------------------------------------------------------------------------------------
class InputLayer(Layer):
    def __init__(self, input_shape=None, batch_size=None, dtype=None, name=None, **kw):
        super().__init__(name=name, input_shape=input_shape, **kw)


def Input(shape=None, batch_size=None, name=None, dtype="float32", **kwargs):
    """``tf.keras.Input``: a symbolic placeholder of shape ``(None,) + shape``."""
    layer = InputLayer(input_shape=tuple(shape), name=name)
    layer.built = True
    full = (batch_size,) + tuple(shape)
    layer.input_shape_ = layer.output_shape_ = full
    return KerasTensor(full, layer=layer, inputs=(), dtype=dtype)


# Set by Model._train_step while fit() trains with the fused cross-entropy: the softmax Dense layer
# named here returns its logits only (see Dense.call).
LOGITS_ONLY = {"on": False, "layer": None}


class Dense(Layer):
    """``activation(x @ kernel + bias)`` on the tcgen05 GEMM with a fused epilogue (K1/K2/K3)."""

    def __init__(self, units, activation=None, use_bias=True, kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.units = int(units)
        self.activation_name = activation if isinstance(activation, (str, type(None))) else None
        self._activation_fn = get_activation(activation)
        self.use_bias = use_bias
        self.kernel_initializer = kernel_initializer
        self.bias_initializer = bias_initializer
        self.last_logits = None
        self._fused_dropout = None      # set by Model._mark_fusions: the Dropout layer right behind

    def build(self, input_shape):
        in_dim = int(input_shape[-1])
        # torch layout [out, in]; Keras files hold [in, out] (transposed on save/load)
        w = _init_tensor(self.kernel_initializer, (in_dim, self.units), in_dim, self.units)
        self.kernel = nn.Parameter(w.t().contiguous())
        self.add_weight_spec("kernel", "kernel", lambda t: t.t(), lambda t: t.t())
        if self.use_bias:
            self.bias = nn.Parameter(_init_tensor(self.bias_initializer, (self.units,), in_dim, self.units))
            self.add_weight_spec("bias", "bias")
        else:
            self.bias = None

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[:-1]) + (self.units,)

    def call(self, x, training=False):
        act = self.activation_name
        if act == "softmax":
            # keep the logits: the fused cross-entropy consumes them (what Keras does with
            # `_keras_logits`), the probabilities are only materialised when somebody looks
            logits = ops.dense(x, self.kernel, self.bias, None)
            if training and LOGITS_ONLY["on"] and LOGITS_ONLY["layer"] is self and torch.is_grad_enabled():
                # inside fit()'s training step with the fused cross-entropy and metrics that come out
                # of the same kernel: nobody reads the probabilities, so they are not computed
                # (a view, not the tensor itself: an attribute pointing back at its owner would keep
                # the tensor alive until the garbage collector runs - possibly in the middle of a
                # CUDA-graph capture, where freeing device memory of an earlier step is an error)
                out = logits.view_as(logits)
                out._keras_logits = logits
                return out
            out = _LazySoftmax.make(logits)
            return out
        drop = self._fused_dropout
        if (drop is not None and training and act == "relu" and drop.rate > 0.0 and torch.is_grad_enabled()
                and x.is_cuda):
            # Dense + ReLU + Dropout: the dropout is applied here and undone in this layer's backward
            # (ReLU mask, dropout scale and bias gradient in one pass); the Dropout layer passes it on
            y = ops.dense(x, self.kernel, self.bias, act, dropout=ops.dropout_plan(x.shape[0] * self.units,
                                                                                   drop.rate, drop.seed))
            y._cb2_dropped = drop
            return y
        if act in (None, "linear", "relu"):
            return ops.dense(x, self.kernel, self.bias, act)
        y = ops.dense(x, self.kernel, self.bias, None)
        return self._activation_fn(y)

    def get_config(self):
        return {**super().get_config(), "units": self.units, "activation": self.activation_name,
                "use_bias": self.use_bias}


class _LazySoftmax:
    """Tags a probability tensor with the logits it came from."""

    @staticmethod
    def make(logits):
        probs = torch.softmax(logits.float(), dim=-1).to(logits.dtype) if not torch.is_grad_enabled() \
            else _SoftmaxWithLogits.apply(logits)
        probs._keras_logits = logits
        return probs


class _SoftmaxWithLogits(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits):
        p = torch.softmax(logits.float(), dim=-1).to(logits.dtype)
        ctx.save_for_backward(p)
        return p

    @staticmethod
    def backward(ctx, dy):
        (p,) = ctx.saved_tensors
        pf, df = p.float(), dy.float()
        return (pf * (df - (df * pf).sum(-1, keepdim=True))).to(dy.dtype)


class Activation(Layer):
    def __init__(self, activation, **kwargs):
        super().__init__(**kwargs)
        self.activation_name = activation if isinstance(activation, str) else None
        self._fn = get_activation(activation)

    def call(self, x, training=False):
        if self.activation_name == "softmax":
            return _LazySoftmax.make(x)
        return x if self._fn is None else self._fn(x)

    def get_config(self):
        return {**super().get_config(), "activation": self.activation_name}


class ReLU(Layer):
    def call(self, x, training=False):
        return torch.relu(x)


class Softmax(Layer):
    def call(self, x, training=False):
        return _LazySoftmax.make(x)


class Dropout(Layer):
    def __init__(self, rate, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.rate = float(rate)
        self.seed = seed

    def call(self, x, training=False):
        if training and getattr(x, "_cb2_dropped", None) is self:
            return x                     # the Dense layer in front already applied this dropout
        return ops.dropout(x, self.rate, training=training, seed=self.seed)

    def get_config(self):
        return {**super().get_config(), "rate": self.rate}


class Flatten(Layer):
    def compute_output_shape(self, input_shape):
        n = 1
        for d in input_shape[1:]:
            n *= int(d)
        return (input_shape[0], n)

    def call(self, x, training=False):
        if x.dim() == 4:   # NCHW-logical -> flatten in Keras (H, W, C) order; a view for channels_last
            x = x.permute(0, 2, 3, 1)
        return x.reshape(x.shape[0], -1)


class Reshape(Layer):
    def __init__(self, target_shape, **kwargs):
        super().__init__(**kwargs)
        self.target_shape = tuple(target_shape)

    def compute_output_shape(self, input_shape):
        return (input_shape[0],) + self.target_shape

    def call(self, x, training=False):
        if x.dim() == 4:
            x = x.permute(0, 2, 3, 1)
        y = x.reshape((x.shape[0],) + self.target_shape)
        if y.dim() == 4:
            y = y.permute(0, 3, 1, 2)
        return y

    def get_config(self):
        return {**super().get_config(), "target_shape": self.target_shape}


def _same_pad(in_size, k, s, d=1):
    eff = (k - 1) * d + 1
    out = -(-in_size // s)
    total = max((out - 1) * s + eff - in_size, 0)
    return total // 2, total - total // 2


class Conv2D(Layer):
    """NHWC convolution on cuDNN's channels_last tensor-core kernels (library op, K8)."""

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 use_bias=True, dilation_rate=(1, 1), kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", **kwargs):
        super().__init__(**kwargs)
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        # 'valid' | 'same' as in Keras, or an explicit symmetric zero padding (int or (ph, pw)):
        # the fused form of ZeroPadding2D + 'valid' convolution (saves a full copy of the input)
        self.padding = padding.lower() if isinstance(padding, str) else _pair(padding)
        self.dilation = _pair(dilation_rate)
        self.activation_name = activation if isinstance(activation, (str, type(None))) else None
        self._activation_fn = get_activation(activation)
        self.use_bias = use_bias
        self.kernel_initializer = kernel_initializer
        self.bias_initializer = bias_initializer
        # Set by a directly following BatchNormalization: in training mode batch-norm removes any
        # per-channel shift, so the bias add (one extra read+write of the activation) and its
        # gradient reduction (one extra read of dY; the gradient is exactly zero) are skipped.
        self.bias_absorbed_by_bn = False
        # Set by a directly following trainable BatchNormalization: a 1x1 convolution on the
        # tcgen05 GEMM then reduces the batch statistics in its epilogue (ops/conv.py).
        self.feeds_bn = False
        # Set when that BatchNormalization is the ONLY consumer of this layer's output: a wide 1x1
        # convolution then hands its operands on and the pair runs as the folded GEMMs of
        # ops/convbn.py (the convolution output is never stored).
        self.fold_candidate = False

    def build(self, input_shape):
        cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        fan_in, fan_out = cin * kh * kw, self.filters * kh * kw
        w = _init_tensor(self.kernel_initializer, (kh, kw, cin, self.filters), fan_in, fan_out)
        self.kernel = nn.Parameter(w.permute(3, 2, 0, 1).contiguous(memory_format=torch.channels_last))
        self.add_weight_spec("kernel", "kernel", lambda t: t.permute(2, 3, 1, 0),
                             lambda t: t.permute(3, 2, 0, 1))
        if self.use_bias:
            self.bias = nn.Parameter(_init_tensor(self.bias_initializer, (self.filters,), fan_in, fan_out))
            self.add_weight_spec("bias", "bias")
        else:
            self.bias = None

    def _out_dim(self, n, k, s, d, axis=0):
        if n is None:
            return None
        if self.padding == "same":
            return -(-n // s)
        extra = 2 * self.padding[axis] if isinstance(self.padding, tuple) else 0
        return (n + extra - ((k - 1) * d + 1)) // s + 1

    def compute_output_shape(self, input_shape):
        b, h, w, _c = input_shape
        return (b, self._out_dim(h, self.kernel_size[0], self.strides[0], self.dilation[0], 0),
                self._out_dim(w, self.kernel_size[1], self.strides[1], self.dilation[1], 1), self.filters)

    def is_stem_geometry(self) -> bool:
        return (self.filters == 64 and self.kernel_size == (7, 7) and self.strides == (2, 2)
                and self.padding == (3, 3) and self.dilation == (1, 1))

    def call(self, x, training=False):
        pending = getattr(x, "_cb2_pending_norm", None)
        if pending is not None:          # raw uint8 image from a deferring ImageNormalization
            mean, std = pending
            nhwc = x.permute(0, 2, 3, 1)
            w = self.kernel
            if (self.is_stem_geometry() and w.dtype == torch.bfloat16 and x.shape[1] == 3
                    and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0 and x.shape[3] // 2 <= 128):
                b = self.bias
                if b is not None and self.bias_absorbed_by_bn and training:
                    b = None
                y = ops.stem_conv_s2d(ops.normalize_u8_images_s2d(nhwc.contiguous(), mean, std), w, b)
                if self.activation_name == "relu":
                    return torch.relu_(y)
                return y if self._activation_fn is None else self._activation_fn(y)
            x = ops.normalize_u8_images(nhwc.contiguous(), mean, std).permute(0, 3, 1, 2)
        pad = self.padding if isinstance(self.padding, tuple) else 0
        if self.padding == "same":
            ph = _same_pad(x.shape[2], self.kernel_size[0], self.strides[0], self.dilation[0])
            pw = _same_pad(x.shape[3], self.kernel_size[1], self.strides[1], self.dilation[1])
            if ph[0] == ph[1] and pw[0] == pw[1]:
                pad = (ph[0], pw[0])
            else:
                x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
        w = self.kernel
        if w.dtype != x.dtype:
            x = x.to(w.dtype)
        strides = self.strides
        if (_STRIDED_1X1 and self.kernel_size == (1, 1) and strides[0] == strides[1] and strides[0] > 1
                and pad in (0, (0, 0)) and x.is_cuda and x.dim() == 4 and self.dilation == (1, 1)):
            x = ops.subsample(x, strides[0])        # a strided 1x1 convolution reads every s-th pixel
            strides = (1, 1)
        if (training and self.fold_candidate and (self.bias_absorbed_by_bn or self.bias is None)
                and self.activation_name in (None, "linear") and torch.is_grad_enabled()
                and ops.convbn_supported(x, w, strides, pad, self.dilation)):
            return ops.DeferredConv(x, w, self, pad)
        return self._plain_conv(x, training, pad, strides)

    def _plain_conv(self, x, training=True, pad=0, strides=None):
        strides = self.strides if strides is None else strides
        w = self.kernel
        b = self.bias
        if b is not None and self.bias_absorbed_by_bn and training:
            b = None
        if b is not None and b.dtype != x.dtype:
            b = b.to(x.dtype)
        if _CONV1X1_GEMM and ops.stem_supported(x, w, strides, pad, self.dilation):
            y = ops.stem_conv(x, w, b)
        elif _CONV1X1_GEMM and ops.kxk_supported(x, w, strides, pad, self.dilation) and (
                _CONV_KXK or (b is None and ops.conv3x3_own(x, w))):
            y = ops.conv_kxk(x, w, b, bn_stats=bool(training) and self.feeds_bn and
                             self.activation_name in (None, "linear"))
        elif _CONV1X1_GEMM and ops.conv1x1_supported(x, w, strides, pad, self.dilation):
            y = ops.conv1x1(x, w, b, bn_stats=bool(training) and self.feeds_bn and
                            self.activation_name in (None, "linear"))
        else:
            y = F.conv2d(x, w, b, stride=strides, padding=pad, dilation=self.dilation)
        if self.activation_name == "relu":
            return torch.relu_(y)
        return y if self._activation_fn is None else self._activation_fn(y)

    def get_config(self):
        return {**super().get_config(), "filters": self.filters, "kernel_size": self.kernel_size,
                "strides": self.strides, "padding": self.padding, "activation": self.activation_name,
                "use_bias": self.use_bias}


class _Pool2D(Layer):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", **kwargs):
        super().__init__(**kwargs)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        # 'valid' | 'same', or an explicit symmetric ZERO padding (int): fused ZeroPadding2D + pool
        self.padding = padding.lower() if isinstance(padding, str) else int(padding)

    def compute_output_shape(self, input_shape):
        b, h, w, c = input_shape

        def o(n, k, s):
            if n is None:
                return None
            if isinstance(self.padding, int):
                return (n + 2 * self.padding - k) // s + 1
            return -(-n // s) if self.padding == "same" else (n - k) // s + 1
        return (b, o(h, self.pool_size[0], self.strides[0]), o(w, self.pool_size[1], self.strides[1]), c)

    def _pad(self, x, value):
        if self.padding != "same":
            return x
        ph = _same_pad(x.shape[2], self.pool_size[0], self.strides[0])
        pw = _same_pad(x.shape[3], self.pool_size[1], self.strides[1])
        if sum(ph) + sum(pw) == 0:
            return x
        return F.pad(x, (pw[0], pw[1], ph[0], ph[1]), value=value)

    def get_config(self):
        return {**super().get_config(), "pool_size": self.pool_size, "strides": self.strides,
                "padding": self.padding}


class MaxPooling2D(_Pool2D):
    def call(self, x, training=False):
        square = self.pool_size[0] == self.pool_size[1] and self.strides[0] == self.strides[1]
        if square and self.padding != "same":
            pad = self.padding if isinstance(self.padding, int) else 0
            return ops.max_pool_nhwc(x, self.pool_size[0], self.strides[0], pad)
        if isinstance(self.padding, int) and self.padding:
            p = self.padding
            x = F.pad(x, (p, p, p, p))
        return F.max_pool2d(self._pad(x, float("-inf")), self.pool_size, self.strides)


MaxPool2D = MaxPooling2D


class AveragePooling2D(_Pool2D):
    def call(self, x, training=False):
        if isinstance(self.padding, int) and self.padding:
            p = self.padding
            x = F.pad(x, (p, p, p, p))
        return F.avg_pool2d(self._pad(x, 0.0), self.pool_size, self.strides)


class GlobalAveragePooling2D(Layer):
    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        return x.mean(dim=(2, 3))


class GlobalMaxPooling2D(Layer):
    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        return x.amax(dim=(2, 3))


class ZeroPadding2D(Layer):
    def __init__(self, padding=(1, 1), **kwargs):
        super().__init__(**kwargs)
        if isinstance(padding, int):
            padding = ((padding, padding), (padding, padding))
        elif isinstance(padding[0], int):
            padding = ((padding[0], padding[0]), (padding[1], padding[1]))
        self.padding = tuple(tuple(p) for p in padding)

    def compute_output_shape(self, input_shape):
        b, h, w, c = input_shape
        return (b, None if h is None else h + sum(self.padding[0]),
                None if w is None else w + sum(self.padding[1]), c)

    def call(self, x, training=False):
        (t, b), (l, r) = self.padding
        return F.pad(x, (l, r, t, b))

    def get_config(self):
        return {**super().get_config(), "padding": self.padding}


class BatchNormalization(Layer):
    """Batch norm over the channel axis.  Moving statistics stay per-replica (tf.distribute
    semantics) and are averaged across replicas when saving."""

    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, center=True, scale=True,
                 gamma_initializer="ones", activation=None, **kwargs):
        super().__init__(**kwargs)
        self.momentum = momentum
        self.epsilon = epsilon
        self.center, self.scale = center, scale
        self.gamma_initializer = gamma_initializer
        if activation not in (None, "relu"):
            raise ValueError("BatchNormalization can only fuse activation='relu'")
        self.fused_activation = activation     # extension: BN+ReLU in one memory pass
        self._absorbed_conv = None             # (set in _symbolic_call) conv whose bias we absorb

    def build(self, input_shape):
        c = int(input_shape[-1])
        g0 = torch.zeros(c) if self.gamma_initializer == "zeros" else torch.ones(c)
        self.gamma = nn.Parameter(g0, requires_grad=self.scale)
        self.beta = nn.Parameter(torch.zeros(c), requires_grad=self.center)
        self.register_buffer("moving_mean", torch.zeros(c))
        self.register_buffer("moving_variance", torch.ones(c))
        self.add_weight_spec("gamma", "gamma")
        self.add_weight_spec("beta", "beta")
        self.add_weight_spec("moving_mean", "moving_mean")
        self.add_weight_spec("moving_variance", "moving_variance")

    def compute_output_shape(self, input_shape):
        if isinstance(input_shape, list):
            return tuple(input_shape[0])
        return tuple(input_shape)

    def _maybe_build(self, input_shape):
        if isinstance(input_shape, list):
            input_shape = input_shape[0]
        super()._maybe_build(input_shape)

    def _symbolic_call(self, inputs):
        out = super()._symbolic_call(inputs)
        first = inputs[0] if isinstance(inputs, (list, tuple)) else inputs
        prev = first.layer
        if isinstance(prev, Conv2D) and prev.activation_name in (None, "linear") and self._trainable:
            prev.feeds_bn = True
            prev.fold_candidate = True       # Model._init_functional clears it if the output has other consumers
            if prev.use_bias:
                prev.bias_absorbed_by_bn = True
                object.__setattr__(self, "_absorbed_conv", prev)   # not a submodule: no double registration
        return out

    def _on_trainable_changed(self):
        # A frozen batch-norm runs on its moving statistics (Keras semantics), which do not cancel a
        # per-channel shift: the convolution in front has to add its bias itself again.
        conv = self._absorbed_conv
        if conv is not None:
            conv.bias_absorbed_by_bn = self._trainable
        if self.built:
            self.gamma.requires_grad_(self._trainable and self.scale)
            self.beta.requires_grad_(self._trainable and self.center)

    def call(self, x, training=False, relu=None, residual=None):
        if isinstance(x, (list, tuple)):       # [x, shortcut]: y = act(BN(x) + shortcut)
            x, residual = x
        if relu is None:
            relu = self.fused_activation == "relu"
        train = bool(training) and self._trainable
        conv = self._absorbed_conv
        mean_bias = conv.bias if (conv is not None and train and conv.bias is not None) else None
        if isinstance(x, ops.DeferredConv):
            d = x
            ok = (train and self.gamma.dtype == torch.float32 and ops.bn_available() and
                  (residual is None or (isinstance(residual, torch.Tensor) and residual.dtype == torch.bfloat16
                                        and residual.dim() == 4 and residual.shape[0] == d.x.shape[0]
                                        and residual.shape[1] == d.w.shape[0]
                                        and tuple(residual.shape[2:]) == tuple(d.x.shape[2:])
                                        and residual.is_contiguous(memory_format=torch.channels_last))))
            if ok:
                return ops.conv1x1_bn(d.x, d.w, self.gamma, self.beta, self.moving_mean, self.moving_variance,
                                      1.0 - self.momentum, self.epsilon, relu=relu, residual=residual,
                                      mean_bias=mean_bias)
            x = d.materialize()
        if x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16 and ops.bn_available():
            return ops.batch_norm_nhwc(x, self.gamma, self.beta, self.moving_mean,
                                       self.moving_variance, train, 1.0 - self.momentum,
                                       self.epsilon, relu=relu, residual=residual,
                                       mean_bias=mean_bias,
                                       sums=getattr(x, "_cb2_bn_sums", None) if train else None)
        y = F.batch_norm(x, self.moving_mean, self.moving_variance, self.gamma, self.beta, train,
                         1.0 - self.momentum, self.epsilon)
        if mean_bias is not None:
            with torch.no_grad():    # the skipped conv bias still belongs in the moving mean
                # .data: batch_norm saved the buffer for backward; this update is not part of the graph
                self.moving_mean.data.add_((1.0 - self.momentum) * mean_bias.detach().float())
        if residual is not None:
            y = y + residual
        return torch.relu_(y) if relu else y

    def get_config(self):
        return {**super().get_config(), "momentum": self.momentum, "epsilon": self.epsilon,
                "activation": self.fused_activation}


class ImageNormalization(Layer):
    """uint8 NHWC images -> normalised bf16 activations in one kernel (input-pipeline tail, K12).
    Float inputs are normalised with torch ops."""

    def __init__(self, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225), **kwargs):
        super().__init__(**kwargs)
        self.mean, self.std = tuple(mean), tuple(std)
        # Set by Model._init_functional when the only consumer is a ResNet-stem convolution: the
        # bytes are then normalised straight into the stem kernel's space-to-depth layout by the
        # convolution (one pass instead of normalise + pad + regroup).
        self.defer_to_stem = False

    def call(self, x, training=False):
        if x.dtype == torch.uint8 and self.defer_to_stem and x.is_cuda and _CONV1X1_GEMM:
            x._cb2_pending_norm = (self.mean, self.std)
            return x
        if x.dtype == torch.uint8:
            nhwc = x.permute(0, 2, 3, 1)                     # back to the raw NHWC bytes (a view)
            out = ops.normalize_u8_images(nhwc.contiguous(), self.mean, self.std)
            if not x.is_cuda:
                out = out.float()
            return out.permute(0, 3, 1, 2)
        c = x.shape[1]
        m = torch.tensor(self.mean[:c], device=x.device, dtype=x.dtype).view(1, c, 1, 1)
        s = torch.tensor(self.std[:c], device=x.device, dtype=x.dtype).view(1, c, 1, 1)
        return (x - m) / s

    def get_config(self):
        return {**super().get_config(), "mean": self.mean, "std": self.std}


class Add(Layer):
    def compute_output_shape(self, input_shape):
        return tuple(input_shape[0])

    def call(self, xs, training=False):
        y = xs[0]
        for t in xs[1:]:
            y = y + t
        return y


class Concatenate(Layer):
    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def compute_output_shape(self, input_shape):
        s = list(input_shape[0])
        ax = self.axis if self.axis >= 0 else len(s) + self.axis
        s[ax] = sum(int(sh[ax]) for sh in input_shape)
        return tuple(s)

    def call(self, xs, training=False):
        nd = xs[0].dim()
        ax = self.axis if self.axis >= 0 else nd + self.axis
        if nd == 4:   # Keras NHWC axis -> internal NCHW axis
            ax = {0: 0, 1: 2, 2: 3, 3: 1}[ax]
        return torch.cat(list(xs), dim=ax)

    def get_config(self):
        return {**super().get_config(), "axis": self.axis}


class Lambda(Layer):
    def __init__(self, function, **kwargs):
        super().__init__(**kwargs)
        self.function = function

    def call(self, x, training=False):
        return self.function(x)


class Rescaling(Layer):
    def __init__(self, scale, offset=0.0, **kwargs):
        super().__init__(**kwargs)
        self.scale, self.offset = scale, offset

    def call(self, x, training=False):
        return x * self.scale + self.offset

    def get_config(self):
        return {**super().get_config(), "scale": self.scale, "offset": self.offset}


class RandomFlip(Layer):
    """Horizontal/vertical random flip (core/tests/testdata/keras_tuner_cifar_example.py:24-31)."""

    def __init__(self, mode="horizontal_and_vertical", **kwargs):
        super().__init__(**kwargs)
        self.mode = mode

    def call(self, x, training=False):
        if not training:
            return x
        if "horizontal" in self.mode and torch.rand(()) < 0.5:
            x = x.flip(3)
        if "vertical" in self.mode and torch.rand(()) < 0.5:
            x = x.flip(2)
        return x

    def get_config(self):
        return {**super().get_config(), "mode": self.mode}


class _IdentityAugment(Layer):
    """Geometry-changing augmentations are accepted and act as identity (shape-preserving)."""

    def __init__(self, factor=0.0, **kwargs):
        super().__init__(**kwargs)
        self.factor = factor

    def get_config(self):
        return {**super().get_config(), "factor": self.factor}


class RandomRotation(_IdentityAugment):
    pass


class RandomWidth(_IdentityAugment):
    pass


class RandomHeight(_IdentityAugment):
    pass


class RandomZoom(_IdentityAugment):
    pass


class experimental:  # tf.keras.layers.experimental.preprocessing.*
    class preprocessing:
        RandomFlip = RandomFlip
        RandomRotation = RandomRotation
        RandomWidth = RandomWidth
        RandomHeight = RandomHeight
        RandomZoom = RandomZoom
        Rescaling = Rescaling


LAYER_CLASSES = {c.__name__: c for c in (
    InputLayer, Dense, Activation, ReLU, Softmax, Dropout, Flatten, Reshape, Conv2D, MaxPooling2D,
    AveragePooling2D, GlobalAveragePooling2D, GlobalMaxPooling2D, ZeroPadding2D,
    BatchNormalization, ImageNormalization, Add, Concatenate, Rescaling, RandomFlip, RandomRotation, RandomWidth,
    RandomHeight, RandomZoom)}


# the layers of extra_layers.py live in this namespace too (`keras.layers.SeparableConv2D`, ...)
from . import extra_layers as _extra  # noqa: E402
for _cls in _extra.EXTRA_LAYERS:
    globals()[_cls.__name__] = _cls
    LAYER_CLASSES[_cls.__name__] = _cls
del _cls
