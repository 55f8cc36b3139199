"""Losses.  Sparse categorical cross-entropy (the loss of every reference workload:
README.md:77, core/tests/testdata/save_and_load.py:82) runs on the fused kernel
(``cloud_b200.ops.softmax_cross_entropy``); the others are plain torch and return the
sum over the local batch so the caller can apply the 1/global_batch scale."""
from __future__ import annotations

import torch
import torch.nn.functional as F


class Reduction:
    NONE = "none"
    SUM = "sum"
    AUTO = "auto"
    SUM_OVER_BATCH_SIZE = "sum_over_batch_size"


class Loss:
    name = "loss"
    fused_xent = False

    def __init__(self, reduction=Reduction.AUTO, name=None):
        self.reduction = reduction
        if name:
            self.name = name

    def per_sample(self, y_true, y_pred):
        raise NotImplementedError

    def __call__(self, y_true, y_pred, sample_weight=None):
        """Keras call convention (y_true, y_pred): mean over the batch unless reduction=NONE."""
        y_true = torch.as_tensor(y_true)
        y_pred = torch.as_tensor(y_pred)
        l = self.per_sample(y_true.to(y_pred.device), y_pred)
        if sample_weight is not None:
            l = l * torch.as_tensor(sample_weight, device=l.device, dtype=l.dtype)
        if self.reduction == Reduction.NONE:
            return l
        if self.reduction == Reduction.SUM:
            return l.sum()
        return l.mean()

    def get_config(self):
        return {"name": self.name, "reduction": self.reduction}


def _logits_of(y_pred, from_logits):
    if from_logits:
        return y_pred
    return getattr(y_pred, "_keras_logits", None)


class SparseCategoricalCrossentropy(Loss):
    name = "sparse_categorical_crossentropy"
    fused_xent = True

    def __init__(self, from_logits=False, reduction=Reduction.AUTO, name=None):
        super().__init__(reduction, name)
        self.from_logits = from_logits

    def logits(self, y_pred):
        return _logits_of(y_pred, self.from_logits)

    def per_sample(self, y_true, y_pred):
        y_true = y_true.reshape(-1).long()
        z = self.logits(y_pred)
        if z is not None:
            return F.cross_entropy(z.float(), y_true, reduction="none")
        p = y_pred.float().clamp_min(1e-7)
        return -torch.log(p.gather(-1, y_true.view(-1, 1)).squeeze(-1))

    def get_config(self):
        return {**super().get_config(), "from_logits": self.from_logits}


class CategoricalCrossentropy(Loss):
    name = "categorical_crossentropy"

    def __init__(self, from_logits=False, reduction=Reduction.AUTO, name=None):
        super().__init__(reduction, name)
        self.from_logits = from_logits

    def per_sample(self, y_true, y_pred):
        z = _logits_of(y_pred, self.from_logits)
        if z is not None:
            return -(y_true.float() * F.log_softmax(z.float(), dim=-1)).sum(-1)
        return -(y_true.float() * torch.log(y_pred.float().clamp_min(1e-7))).sum(-1)


class BinaryCrossentropy(Loss):
    name = "binary_crossentropy"

    def __init__(self, from_logits=False, reduction=Reduction.AUTO, name=None):
        super().__init__(reduction, name)
        self.from_logits = from_logits

    def per_sample(self, y_true, y_pred):
        y_true = y_true.float().reshape(y_pred.shape)
        if self.from_logits:
            l = F.binary_cross_entropy_with_logits(y_pred.float(), y_true, reduction="none")
        else:
            l = F.binary_cross_entropy(y_pred.float().clamp(1e-7, 1 - 1e-7), y_true, reduction="none")
        return l.reshape(l.shape[0], -1).mean(-1)


class MeanSquaredError(Loss):
    name = "mean_squared_error"

    def per_sample(self, y_true, y_pred):
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)) ** 2
        return d.reshape(d.shape[0], -1).mean(-1)


class MeanAbsoluteError(Loss):
    name = "mean_absolute_error"

    def per_sample(self, y_true, y_pred):
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)).abs()
        return d.reshape(d.shape[0], -1).mean(-1)


_BY_NAME = {
    "sparse_categorical_crossentropy": SparseCategoricalCrossentropy,
    "categorical_crossentropy": CategoricalCrossentropy,
    "binary_crossentropy": BinaryCrossentropy,
    "mse": MeanSquaredError, "mean_squared_error": MeanSquaredError,
    "mae": MeanAbsoluteError, "mean_absolute_error": MeanAbsoluteError,
}


def get(identifier):
    if isinstance(identifier, Loss):
        return identifier
    if isinstance(identifier, str) and identifier.lower() in _BY_NAME:
        return _BY_NAME[identifier.lower()]()
    if isinstance(identifier, dict):
        return deserialize(identifier)
    if callable(identifier):
        return _FunctionLoss(identifier)
    raise ValueError(f"Could not interpret loss function identifier: {identifier}")


class _FunctionLoss(Loss):
    def __init__(self, fn):
        super().__init__()
        self.fn = fn
        self.name = getattr(fn, "__name__", "loss")

    def per_sample(self, y_true, y_pred):
        l = self.fn(y_true, y_pred)
        return l if l.dim() else l.expand(y_pred.shape[0])


def serialize(loss: Loss) -> dict:
    return {"class_name": type(loss).__name__, "config": loss.get_config()}


def deserialize(d: dict) -> Loss:
    classes = {c.__name__: c for c in _BY_NAME.values()}
    cfg = dict(d.get("config", {}))
    cfg.pop("name", None)
    return classes[d["class_name"]](**cfg)
