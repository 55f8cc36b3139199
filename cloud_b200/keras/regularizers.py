"""``keras.regularizers``: weight penalties added to the training loss.

Reference usage: ``kernel_regularizer=tf.keras.regularizers.l2(0.01)`` in the cloud_fit test model
(experimental/cloud_fit/tests/unit/client_test.py:69-90 and tests/examples/cloud_fit.ipynb).  A
penalty is ``l1 * sum(|w|) + l2 * sum(w^2)`` of the layer's kernel or bias; ``Model`` adds the
penalties of all its layers to the loss it minimises and reports (``Model.losses``).
"""
from __future__ import annotations

import torch


class Regularizer:
    def __call__(self, w: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def get_config(self) -> dict:
        return {}

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class L1L2(Regularizer):
    def __init__(self, l1=0.0, l2=0.0):
        self.l1 = float(l1)
        self.l2 = float(l2)

    def __call__(self, w):
        w = w.float()
        out = w.new_zeros(())
        if self.l1:
            out = out + self.l1 * w.abs().sum()
        if self.l2:
            out = out + self.l2 * (w * w).sum()
        return out

    def get_config(self):
        return {"l1": self.l1, "l2": self.l2}

    def __repr__(self):
        return "L1L2(l1={}, l2={})".format(self.l1, self.l2)


class L1(L1L2):
    def __init__(self, l1=0.01, **kwargs):
        super().__init__(l1=kwargs.pop("l", l1))


class L2(L1L2):
    def __init__(self, l2=0.01, **kwargs):
        super().__init__(l2=kwargs.pop("l", l2))


def l1(l=0.01):          # noqa: E741  (Keras' own argument name)
    return L1L2(l1=l)


def l2(l=0.01):          # noqa: E741
    return L1L2(l2=l)


def l1_l2(l1=0.01, l2=0.01):
    return L1L2(l1=l1, l2=l2)


def serialize(reg):
    if reg is None:
        return None
    return {"class_name": "L1L2", "config": reg.get_config()}


def get(identifier):
    if identifier is None or isinstance(identifier, Regularizer):
        return identifier
    if isinstance(identifier, str):
        table = {"l1": l1, "l2": l2, "l1_l2": l1_l2}
        if identifier not in table:
            raise ValueError("Unknown regularizer: {}".format(identifier))
        return table[identifier]()
    if isinstance(identifier, dict):
        return L1L2(**identifier.get("config", identifier))
    if isinstance(identifier, tuple) and len(identifier) == 2:      # ("l2", 0.01)
        return L1L2(**{identifier[0]: identifier[1]})
    if callable(identifier):
        fn = identifier

        class _Fn(Regularizer):
            def __call__(self, w):
                return fn(w)
        return _Fn()
    raise ValueError("Could not interpret regularizer identifier: {}".format(identifier))
