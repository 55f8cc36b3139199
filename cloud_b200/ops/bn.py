"""Fused NHWC batch normalisation (+ residual add + ReLU) - filled in by csrc/kernels/bn.cu."""
from __future__ import annotations

_launches = 0


def launches() -> int:
    return _launches


def bn_available() -> bool:
    return False


def batch_norm_nhwc(*args, **kwargs):
    raise NotImplementedError
