"""``tf.nn`` helpers that custom training loops call under a distribution strategy
(reference usage: ``tf.nn.compute_average_loss(per_example_loss, global_batch_size=...)`` in
core/tests/testdata/mnist_example_using_ctl.py:97-101)."""
from __future__ import annotations

import torch

from .distribute import get_strategy


def compute_average_loss(per_example_loss, sample_weight=None, global_batch_size=None):
    """Sum of the per-example losses of THIS replica divided by the GLOBAL batch size, so that the
    sum over replicas - which is what the gradient all-reduce computes - is the global mean."""
    per_example_loss = torch.as_tensor(per_example_loss)
    if sample_weight is not None:
        per_example_loss = per_example_loss * torch.as_tensor(
            sample_weight, dtype=per_example_loss.dtype, device=per_example_loss.device)
    if global_batch_size is None:
        global_batch_size = per_example_loss.shape[0] * get_strategy().num_replicas_in_sync
    return per_example_loss.sum() / float(global_batch_size)


def scale_regularization_loss(regularization_loss):
    """A weight penalty is the same on every replica: each contributes 1 / num_replicas of it."""
    losses = regularization_loss if isinstance(regularization_loss, (list, tuple)) else [regularization_loss]
    total = sum(torch.as_tensor(l) for l in losses)
    return total / float(get_strategy().num_replicas_in_sync)


def softmax(logits, axis=-1):
    return torch.softmax(torch.as_tensor(logits), dim=axis)


def relu(x):
    return torch.relu(torch.as_tensor(x))
