"""Fused sparse softmax cross-entropy (``csrc/kernels/elementwise.cu::softmax_xent_kernel``).

One pass over the logits produces the summed loss, the gradient (already scaled by
``1/global_batch``, the Keras/`tf.distribute` convention) and the accuracy hit count, so the
backward pass of the loss costs nothing and the metric needs no extra kernel.

Reference workload: ``loss='sparse_categorical_crossentropy'`` + ``metrics=['accuracy']``
(README.md:77-79); global-batch scaling as in testdata/mnist_example_using_ctl.py:97-101.
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn.functional as F

from . import _lib

_launches = 0


def launches() -> int:
    return _launches


def softmax_cross_entropy_reference(logits, labels, grad_scale=None):
    """fp32 PyTorch oracle: returns (loss_sum, dlogits, correct)."""
    z = logits.float()
    lse = torch.logsumexp(z, dim=-1)
    loss = lse - z.gather(-1, labels.view(-1, 1).long()).squeeze(-1)
    p = torch.softmax(z, dim=-1)
    onehot = F.one_hot(labels.long(), z.shape[-1]).float()
    scale = 1.0 / z.shape[0] if grad_scale is None else grad_scale
    d = (p - onehot) * scale
    correct = (z.argmax(-1) == labels).sum().float()
    return loss.sum(), d, correct


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, grad_scale, stats, add_count, unit_grad):
        global _launches
        B, C = logits.shape
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        # same (possibly padded) row pitch as the logits: the gradient goes into the weight- and
        # input-gradient GEMMs as it is
        dlogits = torch.empty_strided(logits.shape, logits.stride(), dtype=logits.dtype, device=logits.device)
        loss = torch.zeros((), dtype=torch.float32, device=logits.device)
        lib = _lib.kernels()
        lib.cb2_softmax_xent_ex.restype = ctypes.c_int
        lib.cb2_softmax_xent_ex.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 4 + \
            [ctypes.c_int] * 3 + [ctypes.c_float, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        dt = 1 if logits.dtype == torch.bfloat16 else 0
        rc = lib.cb2_softmax_xent_ex(logits.data_ptr(), dt, labels.data_ptr(), dlogits.data_ptr(), None,
                                     None if stats is None else stats.data_ptr(), B, C, logits.stride(0),
                                     float(grad_scale), loss.data_ptr(),
                                     int(bool(add_count) and stats is not None),
                                     torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "softmax_xent")
        _launches += 1
        ctx.save_for_backward(dlogits)
        ctx.grad_scale = grad_scale
        ctx.unit_grad = bool(unit_grad)
        # mean loss over the *global* batch this call was scaled for
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        if ctx.unit_grad:        # the caller differentiates the loss itself: d loss / d loss = 1
            return dlogits, None, None, None, None, None
        return dlogits * g.to(dlogits.dtype), None, None, None, None, None


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor,
                          grad_scale: float | None = None,
                          stats: torch.Tensor | None = None, count: bool = False,
                          unit_grad: bool = False) -> torch.Tensor:
    """Mean sparse softmax cross-entropy with fused backward.

    ``grad_scale`` defaults to ``1/batch`` (per-replica callers pass ``1/global_batch``).  If
    ``stats`` (fp32[>=2], device) is given, the summed loss and the number of correct arg-max
    predictions are *accumulated* into ``stats[0]`` and ``stats[1]`` (and, with ``count``, the number
    of rows into ``stats[2]``).  ``unit_grad``: the caller promises to call ``backward()`` on this
    loss (or on a sum it is a plain term of), so the stored gradient is returned without a multiply.
    """
    if logits.dim() != 2:
        logits = logits.reshape(-1, logits.shape[-1])
        labels = labels.reshape(-1)
    if grad_scale is None:
        grad_scale = 1.0 / logits.shape[0]
    if logits.is_cuda:
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits = logits.float()
        if labels.dtype != torch.int64:
            labels = labels.long()
        return _XentFn.apply(logits, labels, float(grad_scale), stats, count, unit_grad)
    z = logits.float()
    loss_sum = F.cross_entropy(z, labels.long(), reduction="sum")
    if stats is not None:
        with torch.no_grad():
            stats[0] += loss_sum.detach()
            stats[1] += (z.argmax(-1) == labels).sum().float()
            if count:
                stats[2] += float(z.shape[0])
    return loss_sum * grad_scale
