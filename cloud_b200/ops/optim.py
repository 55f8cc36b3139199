"""Flat-arena optimizers (``csrc/kernels/elementwise.cu::optim_flat_kernel`` and the fused
all-reduce epilogues in ``csrc/kernels/comm.cu``).

All hyper-parameters live in a 16-float device vector so that a whole training step - including
the step counter and Adam bias corrections - can be replayed from a CUDA graph.

Reference workloads: ``optimizer=Adam()`` README.md:78 (K6); ``SGD(0.002)``
experimental/cloud_fit/tests/unit/client_test.py:87-89 and ``RMSprop`` +
``ExponentialDecay`` core/tests/testdata/keras_tuner_cifar_example.py:76 (K7).
"""
from __future__ import annotations

import math

import torch

from . import _lib

OPT_CODES = {"none": 0, "adam": 1, "adam_keras": 1, "adam_torch": 2, "adamw": 3, "sgd": 4,
             "rmsprop": 5}
HYPER_LEN = 16
_launches = 0


def launches() -> int:
    return _launches


def make_hyper(device, *, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, weight_decay=0.0,
               grad_scale=1.0, momentum=0.0, nesterov=False, rms_momentum=0.0) -> torch.Tensor:
    """Device hyper-parameter vector (layout documented in include/cloud_b200_kernels.h)."""
    h = torch.zeros(HYPER_LEN, dtype=torch.float32)
    h[0], h[1], h[2], h[3], h[4] = lr, beta1, beta2, eps, weight_decay
    h[5] = 0.0
    h[6] = 1.0
    h[7] = 1.0
    h[8] = grad_scale
    h[9] = momentum
    h[10] = 1.0 if nesterov else 0.0
    h[11] = rms_momentum
    return h.to(device)


def optim_tick(hyper: torch.Tensor) -> None:
    """step += 1 and refresh the Adam bias corrections, on the device."""
    global _launches
    if hyper.is_cuda:
        rc = _lib.kernels().cb2_optim_tick(hyper.data_ptr(), torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "optim_tick")
        _launches += 1
    else:
        t = float(hyper[5]) + 1.0
        hyper[5] = t
        hyper[6] = 1.0 - float(hyper[1]) ** t
        hyper[7] = 1.0 - float(hyper[2]) ** t


def _reference_update(opt: int, p, g, s1, s2, h):
    """PyTorch restatement of ``opt_update`` (csrc/kernels/optim.cuh); also the CPU path."""
    lr, b1, b2, eps, wd = (float(h[i]) for i in range(5))
    bc1 = float(h[6]) if float(h[6]) > 0 else 1.0
    bc2 = float(h[7]) if float(h[7]) > 0 else 1.0
    g = g.float() * float(h[8])
    mom, nesterov, rms_mom = float(h[9]), float(h[10]), float(h[11])
    if opt == 0:
        return g
    if opt in (1, 2, 3):
        if opt == 3:
            p = p * (1.0 - lr * wd)
        elif wd != 0.0:
            g = g + wd * p
        s1.mul_(b1).add_(g, alpha=1.0 - b1)
        s2.mul_(b2).addcmul_(g, g, value=1.0 - b2)
        if opt == 1:
            return p - (lr * math.sqrt(bc2) / bc1) * s1 / (s2.sqrt() + eps)
        return p - (lr / bc1) * s1 / (s2.sqrt() / math.sqrt(bc2) + eps)
    if wd != 0.0:
        g = g + wd * p
    if opt == 4:
        if mom != 0.0:
            s1.mul_(mom).add_(g)
            g = g + mom * s1 if nesterov else s1
        return p - lr * g
    if opt == 5:
        s1.mul_(mom).addcmul_(g, g, value=1.0 - mom)
        upd = lr * g / torch.sqrt(s1 + eps)
        if rms_mom != 0.0:
            s2.mul_(rms_mom).add_(upd)
            return p - s2
        return p - upd
    raise ValueError(f"unknown optimizer code {opt}")


def optim_flat(opt: int, *, master: torch.Tensor | None, grad: torch.Tensor,
               s1: torch.Tensor | None, s2: torch.Tensor | None,
               param_out: torch.Tensor | None, hyper: torch.Tensor,
               zero_grad: bool = False) -> None:
    """One optimizer step over flat 1-D arenas (in place)."""
    global _launches
    n = grad.numel()
    if grad.is_cuda:
        gd = 1 if grad.dtype == torch.bfloat16 else 0
        pd = 0
        if param_out is not None:
            pd = 1 if param_out.dtype == torch.bfloat16 else 0
        rc = _lib.kernels().cb2_optim_flat(
            opt, None if master is None else master.data_ptr(), grad.data_ptr(), gd,
            None if s1 is None else s1.data_ptr(), None if s2 is None else s2.data_ptr(),
            None if param_out is None else param_out.data_ptr(), pd, n, hyper.data_ptr(),
            int(zero_grad), torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "optim_flat")
        _launches += 1
        return
    with torch.no_grad():
        p = master if master is not None else param_out
        a = s1 if s1 is not None else torch.zeros(0)
        b = s2 if s2 is not None else torch.zeros(0)
        new_p = _reference_update(opt, p.float(), grad, a, b, hyper)
        if master is not None:
            master.copy_(new_p)
        if param_out is not None:
            param_out.copy_(new_p.to(param_out.dtype))
        if zero_grad:
            grad.zero_()


class FlatOptimizerState:
    """fp32 master / moment arenas for one flat parameter group."""

    def __init__(self, n: int, device, opt: int, need_master: bool):
        self.n = n
        self.opt = opt
        z = lambda: torch.zeros(n, dtype=torch.float32, device=device)  # noqa: E731
        self.master = z() if need_master else None
        self.s1 = z() if opt != 0 else None
        self.s2 = z() if opt in (1, 2, 3, 5) else None
