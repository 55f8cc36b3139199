"""NHWC max pooling on the hand-written kernels (``csrc/kernels/pool.cu``): forward keeps a
one-byte arg-max per output element, backward is a gather (no atomics).  Zero padding is implicit,
which is exactly Keras' ``ZeroPadding2D`` + ``MaxPooling2D('valid')`` pair of the ResNet stem
(K8 in SURVEY.md section 2.5)."""
from __future__ import annotations

import ctypes

import torch
import torch.nn.functional as F

from . import _lib

_launches = 0
_declared = False


def launches() -> int:
    return _launches


def _k():
    global _declared
    lib = _lib.kernels()
    if not _declared:
        vp, i = ctypes.c_void_p, ctypes.c_int
        lib.cb2_maxpool_nhwc_forward.restype = i
        lib.cb2_maxpool_nhwc_forward.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        lib.cb2_maxpool_nhwc_backward.restype = i
        lib.cb2_maxpool_nhwc_backward.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        _declared = True
    return lib


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, pad):
        global _launches
        N, C, H, W = x.shape
        OH, OW = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
        y = torch.empty((N, C, OH, OW), dtype=x.dtype, device=x.device,
                        memory_format=torch.channels_last)
        idx = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=x.device)
        rc = _k().cb2_maxpool_nhwc_forward(x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, C,
                                           OH, OW, k, s, pad, torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "maxpool_forward")
        _launches += 1
        ctx.save_for_backward(idx)
        ctx.geom = (N, H, W, C, OH, OW, k, s, pad)
        return y

    @staticmethod
    def backward(ctx, dy):
        global _launches
        (idx,) = ctx.saved_tensors
        N, H, W, C, OH, OW, k, s, pad = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, C, H, W), dtype=dy.dtype, device=dy.device,
                         memory_format=torch.channels_last)
        rc = _k().cb2_maxpool_nhwc_backward(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C,
                                            OH, OW, k, s, pad, torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "maxpool_backward")
        _launches += 1
        return dx, None, None, None


def max_pool_nhwc(x: torch.Tensor, kernel: int, stride: int, pad: int = 0) -> torch.Tensor:
    """Max pooling with implicit ZERO padding of ``pad`` pixels on every side."""
    if (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[1] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last) and 0 <= pad < kernel):
        return _MaxPoolFn.apply(x, int(kernel), int(stride), int(pad))
    if pad:
        x = F.pad(x, (pad, pad, pad, pad))
    return F.max_pool2d(x, kernel, stride)
