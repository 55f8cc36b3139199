"""Memory-bound helper ops: ReLU-backward + bias gradient, Philox dropout, input normalisation.

Kernels: ``csrc/kernels/elementwise.cu``.  Reference workloads: ``Dropout(0.2)`` README.md:73 (K5),
bias gradients of every Dense layer (K3), the ``tf.data`` map that scales images to floats
(core/tests/testdata/mnist_example_using_fit.py:49-51, K12).
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_launches = 0


def launches() -> int:
    return _launches


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def relu_bwd_bias_grad(dy: torch.Tensor, y: torch.Tensor | None, want_db: bool = True,
                       db_dtype: torch.dtype = torch.float32, scale: float = 1.0,
                       db_out: torch.Tensor | None = None):
    """Returns ``(dz, db)`` with ``dz = scale * dy * (y > 0)`` (``dz = scale * dy`` if ``y`` is None)
    and ``db = dz.sum(0)``; one pass over ``dy``.  ``db_out`` (fp32): the sums are ADDED into it
    (a gradient arena) and ``db`` comes back as None.  Row pitches may be padded (unit inner stride)."""
    global _launches
    M, N = dy.shape
    if not dy.is_cuda:
        dz = dy if y is None else torch.where(y > 0, dy, torch.zeros_like(dy))
        if scale != 1.0:
            dz = dz * scale
        db = dz.float().sum(0).to(db_dtype) if want_db else None
        if db_out is not None and db is not None:
            db_out.add_(db)
            db = None
        return dz, db
    if y is None and not want_db and scale == 1.0:
        return dy, None
    if dy.dtype != torch.bfloat16:
        raise TypeError("relu_bwd_bias_grad expects bfloat16 gradients")
    if dy.stride(1) != 1 or (y is not None and (y.stride(1) != 1 or y.stride(0) != dy.stride(0))):
        dy = dy.contiguous()
        y = None if y is None else y.contiguous()
    write_dz = y is not None or scale != 1.0
    dz = torch.empty_strided(dy.shape, dy.stride(), dtype=dy.dtype, device=dy.device) if write_dz else dy
    db = None
    if want_db:
        db = db_out if db_out is not None else torch.empty(N, dtype=db_dtype, device=dy.device)
    lib = _lib.kernels()
    lib.cb2_relu_bwd_bias_grad_ex.restype = ctypes.c_int
    lib.cb2_relu_bwd_bias_grad_ex.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_float,
                                                                                       ctypes.c_void_p]
    rc = lib.cb2_relu_bwd_bias_grad_ex(
        dy.data_ptr(), None if y is None else y.data_ptr(), dz.data_ptr() if write_dz else None,
        None if db is None else db.data_ptr(), 0 if (db is None or db.dtype == torch.float32) else 1, M, N,
        dy.stride(0), 1 if db_out is not None else 0, float(scale), _stream())
    _lib.check(rc, "relu_bwd_bias_grad")
    _launches += 1
    return dz, (None if db_out is not None else db)


_step = {"tensor": None}


def set_step_tensor(t) -> None:
    """Device float holding the current training step (the engine's hyper[5]); dropout mixes it
    into the Philox counter so graph replays do not repeat masks."""
    _step["tensor"] = t


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed, offset):
        ctx.p, ctx.seed, ctx.offset = p, seed, offset
        ctx.step = _step["tensor"] if (_step["tensor"] is not None and
                                       _step["tensor"].device == x.device) else None
        return _dropout_apply(x, p, seed, offset, ctx.step)

    @staticmethod
    def backward(ctx, dy):
        # same (seed, offset, step) => same mask: nothing saved between forward and backward
        return _dropout_apply(dy.contiguous(), ctx.p, ctx.seed, ctx.offset, ctx.step), None, None, None


def _dropout_apply(x, p, seed, offset, step=None):
    global _launches
    xc = x if x.is_contiguous() else x.contiguous()
    y = torch.empty_like(xc)
    if xc.dtype == torch.float32:
        dt = 0
    elif xc.dtype == torch.bfloat16:
        dt = 1
    else:
        raise TypeError("dropout supports float32 and bfloat16")
    rc = _lib.kernels().cb2_dropout(xc.data_ptr(), y.data_ptr(), xc.numel(), float(p),
                                    ctypes.c_uint64(seed), ctypes.c_uint64(offset),
                                    None if step is None else step.data_ptr(), dt, _stream())
    _lib.check(rc, "dropout")
    _launches += 1
    return y


_dropout_counter = {"offset": 0}


def dropout_plan(numel: int, p: float, seed: int | None = None) -> tuple:
    """``(rate, seed, offset)`` for a dropout over ``numel`` elements, taking its stretch of the
    Philox counter exactly like ``dropout`` would (for kernels that apply the mask themselves)."""
    if seed is None:
        seed = torch.initial_seed() & 0xFFFFFFFFFFFFFFFF
    offset = _dropout_counter["offset"]
    _dropout_counter["offset"] += (numel + 3) // 4
    return float(p), int(seed), int(offset)


def dropout(x: torch.Tensor, p: float, training: bool = True, seed: int | None = None,
            offset: int | None = None) -> torch.Tensor:
    """Inverted dropout.  CUDA: Philox mask recomputed in backward (no mask tensor)."""
    if not training or p <= 0.0:
        return x
    if not x.is_cuda:
        return torch.nn.functional.dropout(x, p, training=True)
    if seed is None:
        seed = torch.initial_seed() & 0xFFFFFFFFFFFFFFFF
    if offset is None:
        offset = _dropout_counter["offset"]
        _dropout_counter["offset"] += (x.numel() + 3) // 4
    return _DropoutFn.apply(x, float(p), int(seed), int(offset))


_IMAGENET_MEAN = (0.485, 0.456, 0.406)
_IMAGENET_STD = (0.229, 0.224, 0.225)


def normalize_u8_images(images: torch.Tensor, mean=_IMAGENET_MEAN, std=_IMAGENET_STD,
                        out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8 NHWC ``[N,H,W,C]`` -> bf16 NHWC ``(x/255 - mean) / std`` (channels_last memory)."""
    global _launches
    if images.dtype != torch.uint8:
        raise TypeError("normalize_u8_images expects uint8")
    C = images.shape[-1]
    if not images.is_cuda:
        m = torch.tensor(mean, dtype=torch.float32)[:C] if C <= 3 else torch.tensor(mean).repeat(C)[:C]
        s = torch.tensor(std, dtype=torch.float32)[:C] if C <= 3 else torch.tensor(std).repeat(C)[:C]
        r = ((images.float() / 255.0 - m) / s).to(torch.bfloat16)
        if out is not None:
            out.copy_(r)
            return out
        return r
    images = images.contiguous()
    if out is None:
        out = torch.empty(images.shape, dtype=torch.bfloat16, device=images.device)
    arr = ctypes.c_float * 3
    m = arr(*[float(v) for v in mean])
    s = arr(*[1.0 / float(v) for v in std])
    n_pixels = images.numel() // C
    rc = _lib.kernels().cb2_u8_to_bf16_normalize(images.data_ptr(), out.data_ptr(), n_pixels, C, m, s,
                                                 _stream())
    _lib.check(rc, "u8_to_bf16_normalize")
    _launches += 1
    return out


def normalize_u8_images_s2d(images: torch.Tensor, mean=_IMAGENET_MEAN, std=_IMAGENET_STD) -> torch.Tensor:
    """uint8 NHWC RGB ``[N,H,W,3]`` -> normalised bf16 ``[N, H/2+3, W/2+3, 16]`` in the space-to-depth
    layout of the ResNet stem (``ops.stem_conv_s2d``): normalisation, zero padding and the 2x2
    regrouping in one pass over the bytes."""
    global _launches
    if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[-1] != 3:
        raise TypeError("normalize_u8_images_s2d expects uint8 [N, H, W, 3]")
    n, h, w, _ = images.shape
    if not images.is_cuda:
        from .conv import space_to_depth_stem
        return space_to_depth_stem(normalize_u8_images(images, mean, std).permute(0, 3, 1, 2))
    images = images.contiguous()
    out = torch.empty((n, h // 2 + 3, w // 2 + 3, 16), dtype=torch.bfloat16, device=images.device)
    arr = ctypes.c_float * 3
    rc = _lib.kernels().cb2_u8_to_s2d_bf16(images.data_ptr(), out.data_ptr(), n, h, w,
                                           arr(*[float(v) for v in mean]),
                                           arr(*[1.0 / float(v) for v in std]), _stream())
    _lib.check(rc, "u8_to_s2d_bf16")
    _launches += 1
    return out
