"""Fused NHWC batch normalisation (+ residual add) (+ ReLU): ``csrc/kernels/bn.cu``.

Forward: one statistics pass + one apply pass (instead of cuDNN BN, add and ReLU kernels each
re-reading the activation); backward: one reduce pass + one apply pass with the ReLU mask and the
shortcut gradient folded in.  Reference workload: ResNet-50's 53 BatchNormalization layers
(core/tests/examples/call_run_within_script_with_keras_fit.py:79-83, K8).
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _lib

_launches = 0
_declared = False


def launches() -> int:
    return _launches


def _k():
    global _declared
    lib = _lib.kernels()
    if not _declared:
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        lib.cb2_bn_supported.restype = i
        lib.cb2_bn_supported.argtypes = [i]
        lib.cb2_bn_forward.restype = i
        lib.cb2_bn_scratch_floats.restype = i
        lib.cb2_bn_scratch_floats.argtypes = [i]
        lib.cb2_bn_forward.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, f, i, i, vp]
        lib.cb2_bn_forward_mask.restype = i
        lib.cb2_bn_forward_mask.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, f, i, i, vp]
        lib.cb2_bn_backward3.restype = i
        lib.cb2_bn_backward3.argtypes = [vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp]
        lib.cb2_bn_backward2.restype = i
        lib.cb2_bn_backward2.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp]
        lib.cb2_bn_backward.restype = i
        lib.cb2_bn_backward.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp]
        _declared = True
    return lib


def bn_available() -> bool:
    if os.environ.get("TFC_DISABLE_FUSED_BN") == "1":
        return False
    try:
        _k()
        return True
    except _lib.KernelsUnavailable:
        return False


def _stream():
    return torch.cuda.current_stream().cuda_stream


# One zero-filled scratch arena per device, handed out slice by slice and re-zeroed with a single
# memset at the start of every training step (DPEngine.begin_step): replaces two torch.zeros
# launches per BatchNormalization layer per step (106 for ResNet-50).
_pool = {}


_SIGN_BITS = os.environ.get("TFC_BN_SIGN_BITS", "1") != "0"


def reset_scratch(device) -> None:
    """Start of a step: zero what the previous step used and rewind."""
    st = _pool.get(device)
    if st is None:
        return
    if st["used"]:
        st["buf"][:st["used"]].zero_()
    st["used"] = 0


def _scratch(n: int, device) -> torch.Tensor:
    st = _pool.get(device)
    if st is None:
        st = {"buf": torch.zeros(32 << 20, dtype=torch.float32, device=device), "used": 0,
              "armed": False}
        _pool[device] = st
    if not st["armed"] or st["used"] + n > st["buf"].numel():
        return torch.zeros(n, dtype=torch.float32, device=device)    # no engine driving the pool
    out = st["buf"][st["used"]:st["used"] + n]
    st["used"] += (n + 63) // 64 * 64
    return out


def arm_scratch(device) -> None:
    """Called by the engine once it promises to call ``reset_scratch`` every step."""
    _scratch(0, device)
    _pool[device]["armed"] = True


def _is_nhwc(t):
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last)


class _BNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, run_mean, run_var, momentum, eps, relu, mean_bias,
                sums=None, mailbox=None):
        # `mailbox`: a dict shared with a `fork` of this layer's output (two consumers): the
        # second consumer's gradient arrives there instead of being added to `dy` by autograd
        ctx.mailbox = mailbox
        global _launches
        N, C, H, W = x.shape
        M = N * H * W
        y = torch.empty_like(x)                       # preserves channels_last
        # `sums`: batch statistics already reduced by the producer of x (ops/conv.py: the GEMM
        # epilogue of a 1x1 convolution) - the statistics pass over x is skipped (training = 2)
        mode = 1 if sums is None else 2
        if sums is None:
            sums = _scratch(_k().cb2_bn_scratch_floats(C), x.device)
        save = torch.empty(2 * C, dtype=torch.float32, device=x.device)
        # shortcut + ReLU: the mask cannot be recomputed from x alone; the forward pass leaves one
        # sign bit per element so that backward does not have to re-read the whole output
        mask = torch.empty(M * C // 8, dtype=torch.uint8, device=x.device) \
            if (relu and residual is not None and _SIGN_BITS) else None
        rc = _k().cb2_bn_forward_mask(x.data_ptr(), None if residual is None else residual.data_ptr(),
                                      y.data_ptr(), sums.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                      run_mean.data_ptr(), run_var.data_ptr(), save.data_ptr(),
                                      save[C:].data_ptr(),
                                      None if mean_bias is None else mean_bias.data_ptr(),
                                      None if mask is None else mask.data_ptr(), M, C,
                                      float(eps), float(momentum), int(relu), mode, _stream())
        ctx.params = (gamma, beta)
        _lib.check(rc, "bn_forward")
        _launches += 3 - mode
        # the ReLU mask is recomputed from x unless a shortcut was added before the ReLU
        keep = mask if mask is not None else (y if (relu and residual is not None) else None)
        ctx.save_for_backward(x, keep, gamma, save, beta)
        ctx.relu, ctx.has_res, ctx.M, ctx.C = relu, residual is not None, M, C
        ctx.y_is_mask = mask is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        global _launches
        x, y, gamma, save, beta = ctx.saved_tensors
        C, M = ctx.C, ctx.M
        if not _is_nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if ctx.has_res and ctx.relu else None
        sums = _scratch(_k().cb2_bn_scratch_floats(C), x.device)
        # gamma / beta gradients go straight into the engine's gradient arena when it exposes one
        # (no temporary, no accumulate kernel); the engine is told through `_cb2_grad_written`
        gparam, bparam = ctx.params
        direct = (getattr(gparam, "_cb2_direct_grad", False) and getattr(bparam, "_cb2_direct_grad", False)
                  and gparam.grad is not None and bparam.grad is not None)
        if direct:
            dbeta_ptr, dgamma_ptr, grads = bparam.grad.data_ptr(), gparam.grad.data_ptr(), None
        else:
            grads = torch.empty(2 * C, dtype=torch.float32, device=x.device)
            dbeta_ptr, dgamma_ptr = grads.data_ptr(), grads[C:].data_ptr()
        dy2 = ctx.mailbox.pop("dy2", None) if ctx.mailbox is not None else None
        if dy2 is not None and not (dy2.shape == dy.shape and dy2.dtype == dy.dtype and _is_nhwc(dy2)):
            dy = dy + dy2                              # odd layout: fall back to an explicit add
            dy2 = None
        rc = _k().cb2_bn_backward3(dy.data_ptr(), None if dy2 is None else dy2.data_ptr(),
                                   None if y is None else y.data_ptr(), int(ctx.y_is_mask), x.data_ptr(),
                                   save.data_ptr(), save[C:].data_ptr(), gamma.data_ptr(),
                                   beta.data_ptr(), sums.data_ptr(), dbeta_ptr, dgamma_ptr,
                                   dx.data_ptr(),
                                   None if dres is None else dres.data_ptr(), M, C, int(ctx.relu),
                                   _stream())
        _lib.check(rc, "bn_backward")
        _launches += 2
        if ctx.has_res and dres is None:
            dres = dy                                 # no ReLU: the shortcut gradient is dy itself
        if direct:
            gparam._cb2_grad_written(gparam)
            bparam._cb2_grad_written(bparam)
            return dx, None, None, dres, None, None, None, None, None, None, None, None
        return dx, grads[C:], grads[:C], dres, None, None, None, None, None, None, None, None


def batch_norm_nhwc(x, gamma, beta, run_mean, run_var, training, momentum, eps, relu=False,
                    residual=None, mean_bias=None, sums=None):
    """``act(BN(x) + residual)`` for bf16 channels_last ``x``; ``momentum`` is the torch-style
    update rate of the running statistics (1 - keras_momentum).  ``mean_bias`` (fp32 [C]) is added
    to the batch mean when the running mean is refreshed: it is the bias of a preceding
    convolution that was skipped in training because batch-norm cancels it exactly.  ``sums``
    (fp32 ``[8, 2, C]``): batch statistics the producer of ``x`` has already reduced."""
    global _launches
    C = x.shape[1]
    ok = (x.is_cuda and x.dtype == torch.bfloat16 and _is_nhwc(x) and _k().cb2_bn_supported(C)
          and gamma.dtype == torch.float32 and (residual is None or (
              residual.dtype == torch.bfloat16 and _is_nhwc(residual) and residual.shape == x.shape)))
    if not ok:
        y = torch.nn.functional.batch_norm(x, run_mean, run_var, gamma, beta, training, momentum, eps)
        if residual is not None:
            y = y + residual
        return torch.relu_(y) if relu else y
    needs_grad = torch.is_grad_enabled() and (
        x.requires_grad or gamma.requires_grad or beta.requires_grad or
        (residual is not None and residual.requires_grad))
    if training and needs_grad:
        mailbox: dict = {}
        y = _BNFn.apply(x, gamma, beta, residual, run_mean, run_var, momentum, eps, relu, mean_bias,
                        sums, mailbox)
        y._cb2_mailbox = mailbox
        return y
    if needs_grad:   # frozen layer inside a trainable network: let autograd handle it
        y = torch.nn.functional.batch_norm(x, run_mean, run_var, gamma, beta, False, momentum, eps)
        if residual is not None:
            y = y + residual
        return torch.relu_(y) if relu else y
    # no autograd: inference from the running statistics (one pass), or statistics + apply
    N, _C, H, W = x.shape
    y = torch.empty_like(x)
    sums_ptr = save_ptr = save2_ptr = None
    if training:
        sums = _scratch(_k().cb2_bn_scratch_floats(C), x.device)
        save = torch.empty(2 * C, dtype=torch.float32, device=x.device)
        sums_ptr, save_ptr, save2_ptr = sums.data_ptr(), save.data_ptr(), save[C:].data_ptr()
    rc = _k().cb2_bn_forward(x.data_ptr(), None if residual is None else residual.data_ptr(),
                             y.data_ptr(), sums_ptr, gamma.data_ptr(), beta.data_ptr(),
                             run_mean.data_ptr(), run_var.data_ptr(), save_ptr, save2_ptr,
                             None if mean_bias is None else mean_bias.data_ptr(),
                             N * H * W, C, float(eps), float(momentum), int(relu), int(training),
                             _stream())
    _lib.check(rc, "bn_forward")
    _launches += 2 if training else 1
    return y


class _ForkFn(torch.autograd.Function):
    """Identity with two outputs for a fused batch-norm output that feeds two consumers.  In
    backward the second gradient goes into the producer's mailbox (its backward kernels read both
    gradient streams and sum them in fp32) instead of being added by an elementwise kernel."""

    @staticmethod
    def forward(ctx, a, mailbox):
        ctx.mailbox = mailbox
        # a handle nobody differentiates through must arrive as None in backward, not as a
        # zero-filled tensor of the activation's size (in the default, non channels_last layout)
        ctx.set_materialize_grads(False)
        return a.view_as(a), a.view_as(a)

    @staticmethod
    def backward(ctx, g1, g2):
        if g1 is None or g2 is None:
            return (g1 if g2 is None else g2), None
        if not _is_nhwc(g1):
            g1 = g1.contiguous(memory_format=torch.channels_last)
        ctx.mailbox["dy2"] = g2
        return g1, None


def fork(a: torch.Tensor):
    """Two handles on ``a`` whose gradients are merged inside the producing batch-norm's backward
    kernels; plain ``(a, a)`` when ``a`` did not come from the fused batch-norm with autograd on."""
    mb = getattr(a, "_cb2_mailbox", None)
    if mb is None or not a.requires_grad:
        return a, a
    h1, h2 = _ForkFn.apply(a, mb)
    # one dictionary behind both handles: work derived from the tensor (its stride-2 subsample, read
    # by the two strided 1x1 convolutions of a stage's first block) is done once
    h1._cb2_shared = h2._cb2_shared = {}
    # the consumers may talk to the producer (ops/convbn.py: the 1x1 convolution that reads one
    # handle can deliver the producer's masked gradient itself)
    h1._cb2_producer = h2._cb2_producer = mb
    return h1, h2
