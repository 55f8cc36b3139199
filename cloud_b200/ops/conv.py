"""1x1 convolutions on the tcgen05 GEMM (``csrc/kernels/gemm_tcgen05.cu``).

With channels_last activations a 1x1 / stride-1 convolution IS a matrix product:
``Y[N*H*W, Cout] = X[N*H*W, Cin] @ W[Cout, Cin]^T`` - no im2col, no layout change.  36 of the 53
convolutions of ResNet-50 (and more than half of its FLOPs) have this form, so they run on the
hand-written kernel instead of cuDNN:

* forward: K-major x K-major GEMM; when a BatchNormalization follows, the epilogue also reduces
  the per-channel sum / sum of squares of the tile it is about to store (``colstats``), which makes
  the separate statistics pass over the activation (one full read) unnecessary;
* input gradient: ``dX = dY @ W`` (B operand MN-major, no transpose);
* weight gradient: ``dW = dY^T @ X`` (both operands MN-major) with split-K - the reduction runs over
  N*H*W, up to 200 704 here, while dW has one to a few dozen tiles - summed by TMA reduce-adds
  into an fp32 workspace and rounded into the bf16 gradient arena by the last-arriving warp of
  the same kernel.

Reference workload: ResNet-50 through ``tf.keras.applications`` in
core/tests/examples/call_run_within_script_with_keras_fit.py (K8 of SURVEY.md section 2.5); the
reference leaves these to cuDNN / cuBLAS inside TensorFlow.
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn.functional as F

import importlib
import os

_bn = importlib.import_module(__package__ + ".bn")
_dense = importlib.import_module(__package__ + ".dense")     # the module (ops.dense is also an op)

_FINALIZE_IN_KERNEL = os.environ.get("TFC_WGRAD_FINALIZE", "0") == "1"


def supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation) -> bool:
    """True when ``conv2d(x, w)`` can run as one GEMM on the tcgen05 kernel."""
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 4
            and tuple(w.shape[2:]) == (1, 1) and tuple(stride) == (1, 1) and tuple(dilation) == (1, 1)
            and padding in (0, (0, 0), "valid") and x.shape[1] % 8 == 0 and w.shape[0] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last))


def _as_matrix(t: torch.Tensor) -> torch.Tensor:
    """[N, C, H, W] channels_last -> its [N*H*W, C] row-major matrix (a view)."""
    n, c, h, w = t.shape
    return t.permute(0, 2, 3, 1).reshape(n * h * w, c)


def mode() -> str:
    """Which of the three GEMMs of a 1x1 convolution run on the tcgen05 kernel
    (``TFC_CONV1X1_GEMM``): ``wgrad`` (default) - only the weight gradient, where the split-K
    kernel beats cuBLAS + its split-K reduction on the ResNet-50 shapes (bench/wgrad_probe.py),
    forward and input gradient stay on cuDNN; ``all`` - all three plus batch-norm statistics from
    the forward epilogue; ``0`` - none."""
    v = os.environ.get("TFC_CONV1X1_GEMM", "wgrad").lower()
    return {"1": "all", "true": "all", "off": "0", "false": "0"}.get(v, v)


class _Conv1x1Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, colstats, all_gemm, prod=None):
        n, cin, h, wd = x.shape
        cout = w.shape[0]
        ctx.prod = prod
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        ctx.all_gemm = all_gemm
        if not all_gemm:
            return F.conv2d(x, w, None if bias is None else bias.to(x.dtype))
        w2 = w.reshape(cout, cin)
        y = torch.empty((n, h, wd, cout), dtype=torch.bfloat16, device=x.device)
        _dense.gemm(_as_matrix(x), w2, bias=bias, out=y.view(n * h * wd, cout), colstats=colstats)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        n, cin, h, wd = x.shape
        cout = w.shape[0]
        m = n * h * wd
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dy2 = _as_matrix(dy)
        dx = dw = db = None
        g_next = ctx.prod.pop("g_next", None) if ctx.prod is not None else None
        if ctx.needs_input_grad[0] and g_next is not None:
            # x is one handle of a folded conv+BN(+shortcut)+ReLU output z whose other handle is the
            # shortcut of the block this convolution belongs to.  dL/dz = dy W + g_next, and the
            # producer wants it masked by its ReLU bits, with its column sums: all of that is the
            # epilogue of this input-gradient GEMM (no separate pass over the largest tensor of the block)
            convbn = importlib.import_module(__package__ + ".convbn")
            G = torch.empty((n, h, wd, cin), dtype=torch.bfloat16, device=x.device)
            gsums = _bn._scratch(16 * cin, x.device)
            convbn.fused_gemm(dy2, w.reshape(cout, cin), G.view(m, cin), b_mn=True, residual=g_next,
                              mask_in=ctx.prod["mask"], colstats=gsums)
            ctx.prod["premasked"] = gsums
            dx = G.permute(0, 3, 1, 2)
        elif ctx.needs_input_grad[0] and not ctx.all_gemm:
            dx = torch.ops.aten.convolution_backward(
                dy, x, w, None, (1, 1), (0, 0), (1, 1), False, (0, 0), 1, (True, False, False))[0]
        elif ctx.needs_input_grad[0]:
            dxm = torch.empty((n, h, wd, cin), dtype=torch.bfloat16, device=x.device)
            _dense.gemm(dy2, w.reshape(cout, cin), b_mn=True, out=dxm.view(m, cin))
            dx = dxm.permute(0, 3, 1, 2)
        if ctx.needs_input_grad[1]:
            direct = getattr(w, "_cb2_direct_grad", False) and w.grad is not None
            if _FINALIZE_IN_KERNEL:
                # the kernel itself rounds the finished split-K sum into the bf16 destination
                dst = w.grad.view(cout, cin) if direct else torch.zeros(
                    (cout, cin), dtype=torch.bfloat16, device=x.device)
                _dense.gemm_splitk_bf16(dy2, _as_matrix(x), dst)
                if not direct:
                    dw = dst.view(cout, cin, 1, 1)
            else:
                # split-K slices meet in an fp32 workspace that the engine zeroes once per step;
                # one add rounds it into the gradient arena (measured 1% faster end to end than
                # the in-kernel finalize, whose serial tail costs more than the extra launch)
                ws = _bn._scratch(cout * cin, x.device).view(cout, cin)
                _dense.gemm(dy2, _as_matrix(x), a_mn=True, b_mn=True, out=ws, accumulate=True, split_k=0)
                if direct:
                    w.grad.view(cout, cin).add_(ws)
                else:
                    dw = ws.to(torch.bfloat16).view(cout, cin, 1, 1)
            if direct:
                w._cb2_grad_written(w)      # straight into the engine's arena: no autograd add
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.float().sum(0)
        return dx, dw, db, None, None, None


def conv1x1(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None,
            bn_stats: bool = False) -> torch.Tensor:
    """1x1 / stride-1 convolution of a channels_last bf16 tensor.  With ``bn_stats`` the result
    carries ``_cb2_bn_sums``: the batch statistics a following fused batch-norm can use as is."""
    if not x.is_cuda:
        return F.conv2d(x, w, bias)
    all_gemm = mode() == "all"
    sums = None
    if all_gemm and bn_stats and bias is None:
        cout = w.shape[0]
        sums = _bn._scratch(16 * cout, x.device)
    if bias is not None and bias.dtype != torch.float32:
        bias = bias.float()
    # the input may be one handle of a forked, folded conv+BN output (ops/convbn.py): tell the
    # producer that this convolution can deliver its masked gradient (see backward)
    prod = getattr(x, "_cb2_producer", None)
    if prod is not None and not ("mask" in prod and x.shape[1] % 64 == 0 and w.shape[0] % 64 == 0
                                 and torch.is_grad_enabled()
                                 and os.environ.get("TFC_BN_FOLD_FUSE_G", "1") != "0"):
        prod = None
    if prod is not None:
        prod["dgrad_consumer"] = True
    y = _Conv1x1Fn.apply(x, w, bias, sums, all_gemm, prod)
    if sums is not None:
        y._cb2_bn_sums = sums
    return y


# ----------------------------------------------------------------------------------------------
# ResNet stem: 7x7 / stride 2 / pad 3 on a 3-channel image (cb2_conv_stem_* in gemm_tcgen05.cu)
# ----------------------------------------------------------------------------------------------
def stem_supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation) -> bool:
    """True for the ImageNet stem geometry on a CUDA bf16 channels_last image that needs no input
    gradient (cuDNN takes 2.5 ms for this layer at batch 256; the implicit GEMM ~0.3 ms)."""
    return (mode() != "0" and x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
            and x.dim() == 4 and tuple(w.shape) == (64, 3, 7, 7) and tuple(stride) == (2, 2)
            and padding in (3, (3, 3)) and tuple(dilation) == (1, 1) and not x.requires_grad
            and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0 and x.shape[3] // 2 <= 128)


def space_to_depth_stem(x: torch.Tensor) -> torch.Tensor:
    """[B, 3, H, W] (channels_last or not) -> [B, H/2+3, W/2+3, 16]: pixel (r, s) holds input pixels
    (2r-4+a, 2s-4+b) as channels (a*2+b)*3 + c; borders and channels 12..15 are zero."""
    b, c, h, w = x.shape
    p = F.pad(x, (4, 2, 4, 2)).permute(0, 2, 3, 1)                  # NHWC view, [B, H+6, W+6, 3]
    r, s = (h + 6) // 2, (w + 6) // 2
    p = p.reshape(b, r, 2, s, 2, c).permute(0, 1, 3, 2, 4, 5).reshape(b, r, s, 4 * c)
    return F.pad(p, (0, 16 - 4 * c)).contiguous()


def stem_weight_matrix(w: torch.Tensor) -> torch.Tensor:
    """Kernel [64, 3, 7, 7] -> [64, 256] in the operand order of ``cb2_conv_stem_forward``:
    k = i*64 + j*16 + (a*2+b)*3 + c for kernel row 2i+a-1, column 2j+b-1 (differentiable)."""
    co, c = w.shape[0], w.shape[1]
    w8 = F.pad(w, (1, 0, 1, 0))                                     # tap -1 (zero) in front: 8 x 8
    w8 = w8.reshape(co, c, 4, 2, 4, 2).permute(0, 2, 4, 3, 5, 1)    # co, i, j, a, b, c
    return F.pad(w8.reshape(co, 4, 4, 4 * c), (0, 16 - 4 * c)).reshape(co, 256).contiguous()


class _StemFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xs, wm):
        b, r, s, _ = xs.shape
        oh, ow = r - 3, s - 3
        y = torch.empty((b, oh, ow, 64), dtype=torch.bfloat16, device=xs.device)
        rc = _dense._lib.kernels().cb2_conv_stem_forward(xs.data_ptr(), wm.data_ptr(), y.data_ptr(),
                                                         b, oh, ow, _dense._stream())
        _dense._lib.check(rc, "conv_stem_forward")
        _dense._launches += 1
        ctx.save_for_backward(xs)
        ctx.geom = (b, oh, ow)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        (xs,) = ctx.saved_tensors
        b, oh, ow = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dwt = _bn._scratch(256 * 64, xs.device).view(256, 64)       # zeroed once per step by the engine
        rc = _dense._lib.kernels().cb2_conv_stem_wgrad(xs.data_ptr(), dy.data_ptr(), dwt.data_ptr(),
                                                       b, oh, ow, _dense._stream())
        _dense._lib.check(rc, "conv_stem_wgrad")
        _dense._launches += 1
        return None, dwt.t().to(torch.bfloat16)


def stem_conv_s2d(xs: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """The stem convolution of an image already in the space-to-depth layout
    (``ops.normalize_u8_images_s2d``)."""
    y = _StemFn.apply(xs, stem_weight_matrix(w))
    if bias is not None:
        y = y + bias.to(y.dtype).view(1, -1, 1, 1)
    return y


def stem_conv(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """``conv2d(x, w, stride=2, padding=3)`` for the ResNet stem on the tcgen05 implicit GEMM."""
    y = _StemFn.apply(space_to_depth_stem(x), stem_weight_matrix(w))
    if bias is not None:
        y = y + bias.to(y.dtype).view(1, -1, 1, 1)
    return y


# ----------------------------------------------------------------------------------------------
# k x k / stride-1 / 'same' convolutions: weight gradient on the tcgen05 kernel (cb2_conv_wgrad:
# shifted pixel boxes through 4-D TMA maps); 3x3 forward and input gradient on the implicit-GEMM
# kernel cb2_conv3x3 (halo strip loaded once per 64 channels, nine tap descriptors into it), other
# kernel sizes and the shapes conv3x3_own() turns down on cuDNN
# ----------------------------------------------------------------------------------------------
def kxk_supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation) -> bool:
    k = w.shape[2]
    return (mode() != "0" and x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
            and x.dim() == 4 and w.shape[2] == w.shape[3] and k in (3, 5, 7) and tuple(stride) == (1, 1)
            and tuple(dilation) == (1, 1) and padding in (k // 2, (k // 2, k // 2))
            and x.shape[1] % 64 == 0 and w.shape[0] % 64 == 0 and x.shape[3] <= 64
            and x.is_contiguous(memory_format=torch.channels_last))


_KXK_WGRAD = os.environ.get("TFC_CONV_KXK_WGRAD", "0") == "1"   # k x k weight gradients on cb2_conv_wgrad
_CONV3 = os.environ.get("TFC_CONV3X3", "1")          # "0": cuDNN forward / input gradient for 3x3 too
_CONV3_MIN_W = int(os.environ.get("TFC_CONV3X3_MIN_W", "14"))
_CONV3_MIN_C = int(os.environ.get("TFC_CONV3X3_MIN_C", "64"))
_conv3_declared = False


def _conv3_lib():
    global _conv3_declared
    lib = _dense._lib.kernels()
    if not _conv3_declared:
        vp, i = ctypes.c_void_p, ctypes.c_int
        lib.cb2_conv3x3.restype = i
        lib.cb2_conv3x3.argtypes = [vp, vp, vp, i, i, i, i, i, i, vp, vp]
        _conv3_declared = True
    return lib


def conv3x3_own(x: torch.Tensor, w: torch.Tensor) -> bool:
    """Whether the 3x3 forward / input gradient of this layer run on ``cb2_conv3x3``
    (csrc/kernels/conv3x3_tcgen05.cu).  The kernel computes on 128-row MMA tiles made of whole image
    rows of W + 2 padded pixels: below 14 columns (the 7x7 stage of ResNet-50) more than half of every
    tile would be padding (measured: bench/conv3_probe.py, 2x cuDNN's time) - those shapes stay on
    cuDNN.  With 64 channels the MMAs are short (N = 64) and the forward pass is ~20 % behind cuDNN,
    which the fused statistics (no separate pass over the 56 x 56 activation) more than pay back."""
    if _CONV3 == "0" or w.shape[2] != 3 or not w.is_contiguous(memory_format=torch.channels_last):
        return False
    wd, cin, cout = x.shape[3], x.shape[1], w.shape[0]
    if _CONV3 == "all":
        return 6 <= wd <= 62
    return _CONV3_MIN_W <= wd <= 62 and cin >= _CONV3_MIN_C and cout >= _CONV3_MIN_C


def _conv3_run(src, w, n_out, input_gradient, sums=None):
    """src: channels_last [B, C, H, W] bf16 (NHWC in memory) -> channels_last [B, n_out, H, W]."""
    b, _, h, wd = src.shape
    cout, cin = w.shape[0], w.shape[1]
    out = torch.empty((b, n_out, h, wd), device=src.device, dtype=torch.bfloat16,
                      memory_format=torch.channels_last)
    rc = _conv3_lib().cb2_conv3x3(src.data_ptr(), w.data_ptr(), out.data_ptr(), b, h, wd, cin, cout,
                                  int(input_gradient), 0 if sums is None else sums.data_ptr(), _dense._stream())
    _dense._lib.check(rc, "conv3x3")
    _dense._launches += 1
    return out


class _ConvKxKFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, sums=None):
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        k = w.shape[2]
        ctx.own = bias is None and conv3x3_own(x, w)
        if ctx.own:
            return _conv3_run(x, w, w.shape[0], False, sums)
        return F.conv2d(x, w, None if bias is None else bias.to(x.dtype), padding=k // 2)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        n, cin, h, wd = x.shape
        cout, k = w.shape[0], w.shape[2]
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if ctx.own:
                dx = _conv3_run(dy, w, cin, True)
            else:
                dx = torch.ops.aten.convolution_backward(
                    dy, x, w, None, (1, 1), (k // 2, k // 2), (1, 1), False, (0, 0), 1, (True, False, False))[0]
        if ctx.needs_input_grad[1] and not _KXK_WGRAD:
            # cuDNN's weight gradient (the shifted-box GEMM below is ~2x slower on these shapes: opt-in)
            dw = torch.ops.aten.convolution_backward(
                dy, x, w, None, (1, 1), (k // 2, k // 2), (1, 1), False, (0, 0), 1, (False, True, False))[1]
        elif ctx.needs_input_grad[1]:
            ws = _bn._scratch(k * k * cin * cout, x.device).view(k * k * cin, cout)   # zeroed per step
            rc = _dense._lib.kernels().cb2_conv_wgrad(x.data_ptr(), dy.data_ptr(), ws.data_ptr(), n, h, wd,
                                                      cin, cout, k, k // 2, 1, _dense._stream())
            _dense._lib.check(rc, "conv_wgrad")
            _dense._launches += 1
            # [(i, j, ci), co] -> the parameter's [co, ci, i, j] (channels_last memory = [co, i, j, ci])
            g = ws.view(k, k, cin, cout).permute(3, 2, 0, 1)
            if getattr(w, "_cb2_direct_grad", False) and w.grad is not None:
                w.grad.add_(g)
                w._cb2_grad_written(w)
            else:
                dw = g.to(torch.bfloat16)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 2, 3))
        return dx, dw, db, None


def conv_kxk(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None,
             bn_stats: bool = False) -> torch.Tensor:
    """k x k / stride-1 / 'same' convolution of a channels_last bf16 tensor.  3x3 layers that
    ``conv3x3_own`` accepts run forward and input gradient on the implicit-GEMM kernel; with
    ``bn_stats`` its epilogue also reduces the batch statistics of the result for the batch-norm
    that follows (``_cb2_bn_sums``, as ``conv1x1`` does)."""
    sums = None
    if bn_stats and bias is None and conv3x3_own(x, w):
        sums = _bn._scratch(16 * w.shape[0], x.device)
    y = _ConvKxKFn.apply(x, w, bias, sums)
    if sums is not None:
        y._cb2_bn_sums = sums
    return y


# ----------------------------------------------------------------------------------------------
# Strided 1x1 convolutions = a 1x1 convolution of every s-th pixel
# ----------------------------------------------------------------------------------------------
def _subsample_kernel(src, dst, b, h, w, c, s, backward):
    lib = _dense._lib.kernels()
    lib.cb2_subsample_nhwc.restype = ctypes.c_int
    lib.cb2_subsample_nhwc.argtypes = [ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int] * 6 + [ctypes.c_void_p]
    rc = lib.cb2_subsample_nhwc(src.data_ptr(), dst.data_ptr(), b, h, w, c, s, int(backward), _dense._stream())
    _dense._lib.check(rc, "subsample_nhwc")
    _dense._launches += 1


class _SubsampleFn(torch.autograd.Function):
    """``x[:, :, ::s, ::s]`` as a dense channels_last tensor; backward writes the full-size gradient
    in one pass (zeros between the sampled pixels), channels_last like everything around it."""

    @staticmethod
    def forward(ctx, x, s):
        ctx.shape, ctx.s = tuple(x.shape), s
        b, c, h, w = x.shape
        fast = (x.is_cuda and x.dtype == torch.bfloat16 and c % 8 == 0
                and x.is_contiguous(memory_format=torch.channels_last))
        ctx.fast = fast
        if not fast:
            return x[:, :, ::s, ::s].contiguous(memory_format=torch.channels_last)
        y = torch.empty((b, c, -(-h // s), -(-w // s)), dtype=x.dtype, device=x.device,
                        memory_format=torch.channels_last)
        _subsample_kernel(x, y, b, h, w, c, s, False)
        return y

    @staticmethod
    def backward(ctx, g):
        b, c, h, w = ctx.shape
        s = ctx.s
        dx = torch.empty(ctx.shape, dtype=g.dtype, device=g.device, memory_format=torch.channels_last)
        if ctx.fast and g.dtype == torch.bfloat16:
            if not g.is_contiguous(memory_format=torch.channels_last):
                g = g.contiguous(memory_format=torch.channels_last)
            _subsample_kernel(g, dx, b, h, w, c, s, True)
        else:
            dx.zero_()
            dx[:, :, ::s, ::s].copy_(g)
        return dx, None


def subsample(x: torch.Tensor, s: int) -> torch.Tensor:
    """Every ``s``-th pixel of a channels_last tensor, computed once per tensor: the two strided 1x1
    convolutions at the head of a ResNet stage (``*_0_conv`` and ``*_1_conv``) then run as plain
    GEMMs on a quarter of the rows, and their input gradients meet in ONE scatter instead of two
    strided-dgrad kernels over the full-resolution tensor."""
    shared = getattr(x, "_cb2_shared", None)
    if shared is None:
        shared = {}
        try:
            x._cb2_shared = shared
        except AttributeError:
            pass
    key = ("sub", s)
    if key not in shared:
        shared[key] = _SubsampleFn.apply(x, s)
    return shared[key]
