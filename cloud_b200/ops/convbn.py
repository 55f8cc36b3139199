"""1x1 convolution + BatchNormalization (+ shortcut) (+ ReLU) folded into GEMMs
(``csrc/kernels/convbn_tcgen05.cu``, ``convbn_aux.cu``).

For the wide 1x1 convolutions of a bottleneck block (Cout = 4 Cin) the convolution output is the
largest tensor of the block, and a separate batch-norm reads and writes it seven more times per
step.  This op never materialises it:

* forward: the batch statistics follow from the Gram matrix and the column sums of the *input*
  (``E[y] = W E[x]``, ``E[y^2] = diag(W E[x x^T] W^T)``); one GEMM then applies scale / shift
  (+ shortcut) (+ ReLU) in its epilogue and writes the block output plus one sign bit per element;
* backward: with ``g`` the upstream gradient masked by those bits, ``P = g^T x`` yields both the
  batch-norm reductions and - after a closed-form correction - the weight gradient; the input
  gradient is one GEMM over two K sources, ``dx = g (s' W) - x (W^T diag(t) W) + bias``.

Per layer and step that is two passes over the big tensor in forward (shortcut in, output out) and
three in backward (mask, two GEMM reads) instead of five and ten.

Reference workload: ``conv*_block*_3_conv`` + ``_3_bn`` + ``_add`` + ``_out`` of Keras' ResNet-50
(core/tests/examples/call_run_within_script_with_keras_fit.py:79-83); the reference leaves them
to cuDNN convolution, batch-norm, add and ReLU kernels inside TensorFlow.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.nn.functional as F

import importlib

from . import _lib

_bn = importlib.import_module(__package__ + ".bn")
_dense = importlib.import_module(__package__ + ".dense")     # the module (ops.dense is also an op)

_declared = False
_launches = 0
_colsums: dict = {}      # data_ptr of a masked gradient -> its column sums, for the next backward in line


def _fuse_g() -> bool:
    return os.environ.get("TFC_BN_FOLD_FUSE_G", "1") != "0"


def launches() -> int:
    return _launches


def _k():
    global _declared
    lib = _lib.kernels()
    if not _declared:
        vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        lib.cb2_convbn_gemm.restype = i
        lib.cb2_convbn_gemm.argtypes = [vp, i, vp, i, i, vp, i, vp, i, i, vp, i, i, i, i, vp, vp, vp, i, i,
                                        vp, vp, vp, i, vp]
        lib.cb2_convbn_fwd_finalize.restype = i
        lib.cb2_convbn_fwd_finalize.argtypes = [vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp,
                                                i, i, i, f, f, i, vp]
        lib.cb2_split_hilo_bf16.restype = i
        lib.cb2_split_hilo_bf16.argtypes = [vp, vp, vp, ctypes.c_size_t, vp]
        lib.cb2_convbn_bwd_finalize.restype = i
        lib.cb2_convbn_bwd_finalize.argtypes = [vp, i, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, vp,
                                                vp, i, i, i, vp]
        lib.cb2_bn_stats.restype = i
        lib.cb2_bn_stats.argtypes = [vp, vp, i, i, vp]
        lib.cb2_bn_mask_grad.restype = i
        lib.cb2_bn_mask_grad.argtypes = [vp, vp, vp, vp, vp, i, i, vp]
        _declared = True
    return lib


def enabled() -> bool:
    return os.environ.get("TFC_BN_FOLD", "1") != "0"


def supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation) -> bool:
    """True when conv(x, w) + batch-norm can run as the folded GEMMs: a 1x1 / stride-1 convolution
    of a channels_last bf16 CUDA tensor whose output is at least twice as wide as its input."""
    if not (enabled() and x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16):
        return False
    if x.dim() != 4 or tuple(w.shape[2:]) != (1, 1) or tuple(stride) != (1, 1) or tuple(dilation) != (1, 1):
        return False
    if padding not in (0, (0, 0), "valid"):
        return False
    cin, cout = x.shape[1], w.shape[0]
    # it pays where the convolution output would not stay in the 126 MB L2 between the kernels that
    # touch it; smaller outputs are cheaper through the separate batch-norm kernels (measured,
    # bench/convbn_probe.py)
    out_mb = x.shape[0] * x.shape[2] * x.shape[3] * cout * 2 / 1e6
    if out_mb < float(os.environ.get("TFC_BN_FOLD_MIN_MB", "100")):
        return False
    return (cin % 64 == 0 and 64 <= cin <= 1024 and cout % 64 == 0 and cout >= 2 * cin
            and x.is_contiguous(memory_format=torch.channels_last) and w.is_contiguous(
                memory_format=torch.channels_last))


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _as_matrix(t: torch.Tensor) -> torch.Tensor:
    n, c, h, w = t.shape
    return t.permute(0, 2, 3, 1).reshape(n * h * w, c)


def fused_gemm(a, b, out, *, b_mn=False, a2=None, b2=None, scale=None, shift=None, residual=None,
               relu=False, mask_in=None, mask_out=None, colstats=None, block_n=0):
    """``out[M,N] = mask_in ? relu?((a @ op(b) + a2 @ op(b2)) * scale + shift + residual) : 0``
    (bf16 operands and output; ``b`` stored ``[N,K]``, or ``[K,N]`` with ``b_mn``).  ``mask_out``
    receives one bit per element (stored value > 0), ``colstats`` ``[8, 2, N]`` the column sums /
    sums of squares of the stored values."""
    global _launches
    M, K1 = a.shape
    N = out.shape[1]
    K2 = 0 if a2 is None else a2.shape[1]
    rc = _k().cb2_convbn_gemm(_ptr(a), a.stride(0), _ptr(b), b.stride(0), int(b_mn), _ptr(a2),
                              0 if a2 is None else a2.stride(0), _ptr(b2), 0 if b2 is None else b2.stride(0),
                              K2, _ptr(out), out.stride(0), M, N, K1, _ptr(scale), _ptr(shift),
                              _ptr(residual), 0 if residual is None else residual.stride(0), int(relu),
                              _ptr(mask_in), _ptr(mask_out), _ptr(colstats), block_n, _stream())
    if rc and os.environ.get("TFC_DEBUG_KERNEL_ARGS"):
        import sys
        print(f"[convbn_gemm failed] M={M} N={N} K1={K1} K2={K2} b_mn={b_mn} a={tuple(a.shape)}/{a.stride()} "
              f"b={tuple(b.shape)}/{b.stride()} out={out.stride()} res={None if residual is None else residual.stride()} "
              f"relu={relu} mask_in={mask_in is not None} mask_out={mask_out is not None} colstats={colstats is not None} "
              f"ptrs a={_ptr(a):#x} b={_ptr(b):#x} out={_ptr(out):#x} res={_ptr(residual) or 0:#x} "
              f"launches={_launches}", file=sys.stderr, flush=True)
    _lib.check(rc, "convbn_gemm")
    _launches += 1
    return out


class _ConvBnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, gamma, beta, residual, run_mean, run_var, momentum, eps, relu, mean_bias,
                mailbox):
        global _launches
        lib = _k()
        n, cin, h, wd = x.shape
        cout = w.shape[0]
        M = n * h * wd
        dev = x.device
        X = _as_matrix(x)
        W2 = w.reshape(cout, cin)
        # statistics of the input: column sums (one bandwidth pass over the small tensor) and Gram matrix
        xsums = _bn._scratch(lib.cb2_bn_scratch_floats(cin), dev)
        _lib.check(lib.cb2_bn_stats(X.data_ptr(), xsums.data_ptr(), M, cin, _stream()), "bn_stats")
        gram = _bn._scratch(cin * cin, dev).view(cin, cin)
        _dense.gemm(X, X, a_mn=True, b_mn=True, out=gram, accumulate=True, split_k=0)
        f32 = dict(dtype=torch.float32, device=dev)
        vec = torch.empty(4 * cout + cin, **f32)
        scale, shift, mean, rstd = (vec[i * cout:(i + 1) * cout] for i in range(4))
        S = vec[4 * cout:]
        U = torch.empty((cout, cin), **f32)
        u_ready = cin >= int(os.environ.get("TFC_CONVBN_U_GEMM_MIN_K", "256"))
        if u_ready:
            # U = W Gram is O(N K^2): from K = 256 on it goes through the tensor cores, Gram as the
            # sum of two bf16 matrices (hi + lo keeps ~16 mantissa bits; the variance is a difference
            # of two such products), accumulated in fp32
            hl = torch.empty((2, cin, cin), dtype=torch.bfloat16, device=dev)
            _lib.check(lib.cb2_split_hilo_bf16(gram.data_ptr(), hl[0].data_ptr(), hl[1].data_ptr(), cin * cin,
                                               _stream()), "split_hilo")
            _dense.gemm(W2, hl[0], out=U)                      # Gram is symmetric: [n, k] = [k, n]
            _dense.gemm(W2, hl[1], out=U, accumulate=True)
            _launches += 1
        rc = lib.cb2_convbn_fwd_finalize(gram.data_ptr(), cin, xsums.data_ptr(), W2.data_ptr(), cin,
                                         gamma.data_ptr(), beta.data_ptr(), run_mean.data_ptr(),
                                         run_var.data_ptr(), _ptr(mean_bias), scale.data_ptr(),
                                         shift.data_ptr(), mean.data_ptr(), rstd.data_ptr(), U.data_ptr(),
                                         S.data_ptr(), cout, cin, M, float(eps), float(momentum), int(u_ready),
                                         _stream())
        _lib.check(rc, "convbn_fwd_finalize")
        z = torch.empty((n, h, wd, cout), dtype=torch.bfloat16, device=dev)
        mask = torch.empty(M * cout // 8, dtype=torch.uint8, device=dev) if relu else None
        R = None if residual is None else _as_matrix(residual)
        fused_gemm(X, W2, z.view(M, cout), scale=scale, shift=shift, residual=R, relu=relu, mask_out=mask)
        _launches += 2
        if mailbox is not None and mask is not None:
            mailbox["mask"] = mask               # for a consumer that masks this layer's gradient itself
        rp = getattr(residual, "_cb2_producer", None) if residual is not None else None
        # the shortcut is one handle of a forked folded output whose other handle feeds a 1x1
        # convolution that registered itself: that convolution's input-gradient GEMM will add this
        # layer's shortcut gradient and apply the producer's ReLU mask in its epilogue
        ctx.res_prod = rp if (rp is not None and rp.get("dgrad_consumer") and _fuse_g()) else None
        ctx.save_for_backward(x, w, gamma, mask, vec, U)
        ctx.params = (gamma, beta)
        ctx.wparam = w
        ctx.geom = (n, cin, h, wd, cout, M)
        ctx.relu, ctx.has_res = relu, residual is not None
        ctx.mailbox = mailbox
        return z.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dz):
        global _launches
        lib = _k()
        x, w, gamma, mask, vec, U = ctx.saved_tensors
        n, cin, h, wd, cout, M = ctx.geom
        dev = x.device
        scale, shift, mean, rstd = (vec[i * cout:(i + 1) * cout] for i in range(4))
        S = vec[4 * cout:]
        X = _as_matrix(x)
        W2 = w.reshape(cout, cin)
        if not dz.is_contiguous(memory_format=torch.channels_last):
            dz = dz.contiguous(memory_format=torch.channels_last)
        dz2 = ctx.mailbox.pop("dy2", None) if ctx.mailbox is not None else None
        if dz2 is not None and not (dz2.shape == dz.shape and dz2.dtype == dz.dtype and
                                    dz2.is_contiguous(memory_format=torch.channels_last)):
            dz = dz + dz2
            dz2 = None
        # g = masked gradient (+ second consumer's gradient), with its column sums
        pre = ctx.mailbox.pop("premasked", None) if ctx.mailbox is not None else None
        cached = _colsums.pop((dz.data_ptr(), M, cout), None) if (mask is None and dz2 is None) else None
        if pre is not None and dz2 is None:
            cached = pre                         # dz already is the masked, summed gradient (see conv.py)
        gsums = cached if cached is not None else _bn._scratch(lib.cb2_bn_scratch_floats(cout), dev)
        if pre is not None and dz2 is None:
            G = _as_matrix(dz)
        elif mask is None and dz2 is None:
            G = _as_matrix(dz)
            if cached is None:
                _lib.check(lib.cb2_bn_stats(G.data_ptr(), gsums.data_ptr(), M, cout, _stream()), "bn_stats")
        else:
            g4 = torch.empty_like(dz)
            G = _as_matrix(g4)
            rc = lib.cb2_bn_mask_grad(_as_matrix(dz).data_ptr(), None if dz2 is None else _as_matrix(dz2).data_ptr(),
                                      _ptr(mask), G.data_ptr(), gsums.data_ptr(), M, cout, _stream())
            _lib.check(rc, "bn_mask_grad")
        P = _bn._scratch(cout * cin, dev).view(cout, cin)
        _dense.gemm(G, X, a_mn=True, b_mn=True, out=P, accumulate=True, split_k=0)
        gparam, bparam = ctx.params
        wparam = ctx.wparam
        direct_bn = (getattr(gparam, "_cb2_direct_grad", False) and getattr(bparam, "_cb2_direct_grad", False)
                     and gparam.grad is not None and bparam.grad is not None)
        if direct_bn:
            dgamma, dbeta = gparam.grad, bparam.grad
        else:
            dgamma = torch.zeros(cout, dtype=torch.float32, device=dev)
            dbeta = torch.zeros(cout, dtype=torch.float32, device=dev)
        direct_w = getattr(wparam, "_cb2_direct_grad", False) and wparam.grad is not None
        dW = wparam.grad.view(cout, cin) if direct_w else torch.zeros((cout, cin), dtype=torch.bfloat16, device=dev)
        Wst = torch.empty((2, cout, cin), dtype=torch.bfloat16, device=dev)
        Ws, Wt = Wst[0], Wst[1]
        nbias = _bn._scratch(cin, dev)                       # accumulated into: starts at zero
        rc = lib.cb2_convbn_bwd_finalize(P.data_ptr(), cin, gsums.data_ptr(), W2.data_ptr(), cin, U.data_ptr(),
                                         S.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                         dgamma.data_ptr(), dbeta.data_ptr(), dW.data_ptr(), dW.stride(0),
                                         Ws.data_ptr(), Wt.data_ptr(), nbias.data_ptr(), cout, cin, M, _stream())
        _lib.check(rc, "convbn_bwd_finalize")
        _launches += 3
        dx = None
        if ctx.needs_input_grad[0]:
            Qn = _dense.gemm(Wt, W2, a_mn=True, b_mn=True)    # -(W^T diag(t) W), [cin, cin]
            dxm = torch.empty((n, h, wd, cin), dtype=torch.bfloat16, device=dev)
            fused_gemm(G, Ws, dxm.view(M, cin), b_mn=True, a2=X, b2=Qn, shift=nbias)
            dx = dxm.permute(0, 3, 1, 2)
        dres = None
        if ctx.has_res and ctx.res_prod is not None:
            ctx.res_prod["g_next"] = G           # picked up by the 1x1 convolution on the other handle
        elif ctx.has_res:
            dres = G.view(n, h, wd, cout).permute(0, 3, 1, 2)
            # the shortcut may itself be a folded convolution (the first block of a stage): its
            # backward needs the column sums of exactly this tensor - already reduced above
            _colsums.clear()
            _colsums[(G.data_ptr(), M, cout)] = gsums
        dw = dg = db = None
        if direct_w:
            wparam._cb2_grad_written(wparam)
        else:
            dw = dW.view(cout, cin, 1, 1)
        if direct_bn:
            gparam._cb2_grad_written(gparam)
            bparam._cb2_grad_written(bparam)
        else:
            dg, db = dgamma, dbeta
        return dx, dw, dg, db, dres, None, None, None, None, None, None, None


def conv1x1_bn(x, w, gamma, beta, run_mean, run_var, momentum, eps, relu=False, residual=None,
               mean_bias=None):
    """``act(BN(conv1x1(x, w)) + residual)`` in training mode (batch statistics) without storing the
    convolution output.  ``momentum`` is the torch-style update rate of the running statistics."""
    mailbox: dict = {}
    z = _ConvBnFn.apply(x, w, gamma, beta, residual, run_mean, run_var, momentum, eps, relu, mean_bias,
                        mailbox)
    z._cb2_mailbox = mailbox
    return z


def reference(x, w, gamma, beta, eps, relu=False, residual=None):
    """fp32 PyTorch oracle of ``conv1x1_bn`` (training mode)."""
    y = F.conv2d(x.float(), w.float())
    y = F.batch_norm(y, None, None, gamma.float(), beta.float(), True, 0.0, eps)
    if residual is not None:
        y = y + residual.float()
    return torch.relu(y) if relu else y


class DeferredConv:
    """What a Conv2D hands to the BatchNormalization behind it when the pair runs folded: the
    convolution's operands instead of its output."""

    def __init__(self, x, w, layer, pad=0):
        self.x, self.w, self.layer, self.pad = x, w, layer, pad

    def materialize(self):
        return self.layer._plain_conv(self.x, True, self.pad, (1, 1))
