"""Dense layer ops on the tcgen05 GEMM (``csrc/kernels/gemm_tcgen05.cu``).

``dense(x, w, b, activation)`` is an autograd function whose forward and both backward GEMMs run on
the hand-written kernel: forward with a fused bias(+ReLU | softmax) epilogue, dgrad with a fused
ReLU-mask epilogue available through ``gemm(..., relu_mask=)``, wgrad from MN-major operands (no
transposes), bias gradient fused with the ReLU backward in one memory pass.

Reference workload: ``Dense(512, activation='relu')`` / ``Dense(10, activation='softmax')`` in
README.md:71-75 (ops K1, K2, K3 of SURVEY.md section 2.5).
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F

from . import _lib

GEMM_RELU = 1
GEMM_OUT_F32 = 2
GEMM_ACCUM = 4
GEMM_SOFTMAX = 8

_launches = 0
_DIRECT_W = os.environ.get("TFC_DENSE_DIRECT_W", "1") != "0"
_DIRECT_B = os.environ.get("TFC_DENSE_DIRECT_B", "1") != "0"


def launches() -> int:
    return _launches


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         bias: torch.Tensor | None = None, relu: bool = False, softmax: bool = False,
         relu_mask: torch.Tensor | None = None, out: torch.Tensor | None = None,
         out_dtype: torch.dtype = torch.bfloat16, accumulate: bool = False,
         block_n: int = 0, split_k: int = 1, colstats: torch.Tensor | None = None) -> torch.Tensor:
    """C[M,N] = op(A) @ op(B)^T with fused epilogue, bf16 operands, fp32 accumulation in TMEM.

    ``a`` is stored ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); ``b`` is stored
    ``[N,K]`` (``b_mn=False``) or ``[K,N]`` (``b_mn=True``).  Row pitches must be multiples of 8.
    ``split_k`` (0 = choose) cuts K into slices summed by TMA reduce-adds; it needs
    ``accumulate=True`` into an fp32 ``out``.  ``colstats`` (fp32 ``[8, 2, N]``) receives the
    per-column sum / sum of squares of the bf16 result (batch-norm statistics).
    """
    global _launches
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise TypeError("gemm operands must be bfloat16")
    if a.dim() != 2 or b.dim() != 2 or a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("gemm operands must be 2-D with unit inner stride")
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    if K != Kb:
        raise ValueError(f"reduction dims differ: {K} vs {Kb}")
    if not a.is_cuda:
        c = _gemm_reference(a, b, a_mn, b_mn, bias, relu, softmax, relu_mask, out, out_dtype,
                            accumulate)
        if colstats is not None:
            cf = c.float()
            colstats[0, 0] += cf.sum(0)
            colstats[0, 1] += (cf * cf).sum(0)
        return c
    if out is None:
        if accumulate:
            raise ValueError("accumulate=True needs out=")
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    flags = 0
    if relu:
        flags |= GEMM_RELU
    if softmax:
        flags |= GEMM_SOFTMAX
    if out.dtype == torch.float32:
        flags |= GEMM_OUT_F32
    elif out.dtype != torch.bfloat16:
        raise TypeError("gemm output must be bfloat16 or float32")
    if accumulate:
        flags |= GEMM_ACCUM
    if bias is not None and bias.dtype != torch.float32:
        bias = bias.float()
    lib = _lib.kernels()
    if colstats is not None and (colstats.dtype != torch.float32 or colstats.numel() != 16 * N
                                 or not colstats.is_contiguous()):
        raise ValueError("colstats must be a contiguous fp32 [8, 2, N] tensor")
    rc = lib.cb2_gemm_bf16_ex(_ptr(a), _ptr(b), _ptr(out), _ptr(bias), _ptr(relu_mask), M, N, K,
                              a.stride(0), b.stride(0), out.stride(0),
                              relu_mask.stride(0) if relu_mask is not None else 0,
                              int(a_mn), int(b_mn), flags, block_n, int(split_k), _ptr(colstats),
                              _stream())
    _lib.check(rc, "gemm_bf16")
    _launches += 1
    return out


_splitk_state: dict = {}      # device -> (fp32 workspace, counters); shared by stream-ordered calls


def gemm_splitk_bf16(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn: bool = True,
                     b_mn: bool = True, block_n: int = 0, split_k: int = 0) -> torch.Tensor:
    """``out[M,N] += bf16(op(A) @ op(B)^T)`` with split-K and the final rounding done inside the
    kernel (``cb2_gemm_splitk_bf16``): the weight-gradient GEMM of a 1x1 convolution, written
    straight into a bf16 gradient buffer.  Default operand layout: both stored ``[K, M]`` / ``[K, N]``."""
    global _launches
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    N = b.shape[1] if b_mn else b.shape[0]
    if out.dtype != torch.bfloat16 or out.shape != (M, N) or out.stride(1) != 1:
        raise ValueError("out must be a bf16 [M, N] matrix with unit inner stride")
    if not a.is_cuda:
        af = a.float().t() if a_mn else a.float()
        bf = b.float() if b_mn else b.float().t()
        out += (af @ bf).to(torch.bfloat16)
        return out
    st = _splitk_state.get(a.device)
    if st is None or st[0].numel() < M * N:
        ws = torch.zeros(max(M * N, 1 << 20), dtype=torch.float32, device=a.device)
        counters = torch.zeros(4096, dtype=torch.int32, device=a.device)
        st = _splitk_state[a.device] = (ws, counters)
    ws, counters = st
    rc = _lib.kernels().cb2_gemm_splitk_bf16(
        _ptr(a), _ptr(b), _ptr(ws), _ptr(out), _ptr(counters), counters.numel(), M, N, K,
        a.stride(0), b.stride(0), N, out.stride(0), int(a_mn), int(b_mn), block_n, int(split_k), _stream())
    _lib.check(rc, "gemm_splitk_bf16")
    _launches += 1
    return out


def _gemm_reference(a, b, a_mn, b_mn, bias, relu, softmax, relu_mask, out, out_dtype, accumulate):
    af = a.float().t() if a_mn else a.float()
    bf = b.float() if b_mn else b.float().t()
    c = af @ bf
    if bias is not None:
        c = c + bias.float()
    if relu:
        c = torch.relu(c)
    if softmax:
        c = torch.softmax(c, dim=-1)
    if relu_mask is not None:
        c = torch.where(relu_mask.float() > 0, c, torch.zeros_like(c))
    if out is not None:
        if accumulate:
            out += c.to(out.dtype)
        else:
            out.copy_(c.to(out.dtype))
        return out
    return c.to(out_dtype)


def dense_reference(x, w, b=None, activation=None):
    """fp32 PyTorch oracle for ``dense``."""
    y = x.float() @ w.float().t()
    if b is not None:
        y = y + b.float()
    if activation == "relu":
        y = torch.relu(y)
    elif activation == "softmax":
        y = torch.softmax(y, dim=-1)
    return y


def _pad_cols(t: torch.Tensor, mult: int = 8) -> torch.Tensor:
    r = (-t.shape[-1]) % mult
    if r == 0:
        return t if t.is_contiguous() else t.contiguous()
    return F.pad(t, (0, r))


def _padded_empty(m: int, n: int, device, dtype=torch.bfloat16) -> torch.Tensor:
    """An [m, n] view of an [m, ceil8(n)] buffer: its row pitch is a multiple of 16 bytes, so TMA
    can load and store it - no pad / un-pad copies around a GEMM whose width is not a multiple of 8."""
    n8 = (n + 7) // 8 * 8
    buf = torch.empty((m, n8), dtype=dtype, device=device)
    return buf if n8 == n else buf[:, :n]


def _tma_ok(t: torch.Tensor) -> bool:
    return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


class _DenseFn(torch.autograd.Function):
    """``act(x @ w.T + b)`` (+ inverted dropout after a ReLU).  Backward in three kernels: ReLU mask
    (x dropout scale) + bias gradient in one pass, input gradient, and the weight gradient stored
    straight into the engine's gradient arena when the parameter lives there."""

    @staticmethod
    def forward(ctx, x, w, b, activation, drop):
        x2 = x.reshape(-1, x.shape[-1])
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        wk = w if w.dtype == torch.bfloat16 else w.to(torch.bfloat16)
        if x2.shape[1] % 8 or not _tma_ok(x2):
            x2 = _pad_cols(x2)           # zero-filled up to a multiple of 8 columns (a copy)
        if wk.shape[1] % 8 or not _tma_ok(wk):
            wk = _pad_cols(wk)
        M, N = x2.shape[0], w.shape[0]
        fused_softmax = activation == "softmax" and N <= 256
        y = gemm(x2, wk, bias=b, relu=(activation == "relu"), softmax=fused_softmax,
                 out=_padded_empty(M, N, x2.device))
        if activation == "softmax" and not fused_softmax:
            y = torch.softmax(y.float(), dim=-1).to(torch.bfloat16)  # row wider than one tile
        ctx.drop_scale = 1.0
        if drop is not None and activation == "relu":
            from .elementwise import _dropout_apply, _step
            p, seed, offset = drop
            step = _step["tensor"] if (_step["tensor"] is not None and _step["tensor"].device == y.device) else None
            y = _dropout_apply(y, p, seed, offset, step)
            ctx.drop_scale = 1.0 / (1.0 - p)
        ctx.activation = activation
        ctx.x_shape = x.shape
        ctx.in_features = w.shape[1]
        ctx.has_bias = b is not None
        ctx.w_dtype = w.dtype
        ctx.b_dtype = b.dtype if b is not None else None
        ctx.x_dtype = x.dtype
        ctx.wparam, ctx.bparam = w, b
        ctx.save_for_backward(x2, wk, y if activation in ("relu", "softmax") else None)
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        from .elementwise import relu_bwd_bias_grad
        x2, wk, y = ctx.saved_tensors
        N = wk.shape[0]
        w, b = ctx.wparam, ctx.bparam
        dy2 = dy.reshape(-1, N)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        if not _tma_ok(dy2):
            padded = _padded_empty(dy2.shape[0], N, dy2.device)
            padded.copy_(dy2)
            dy2 = padded
        need_db = ctx.has_bias and ctx.needs_input_grad[2]
        # the bias gradient goes straight into the engine's fp32 gradient arena when it exposes one
        direct_b = need_db and _DIRECT_B and getattr(b, "_cb2_direct_grad", False) and b.grad is not None \
            and b.grad.dtype == torch.float32 and b.grad.is_contiguous()
        db = None
        if ctx.activation == "softmax":
            yf = y.float()
            dzf = yf * (dy2.float() - (dy2.float() * yf).sum(-1, keepdim=True))
            dz = _padded_empty(dy2.shape[0], N, dy2.device)
            dz.copy_(dzf)
            if need_db:
                db = dzf.sum(0)
                if direct_b:
                    b.grad.add_(db)
        else:
            dz, db = relu_bwd_bias_grad(dy2, y if ctx.activation == "relu" else None, want_db=need_db,
                                        scale=ctx.drop_scale, db_out=b.grad if direct_b else None)
        if direct_b:
            b._cb2_grad_written(b)
            db = None
        dx = dw = None
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dZ[M,N] @ W[N,K] : B is stored [reduce=N, K] -> MN-major
            dx = gemm(dz, wk, b_mn=True)
            dx = dx[:, :ctx.in_features].reshape(ctx.x_shape).to(ctx.x_dtype)
        if ctx.needs_input_grad[1]:
            # dW[N,K] = dZ^T[N,M] @ X[M,K] : both stored [reduce=M, *] -> MN-major
            direct_w = (_DIRECT_W and getattr(w, "_cb2_direct_grad", False) and w.grad is not None
                        and w.grad.dtype == torch.bfloat16 and w.grad.dim() == 2 and _tma_ok(w.grad)
                        and x2.shape[1] == ctx.in_features and not getattr(w, "_cb2_grad_stored", False))
            if direct_w:
                # the arena is zero at this point of the step (the bucket kernel zeroes what it
                # reduced), and this is the only writer of this parameter's gradient: a plain store
                gemm(dz, x2, a_mn=True, b_mn=True, out=w.grad)
                w._cb2_grad_written(w)
            else:
                dw = gemm(dz, x2, a_mn=True, b_mn=True,
                          out_dtype=torch.float32 if ctx.w_dtype == torch.float32 else torch.bfloat16)
                dw = dw[:N, :ctx.in_features]
                if dw.dtype != ctx.w_dtype:
                    dw = dw.to(ctx.w_dtype)
        if db is not None and ctx.b_dtype is not None and db.dtype != ctx.b_dtype:
            db = db.to(ctx.b_dtype)
        return dx, dw, db, None, None


def dense(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor | None = None,
          activation: str | None = None, dropout: tuple | None = None) -> torch.Tensor:
    """``activation(x @ w.T + b)`` with ``w`` in torch layout ``[out, in]``.

    CUDA: bf16 tcgen05 GEMM with fused epilogue (output bf16).  CPU: PyTorch in the input dtype.
    ``dropout = (rate, seed, offset)``: inverted dropout behind a ReLU, undone in backward by the
    same pass that applies the ReLU mask and reduces the bias gradient.
    """
    if activation not in (None, "linear", "relu", "softmax"):
        raise ValueError(f"unsupported fused activation: {activation}")
    if activation == "linear":
        activation = None
    if x.is_cuda:
        return _DenseFn.apply(x, w, b, activation, dropout)
    y = F.linear(x, w.to(x.dtype), None if b is None else b.to(x.dtype))
    if activation == "relu":
        y = torch.relu(y)
    elif activation == "softmax":
        y = torch.softmax(y, dim=-1)
    if dropout is not None:
        y = F.dropout(y, dropout[0], training=True)
    return y
