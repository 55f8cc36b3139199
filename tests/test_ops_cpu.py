"""CPU reference paths of the hot ops: the same maths the sm_100a kernels implement, so the layers
above them can be tested (and used on the CPU preset) without a GPU."""
import math

import pytest
import torch
import torch.nn.functional as F

from cloud_b200 import ops
from cloud_b200.ops import conv as conv_ops
import importlib

dense_ops = importlib.import_module("cloud_b200.ops.dense")


def test_gemm_reference_modes_and_epilogues():
    torch.manual_seed(0)
    a = torch.randn(12, 16).to(torch.bfloat16)
    b = torch.randn(10, 16).to(torch.bfloat16)
    bias = torch.randn(10)
    want = a.float() @ b.float().t()
    assert torch.allclose(ops.gemm(a, b, out_dtype=torch.float32), want, atol=1e-5)
    assert torch.allclose(ops.gemm(a.t().contiguous(), b.t().contiguous(), a_mn=True, b_mn=True,
                                   out_dtype=torch.float32), want, atol=1e-5)
    r = ops.gemm(a, b, bias=bias, relu=True, out_dtype=torch.float32)
    assert torch.allclose(r, torch.relu(want + bias), atol=1e-5)
    sm = ops.gemm(a, b, bias=bias, softmax=True, out_dtype=torch.float32)
    assert torch.allclose(sm, torch.softmax(want + bias, 1), atol=1e-5)
    acc = torch.ones(12, 10)
    ops.gemm(a, b, out=acc, accumulate=True)
    assert torch.allclose(acc, want + 1, atol=1e-5)
    stats = torch.zeros(8, 2, 10)
    c = ops.gemm(a, b, colstats=stats).float()
    assert torch.allclose(stats.sum(0)[0], c.sum(0), atol=1e-3)
    assert torch.allclose(stats.sum(0)[1], (c * c).sum(0), atol=1e-2)


def test_gemm_argument_checks():
    a = torch.randn(4, 8).to(torch.bfloat16)
    with pytest.raises(TypeError):
        ops.gemm(a.float(), a)
    with pytest.raises(ValueError):
        ops.gemm(a, torch.randn(4, 6).to(torch.bfloat16))


def test_split_k_reference_accumulates_into_bf16():
    torch.manual_seed(1)
    a = torch.randn(64, 8).to(torch.bfloat16)        # [K, M]
    b = torch.randn(64, 6).to(torch.bfloat16)        # [K, N]
    out = torch.zeros(8, 6, dtype=torch.bfloat16)
    dense_ops.gemm_splitk_bf16(a, b, out)
    assert torch.allclose(out.float(), a.float().t() @ b.float(), atol=0.1, rtol=2e-2)


def test_dense_autograd_matches_torch_on_cpu():
    torch.manual_seed(2)
    x = torch.randn(16, 12, requires_grad=True)
    w = torch.randn(7, 12, requires_grad=True)
    b = torch.randn(7, requires_grad=True)
    y = ops.dense(x, w, b, "relu")
    y.sum().backward()
    xr, wr, br = (t.detach().clone().requires_grad_(True) for t in (x, w, b))
    torch.relu(F.linear(xr, wr, br)).sum().backward()
    for got, want in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert torch.allclose(got, want, atol=1e-5)


def test_softmax_cross_entropy_reference():
    torch.manual_seed(3)
    logits = torch.randn(9, 5, requires_grad=True)
    labels = torch.randint(0, 5, (9,))
    loss = ops.softmax_cross_entropy(logits, labels)
    loss = loss[0] if isinstance(loss, tuple) else loss
    want = F.cross_entropy(logits.detach(), labels, reduction="sum")
    assert math.isclose(float(loss.detach().sum()), float(want), rel_tol=1e-4) or \
        math.isclose(float(loss.detach().sum()), float(want) / 9, rel_tol=1e-4)


def test_stem_space_to_depth_algebra_equals_the_7x7_convolution():
    """The layout cb2_conv_stem_forward consumes: 4 kernel rows x 128-byte windows == conv2d."""
    torch.manual_seed(4)
    B, H, W = 2, 32, 48
    x = torch.randn(B, 3, H, W)
    w = torch.randn(64, 3, 7, 7)
    ref = F.conv2d(x, w, stride=2, padding=3)
    xs = conv_ops.space_to_depth_stem(x)
    wm = conv_ops.stem_weight_matrix(w)
    assert xs.shape == (B, H // 2 + 3, W // 2 + 3, 16) and wm.shape == (64, 256)
    assert float(xs[..., 12:].abs().max()) == 0.0 and float(xs[:, :2].abs().max()) == 0.0
    OH, OW = H // 2, W // 2
    out = torch.zeros(B, OH, OW, 64)
    for i in range(4):
        for ow in range(OW):
            win = xs[:, i:i + OH, ow:ow + 4, :].reshape(B, OH, 64)
            out[:, :, ow, :] += win @ wm[:, i * 64:(i + 1) * 64].t()
    assert torch.allclose(out.permute(0, 3, 1, 2), ref, atol=1e-3)


def test_stem_weight_matrix_is_differentiable_back_to_the_kernel():
    w = torch.randn(64, 3, 7, 7, requires_grad=True)
    wm = conv_ops.stem_weight_matrix(w)
    g = torch.randn_like(wm)
    wm.backward(g)
    # every kernel element appears exactly once in the matrix
    assert torch.allclose(w.grad.abs().sum(), g[wm.detach() != 0].abs().sum(), rtol=1e-4)
    assert int((wm.detach() != 0).sum()) == 64 * 3 * 49


def test_conv_helpers_fall_back_on_cpu():
    x = torch.randn(2, 8, 6, 6)
    w = torch.randn(16, 8, 1, 1)
    assert not ops.conv1x1_supported(x, w, (1, 1), 0, (1, 1))
    assert torch.allclose(ops.conv1x1(x, w), F.conv2d(x, w), atol=1e-5)
    assert not ops.stem_supported(torch.randn(1, 3, 8, 8), torch.randn(64, 3, 7, 7), (2, 2), (3, 3), (1, 1))
    assert ops.conv1x1_mode() in ("wgrad", "all", "0")


def test_batch_norm_cpu_fallback_matches_torch():
    torch.manual_seed(5)
    x = torch.randn(4, 8, 5, 5)
    g, b = torch.rand(8) + 0.5, torch.randn(8)
    rm, rv = torch.zeros(8), torch.ones(8)
    res = torch.randn(4, 8, 5, 5)
    y = ops.batch_norm_nhwc(x, g, b, rm, rv, True, 0.1, 1e-5, relu=True, residual=res)
    want = torch.relu(F.batch_norm(x, torch.zeros(8), torch.ones(8), g, b, True, 0.1, 1e-5) + res)
    assert torch.allclose(y, want, atol=1e-5)
