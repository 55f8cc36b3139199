"""GPU numerics tests: every hand-written kernel against a plain PyTorch fp32 oracle of the same op.

Run on a B200 with ``python -m pytest tests -m gpu``.  (The reference has no GPU tests at all -
SURVEY.md section 4 - so this tier is new.)
"""
import math

import pytest
import torch

from cloud_b200 import ops
import importlib

dense_ops = importlib.import_module('cloud_b200.ops.dense')
from cloud_b200.ops.optim import _reference_update

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _rand_bf16(*shape, scale=1.0):
    return (torch.randn(*shape, device=_dev()) * scale).to(torch.bfloat16)


def _check_close(got, want, rtol, atol, what=""):
    got = got.float()
    want = want.float()
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{what}: {bad}/{err.numel()} mismatches, max err {err.max().item():.4g}"


GEMM_SHAPES = [
    # M, N, K
    (128, 128, 64),
    (128, 512, 784),     # README MLP layer 1 (README.md:72)
    (128, 10, 512),      # README MLP layer 2, N not a multiple of 8
    (256, 256, 4096),
    (1000, 136, 520),    # ragged tiles in every dimension
    (4096, 2048, 1024),
    (64, 120, 2048),     # ResNet head (call_run_within_script_with_keras_fit.py:83-85)
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("block_n", [0, 64, 128, 256])
def test_gemm_kmajor_forward(M, N, K, block_n):
    torch.manual_seed(0)
    a = _rand_bf16(M, K)
    b = _rand_bf16(N, K)
    bias = torch.randn(N, device=_dev())
    got = ops.gemm(a, b, bias=bias, relu=True, block_n=block_n)
    want = torch.relu(a.float() @ b.float().t() + bias)
    _check_close(got, want, rtol=1e-2, atol=1e-2 * math.sqrt(K), what=f"fwd {M}x{N}x{K}")


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 512, 128), (1000, 136, 520), (512, 784, 512)])
def test_gemm_b_mn_major_dgrad(M, N, K):
    # C[M,N] = A[M,K] @ Bst[K,N]   (B stored [K,N] -> MN-major)
    torch.manual_seed(1)
    a = _rand_bf16(M, K)
    bst = _rand_bf16(K, N)
    got = ops.gemm(a, bst, b_mn=True)
    want = a.float() @ bst.float()
    _check_close(got, want, rtol=1e-2, atol=1e-2 * math.sqrt(K), what="dgrad")


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (512, 784, 128), (136, 520, 1000), (64, 2048, 4096)])
def test_gemm_both_mn_major_wgrad(M, N, K):
    # C[M,N] = Ast[K,M]^T @ Bst[K,N]
    torch.manual_seed(2)
    ast = _rand_bf16(K, M)
    bst = _rand_bf16(K, N)
    got = ops.gemm(ast, bst, a_mn=True, b_mn=True, out_dtype=torch.float32)
    want = ast.float().t() @ bst.float()
    _check_close(got, want, rtol=1e-2, atol=1e-2 * math.sqrt(K), what="wgrad")


def test_gemm_accumulate_and_mask():
    torch.manual_seed(3)
    M, N, K = 256, 192, 320
    a, b = _rand_bf16(M, K), _rand_bf16(N, K)
    out = torch.randn(M, N, device=_dev())
    base = out.clone()
    ops.gemm(a, b, out=out, accumulate=True)
    _check_close(out, base + a.float() @ b.float().t(), 1e-2, 0.3, "accumulate")
    mask = _rand_bf16(M, N)
    got = ops.gemm(a, b, relu_mask=mask)
    want = torch.where(mask.float() > 0, a.float() @ b.float().t(), torch.zeros(M, N, device=_dev()))
    _check_close(got, want, 1e-2, 0.3, "mask")


def test_gemm_softmax_epilogue():
    torch.manual_seed(4)
    M, N, K = 300, 10, 512
    a, b = _rand_bf16(M, K, scale=0.1), _rand_bf16(N, K, scale=0.1)
    bias = torch.randn(N, device=_dev())
    got = ops.gemm(a, b, bias=bias, softmax=True, out_dtype=torch.float32)
    want = torch.softmax(a.float() @ b.float().t() + bias, dim=-1)
    _check_close(got, want, 1e-2, 1e-3, "softmax")
    assert torch.allclose(got.sum(-1), torch.ones(M, device=_dev()), atol=1e-3)


@pytest.mark.parametrize("activation", [None, "relu", "softmax"])
@pytest.mark.parametrize("M,K,N", [(128, 784, 512), (96, 512, 10), (64, 2048, 120)])
def test_dense_autograd(activation, M, K, N):
    torch.manual_seed(5)
    x = (torch.randn(M, K, device=_dev()) * 0.5).to(torch.bfloat16).requires_grad_(True)
    w = (torch.randn(N, K, device=_dev()) / math.sqrt(K)).to(torch.bfloat16).requires_grad_(True)
    b = (torch.randn(N, device=_dev()) * 0.1).requires_grad_(True)
    y = ops.dense(x, w, b, activation)
    gy = torch.randn_like(y.float()).to(torch.bfloat16)
    y.backward(gy)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True)
    yr = ops.dense_reference(xr, wr, br, activation)
    yr.backward(gy.float())
    _check_close(y, yr, 2e-2, 2e-2, "dense fwd")
    _check_close(x.grad, xr.grad, 3e-2, 3e-2 * math.sqrt(N / 64 + 1), "dense dX")
    _check_close(w.grad, wr.grad, 3e-2, 3e-2 * math.sqrt(M / 64 + 1), "dense dW")
    _check_close(b.grad, br.grad, 3e-2, 3e-2 * math.sqrt(M / 64 + 1), "dense db")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,C", [(128, 10), (64, 120), (257, 1000), (8, 33)])
def test_softmax_xent(dtype, B, C):
    torch.manual_seed(6)
    logits = (torch.randn(B, C, device=_dev()) * 3).to(dtype).requires_grad_(True)
    labels = torch.randint(0, C, (B,), device=_dev())
    stats = torch.zeros(4, device=_dev())
    loss = ops.softmax_cross_entropy(logits, labels, stats=stats)
    loss.backward()
    ls, d, correct = ops.softmax_cross_entropy_reference(logits.detach(), labels)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert abs(loss.item() - ls.item() / B) < 1e-3 * max(1.0, abs(ls.item() / B))
    _check_close(logits.grad, d, tol, tol / B, "dlogits")
    assert stats[1].item() == correct.item()
    assert abs(stats[0].item() - ls.item()) < 1e-3 * max(1.0, abs(ls.item()))


def test_relu_bwd_bias_grad():
    torch.manual_seed(7)
    for M, N in [(128, 512), (1000, 136), (64, 10), (5000, 64)]:
        dy, y = _rand_bf16(M, N), _rand_bf16(M, N)
        dz, db = ops.relu_bwd_bias_grad(dy, y)
        want = torch.where(y.float() > 0, dy.float(), torch.zeros_like(dy.float()))
        assert torch.equal(dz.float(), want)
        _check_close(db, want.sum(0), 1e-3, 1e-2, "db")
        _, db2 = ops.relu_bwd_bias_grad(dy, None)
        _check_close(db2, dy.float().sum(0), 1e-3, 1e-2, "db no mask")


def test_dropout_mask_replay():
    torch.manual_seed(8)
    x = torch.randn(1 << 20, device=_dev(), requires_grad=True)
    y = ops.dropout(x, 0.2, seed=123, offset=77)
    keep = (y != 0).float().mean().item()
    assert abs(keep - 0.8) < 5e-3
    nz = y != 0
    assert torch.allclose(y[nz], x[nz] / 0.8)
    y.sum().backward()
    assert torch.equal(x.grad != 0, nz)           # backward regenerates the same mask
    y2 = ops.dropout(x.detach().to(torch.bfloat16), 0.5, seed=123, offset=0)
    assert abs((y2 != 0).float().mean().item() - 0.5) < 5e-3


def test_normalize_u8_images():
    img = torch.randint(0, 256, (8, 32, 32, 3), dtype=torch.uint8, device=_dev())
    got = ops.normalize_u8_images(img)
    want = ops.normalize_u8_images(img.cpu()).to(_dev())
    _check_close(got, want, 1e-2, 1e-2, "normalize")


@pytest.mark.parametrize("opt", ["adam_keras", "adam_torch", "adamw", "sgd", "rmsprop"])
@pytest.mark.parametrize("gdt,pdt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16)])
def test_optim_flat_vs_reference(opt, gdt, pdt):
    torch.manual_seed(9)
    n = 100_003
    code = ops.OPT_CODES[opt]
    hyper = ops.make_hyper(_dev(), lr=1e-2, weight_decay=0.01, grad_scale=0.5, momentum=0.9,
                           nesterov=True, rms_momentum=0.5)
    master = torch.randn(n, device=_dev())
    s1 = torch.zeros(n, device=_dev())
    s2 = torch.zeros(n, device=_dev())
    pout = master.to(pdt).clone()
    r_master, r_s1, r_s2 = master.cpu().clone(), s1.cpu().clone(), s2.cpu().clone()
    for _ in range(3):
        g = torch.randn(n, device=_dev()).to(gdt)
        ops.optim_tick(hyper)
        ops.optim_flat(code, master=master, grad=g.clone(), s1=s1, s2=s2, param_out=pout,
                       hyper=hyper, zero_grad=False)
        r_master = _reference_update(code, r_master, g.cpu(), r_s1, r_s2, hyper.cpu())
    _check_close(master, r_master.to(_dev()), 1e-4, 1e-5, f"{opt} master")
    _check_close(pout, master, 1e-2 if pdt == torch.bfloat16 else 0, 1e-6, f"{opt} param_out")
    assert hyper[5].item() == 3.0


def test_adam_torch_matches_torch_optim():
    torch.manual_seed(10)
    n = 4096
    p = torch.randn(n, device=_dev())
    ref = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1e-3, eps=1e-8)
    hyper = ops.make_hyper(_dev(), lr=1e-3, eps=1e-8)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for _ in range(5):
        g = torch.randn(n, device=_dev())
        ref.grad = g.clone()
        opt.step()
        ops.optim_tick(hyper)
        ops.optim_flat(ops.OPT_CODES["adam_torch"], master=None, grad=g, s1=m, s2=v, param_out=p,
                       hyper=hyper, zero_grad=True)
        assert g.abs().sum().item() == 0.0
    _check_close(p, ref.detach(), 1e-5, 1e-6, "adam vs torch")


def test_kernel_launch_counters_move():
    before = dense_ops.launches()
    ops.gemm(_rand_bf16(128, 64), _rand_bf16(64, 64))
    assert dense_ops.launches() == before + 1


@pytest.mark.parametrize("C,HW,N", [(64, 56, 8), (256, 14, 16), (2048, 7, 4), (512, 28, 3)])
@pytest.mark.parametrize("relu,with_res", [(False, False), (True, False), (True, True), (False, True)])
def test_fused_batch_norm(C, HW, N, relu, with_res):
    torch.manual_seed(11)
    dev = _dev()
    x = (torch.randn(N, C, HW, HW, device=dev) * 2 + 0.5).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    res = None
    if with_res:
        res = torch.randn(N, C, HW, HW, device=dev).to(torch.bfloat16) \
            .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    gamma = (torch.rand(C, device=dev) + 0.5).requires_grad_(True)
    beta = (torch.randn(C, device=dev) * 0.1).requires_grad_(True)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    y = ops.batch_norm_nhwc(x, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=relu, residual=res)
    gy = torch.randn_like(y)
    y.backward(gy)
    # fp32 oracle
    xr = x.detach().float().requires_grad_(True)
    rr = res.detach().float().requires_grad_(True) if with_res else None
    gr, br = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    rm2, rv2 = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    yr = torch.nn.functional.batch_norm(xr, rm2, rv2, gr, br, True, 0.1, 1e-5)
    if with_res:
        yr = yr + rr
    if relu:
        yr = torch.relu(yr)
    yr.backward(gy.float())
    _check_close(y, yr, 2e-2, 2e-2, "bn fwd")
    _check_close(rm, rm2, 1e-3, 1e-3, "running mean")
    _check_close(rv, rv2, 1e-2, 1e-2, "running var")
    # compare gradients where the bf16 ReLU mask agrees with the fp32 one
    _check_close(beta.grad, br.grad, 3e-2, 0.05 * (N * HW * HW) ** 0.5, "dbeta")
    _check_close(gamma.grad, gr.grad, 3e-2, 0.08 * (N * HW * HW) ** 0.5, "dgamma")
    err = (x.grad.float() - xr.grad).abs()
    assert (err > 0.05 + 0.05 * xr.grad.abs()).float().mean().item() < 2e-3, "bn dX"
    if with_res:
        err = (res.grad.float() - rr.grad).abs()
        assert (err > 0.02 + 0.02 * rr.grad.abs()).float().mean().item() < 2e-3, "bn dRes"
    # inference path
    with torch.no_grad():
        yi = ops.batch_norm_nhwc(x.detach(), gamma.detach(), beta.detach(), rm, rv, False, 0.1, 1e-5,
                                 relu=relu, residual=None if res is None else res.detach())
        yo = torch.nn.functional.batch_norm(x.detach().float(), rm, rv, gamma, beta, False, 0.1, 1e-5)
        if with_res:
            yo = yo + res.detach().float()
        if relu:
            yo = torch.relu(yo)
    _check_close(yi, yo, 2e-2, 2e-2, "bn inference")


@pytest.mark.parametrize("M,N,K,split", [(256, 64, 8192, 0), (64, 256, 200704 // 8, 0), (512, 2048, 3136, 4),
                                         (136, 520, 1000, 3), (128, 128, 256, 8)])
def test_gemm_split_k_wgrad(M, N, K, split):
    """dW = dY^T @ X with the reduction cut into K slices summed by TMA reduce-adds."""
    torch.manual_seed(5)
    a = _rand_bf16(K, M, scale=0.5)          # stored [K, M]  (A MN-major)
    b = _rand_bf16(K, N, scale=0.5)          # stored [K, N]  (B MN-major)
    out = torch.zeros(M, N, device=_dev())
    ops.gemm(a, b, a_mn=True, b_mn=True, out=out, accumulate=True, split_k=split)
    want = a.float().t() @ b.float()
    _check_close(out, want, 1e-3, 1e-3 * math.sqrt(K), "split-K wgrad")
    # a second call accumulates on top
    ops.gemm(a, b, a_mn=True, b_mn=True, out=out, accumulate=True, split_k=split)
    _check_close(out, 2 * want, 1e-3, 2e-3 * math.sqrt(K), "split-K wgrad, second accumulate")


@pytest.mark.parametrize("M,N,K", [(200704 // 16, 256, 64), (3136, 2048, 512), (1000, 520, 136), (128, 64, 64),
                                   (50176 // 4, 128, 512)])
def test_gemm_column_statistics(M, N, K):
    """The colstats epilogue returns sum / sum of squares of the bf16 output per column."""
    torch.manual_seed(6)
    a, b = _rand_bf16(M, K), _rand_bf16(N, K, scale=0.2)
    stats = torch.zeros(8, 2, N, device=_dev())
    c = ops.gemm(a, b, colstats=stats)
    want_c = a.float() @ b.float().t()
    _check_close(c, want_c, 2e-2, 2e-2, "gemm with colstats")
    cf = c.float()
    tot = stats.sum(0)
    _check_close(tot[0], cf.sum(0), 1e-3, 1e-3 * math.sqrt(M) * cf.abs().max().item(), "column sums")
    _check_close(tot[1], (cf * cf).sum(0), 1e-3, 1e-2, "column sums of squares")


@pytest.mark.parametrize("gemm_mode", ["all", "wgrad"])
@pytest.mark.parametrize("N,H,Cin,Cout,bias", [(8, 56, 64, 256, False), (4, 14, 1024, 256, False),
                                               (2, 7, 512, 2048, True), (3, 28, 128, 512, False)])
def test_conv1x1_on_gemm_matches_conv2d(N, H, Cin, Cout, bias, gemm_mode, monkeypatch):
    monkeypatch.setenv("TFC_CONV1X1_GEMM", gemm_mode)
    torch.manual_seed(7)
    dev = _dev()
    x = torch.randn(N, Cin, H, H, device=dev).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    w = (torch.randn(Cout, Cin, 1, 1, device=dev) / math.sqrt(Cin)).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    b = (torch.randn(Cout, device=dev) * 0.1).requires_grad_(True) if bias else None
    assert ops.conv1x1_supported(x, w, (1, 1), 0, (1, 1))
    y = ops.conv1x1(x, w, b, bn_stats=not bias)
    assert y.shape == (N, Cout, H, H) and y.is_contiguous(memory_format=torch.channels_last)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True) if bias else None
    yr = torch.nn.functional.conv2d(xr, wr, br)
    yr.backward(gy.float())
    _check_close(y, yr, 2e-2, 2e-2, "conv1x1 fwd")
    _check_close(x.grad, xr.grad, 2e-2, 2e-2 * math.sqrt(Cout / 64), "conv1x1 dX")
    _check_close(w.grad, wr.grad, 2e-2, 2e-2 * math.sqrt(N * H * H), "conv1x1 dW")
    if bias:
        _check_close(b.grad, br.grad, 1e-2, 1e-2 * math.sqrt(N * H * H), "conv1x1 db")
    elif gemm_mode == "all":
        sums = y._cb2_bn_sums.view(8, 2, Cout).sum(0)
        yf = y.detach().float()
        _check_close(sums[0], yf.sum((0, 2, 3)), 1e-3, 1e-2 * math.sqrt(N * H * H), "conv1x1 channel sums")
        _check_close(sums[1], (yf * yf).sum((0, 2, 3)), 1e-3, 1e-1, "conv1x1 channel sums of squares")


def test_batch_norm_with_statistics_from_the_conv_epilogue(monkeypatch):
    """conv1x1(bn_stats) -> fused BN must equal conv -> BN with its own statistics pass."""
    monkeypatch.setenv("TFC_CONV1X1_GEMM", "all")
    torch.manual_seed(8)
    dev = _dev()
    N, H, Cin, Cout = 8, 28, 128, 256
    x = torch.randn(N, Cin, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(Cout, Cin, 1, 1, device=dev) / math.sqrt(Cin)).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last)
    gamma = (torch.rand(Cout, device=dev) + 0.5).requires_grad_(True)
    beta = (torch.randn(Cout, device=dev) * 0.1).requires_grad_(True)
    outs = []
    for fused in (True, False):
        rm, rv = torch.zeros(Cout, device=dev), torch.ones(Cout, device=dev)
        y = ops.conv1x1(x, w, None, bn_stats=fused)
        sums = getattr(y, "_cb2_bn_sums", None)
        assert (sums is not None) == fused
        z = ops.batch_norm_nhwc(y, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=True, sums=sums)
        outs.append((z.detach().float(), rm, rv))
    _check_close(outs[0][0], outs[1][0], 1e-2, 1e-2, "BN output")
    _check_close(outs[0][1], outs[1][1], 1e-4, 1e-4, "running mean")
    _check_close(outs[0][2], outs[1][2], 1e-3, 1e-4, "running var")


@pytest.mark.parametrize("M,N,K", [(256, 64, 8192), (64, 256, 25088), (512, 2048, 3136), (136, 520, 1000),
                                   (1024, 256, 12544), (64, 64, 200704)])
def test_gemm_split_k_finalized_in_kernel(M, N, K):
    """cb2_gemm_splitk_bf16: bf16 out += A^T B, workspace and counters left clean for the next call."""
    torch.manual_seed(9)
    a = _rand_bf16(K, M, scale=0.5)
    b = _rand_bf16(K, N, scale=0.5)
    want = a.float().t() @ b.float()
    out = torch.zeros(M, N, device=_dev(), dtype=torch.bfloat16)
    for rep in (1, 2, 3):           # repeated calls reuse the shared workspace: it must come back zeroed
        dense_ops.gemm_splitk_bf16(a, b, out)
        if rep == 1:
            _check_close(out, want, 1e-2, 1e-2 * math.sqrt(K), "split-K finalize")
    # three accumulations, each rounded to bf16 on the way in
    _check_close(out, 3 * want, 2e-2, 3e-2 * math.sqrt(K), "split-K finalize, accumulated")
    ws, counters = dense_ops._splitk_state[_dev()]
    assert int(counters.abs().sum().item()) == 0
    assert float(ws.abs().max().item()) == 0.0


@pytest.mark.parametrize("B,H,W", [(4, 224, 224), (3, 64, 96), (2, 32, 256)])
def test_stem_conv_implicit_gemm_matches_conv2d(B, H, W):
    """7x7/2 stem on the tcgen05 implicit GEMM (space-to-depth input, 4-D TMA map) vs conv2d."""
    torch.manual_seed(10)
    dev = _dev()
    x = torch.randn(B, 3, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, 3, 7, 7, device=dev) / math.sqrt(147)).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    assert ops.stem_supported(x, w, (2, 2), (3, 3), (1, 1))
    y = ops.stem_conv(x, w)
    assert y.shape == (B, 64, H // 2, W // 2) and y.is_contiguous(memory_format=torch.channels_last)
    gy = torch.randn_like(y)
    y.backward(gy)
    wr = w.detach().float().requires_grad_(True)
    yr = torch.nn.functional.conv2d(x.float(), wr, None, stride=2, padding=3)
    yr.backward(gy.float())
    _check_close(y, yr, 2e-2, 2e-2, "stem fwd")
    _check_close(w.grad, wr.grad, 2e-2, 2e-2 * math.sqrt(B * H * W / 4), "stem dW")


def test_u8_images_straight_into_the_stem_layout():
    """normalize_u8_images_s2d == space_to_depth(normalize_u8_images) and the Keras stem uses it."""
    from cloud_b200.ops import conv as conv_ops
    torch.manual_seed(11)
    dev = _dev()
    img = torch.randint(0, 256, (3, 64, 96, 3), dtype=torch.uint8, device=dev)
    got = ops.normalize_u8_images_s2d(img)
    want = conv_ops.space_to_depth_stem(ops.normalize_u8_images(img).permute(0, 3, 1, 2))
    assert got.shape == want.shape == (3, 35, 51, 16)
    assert torch.equal(got, want)

    import cloud_b200 as tfc
    from cloud_b200.keras import layers
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
    inp = keras.Input(shape=(64, 96, 3))
    x = layers.ImageNormalization(name="norm")(inp)
    x = layers.Conv2D(64, 7, strides=2, padding=3, use_bias=False, name="stem")(x)
    out = layers.GlobalAveragePooling2D()(x)
    m = keras.Model(inp, out)
    assert m.get_layer("norm").defer_to_stem
    m.compile(optimizer="sgd", loss="mse")
    y = m.predict(img.cpu().numpy(), batch_size=3)
    w = torch.tensor(m.get_layer("stem").get_weights()[0], device=dev).permute(3, 2, 0, 1)   # HWIO -> OIHW
    ref = torch.nn.functional.conv2d(ops.normalize_u8_images(img).permute(0, 3, 1, 2).float(), w.float(),
                                     stride=2, padding=3).mean((2, 3))
    _check_close(torch.tensor(y, device=dev), ref, 3e-2, 3e-2, "keras stem on raw bytes")


@pytest.mark.parametrize("B,H,W,Cin,Cout,k", [(4, 56, 56, 64, 64, 3), (3, 28, 28, 128, 128, 3),
                                              (5, 14, 14, 256, 256, 3), (6, 7, 7, 512, 512, 3),
                                              (2, 20, 12, 64, 128, 5)])
def test_conv_kxk_weight_gradient_on_shifted_tma_boxes(B, H, W, Cin, Cout, k):
    torch.manual_seed(12)
    dev = _dev()
    x = torch.randn(B, Cin, H, W, device=dev).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    w = (torch.randn(Cout, Cin, k, k, device=dev) / math.sqrt(Cin * k * k)).to(torch.bfloat16) \
        .contiguous(memory_format=torch.channels_last).requires_grad_(True)
    assert ops.kxk_supported(x, w, (1, 1), k // 2, (1, 1))
    y = ops.conv_kxk(x, w)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr, padding=k // 2)
    yr.backward(gy.float())
    _check_close(y, yr, 2e-2, 2e-2, "conv fwd")
    _check_close(x.grad, xr.grad, 3e-2, 3e-2, "conv dX")
    _check_close(w.grad, wr.grad, 2e-2, 2e-2 * math.sqrt(B * H * W), "conv dW")


def test_forked_batch_norm_output_merges_gradients_in_its_backward():
    """y = BN(x) feeds two consumers: with ops.bn_fork the two gradients are summed inside the BN
    backward kernels; the result must equal plain autograd accumulation."""
    torch.manual_seed(13)
    dev = _dev()
    C = 128
    x0 = (torch.randn(6, C, 14, 14, device=dev) * 2).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w1 = torch.randn(C, C, 1, 1, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) / 12
    gamma0 = torch.rand(C, device=dev) + 0.5
    grad_out = torch.randn(6, C, 14, 14, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res_out = {}
    for forked in (True, False):
        x = x0.clone().requires_grad_(True)
        gamma = gamma0.clone().requires_grad_(True)
        beta = torch.zeros(C, device=dev, requires_grad=True)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        a = ops.batch_norm_nhwc(x, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=True)
        a1, a2 = ops.bn_fork(a) if forked else (a, a)
        assert (a1 is not a) == forked
        out = torch.nn.functional.conv2d(a1, w1) * 0.5 + a2 * 2.0
        out.backward(grad_out)
        res_out[forked] = (x.grad.float(), gamma.grad.clone(), beta.grad.clone())
    for got, want, name in zip(res_out[True], res_out[False], ("dx", "dgamma", "dbeta")):
        scale = want.abs().max().item()
        assert (got - want).abs().max().item() <= 2e-2 * scale + 1e-3, name
