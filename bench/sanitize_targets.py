"""One small launch of every hand-written single-GPU kernel, for compute-sanitizer
(SURVEY.md section 5.2):

    python bench/sanitize_targets.py && \
        compute-sanitizer --tool memcheck --error-exitcode 1 python bench/sanitize_targets.py

Shapes are tiny on purpose (the sanitizer slows kernels down 10-100x) but ragged, so every
bounds-handling path (TMA clipping, row/column tails, split-K tails, partial pixel boxes) runs.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TFC_CONV1X1_GEMM", "all")
from cloud_b200 import ops  # noqa: E402
import importlib  # noqa: E402

dense_ops = importlib.import_module("cloud_b200.ops.dense")
dev = torch.device("cuda:0")
torch.manual_seed(0)


def bf(*shape):
    return torch.randn(*shape, device=dev).to(torch.bfloat16)


def nhwc(*shape):
    return bf(*shape).contiguous(memory_format=torch.channels_last)


# GEMM: all operand-major modes, ragged tiles, epilogues, split-K, column statistics, finalize
for (a_mn, b_mn) in ((False, False), (False, True), (True, True), (True, False)):
    a = bf(264, 200) if a_mn else bf(200, 264)
    b = bf(264, 136) if b_mn else bf(136, 264)
    ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, bias=torch.randn(136, device=dev), relu=True)
ops.gemm(bf(96, 72), bf(10, 72), bias=torch.zeros(10, device=dev), softmax=True)
acc = torch.zeros(200, 136, device=dev)
ops.gemm(bf(1000, 200), bf(1000, 136), a_mn=True, b_mn=True, out=acc, accumulate=True, split_k=3)
ops.gemm(bf(520, 72), bf(136, 72), colstats=torch.zeros(8, 2, 136, device=dev))
dense_ops.gemm_splitk_bf16(bf(1000, 200), bf(1000, 136), torch.zeros(200, 136, device=dev, dtype=torch.bfloat16))

# dense layer autograd (fwd, dgrad with mask, wgrad, bias grad)
x = bf(40, 72).requires_grad_(True)
w = bf(24, 72).requires_grad_(True)
bb = torch.zeros(24, device=dev, requires_grad=True)
ops.dense(x, w, bb, "relu").float().sum().backward()

# convolutions on the GEMM: 1x1 (all three GEMMs + statistics), stem, k x k wgrad, strided 1x1 wgrad
xc = nhwc(2, 64, 10, 6).requires_grad_(True)
wc = nhwc(128, 64, 1, 1).requires_grad_(True)
ops.conv1x1(xc, wc, None, bn_stats=True).float().sum().backward()
img = torch.randint(0, 256, (2, 20, 28, 3), dtype=torch.uint8, device=dev)
ws = (nhwc(64, 3, 7, 7) / 12).requires_grad_(True)
ops.stem_conv_s2d(ops.normalize_u8_images_s2d(img), ws).float().sum().backward()
ops.stem_conv(nhwc(1, 3, 12, 16), ws.detach().requires_grad_(True)).float().sum().backward()
xk = nhwc(2, 64, 9, 7).requires_grad_(True)
wk = nhwc(64, 64, 3, 3).requires_grad_(True)
ops.conv_kxk(xk, wk).float().sum().backward()
from cloud_b200.ops import _lib  # noqa: E402
xs2, dys2 = nhwc(2, 64, 12, 8), nhwc(2, 128, 6, 4)
wsb = torch.zeros(64, 128, device=dev)
_lib.check(_lib.kernels().cb2_conv_wgrad(xs2.data_ptr(), dys2.data_ptr(), wsb.data_ptr(), 2, 6, 4, 64, 128, 1, 0, 2,
                                         torch.cuda.current_stream().cuda_stream), "conv_wgrad")

# batch norm (+shortcut, +ReLU, forked gradient), pooling, loss, dropout, normalisation, optimizer
for c, relu, with_res in ((64, True, True), (8, False, False), (192, True, False)):
    xb = nhwc(3, c, 5, 7).requires_grad_(True)
    res = nhwc(3, c, 5, 7).requires_grad_(True) if with_res else None
    g = torch.ones(c, device=dev, requires_grad=True)
    be = torch.zeros(c, device=dev, requires_grad=True)
    y = ops.batch_norm_nhwc(xb, g, be, torch.zeros(c, device=dev), torch.ones(c, device=dev), True, 0.1, 1e-5,
                            relu=relu, residual=res)
    y1, y2 = ops.bn_fork(y)
    (y1.float() * 2 + y2.float()).sum().backward()
xp = nhwc(2, 16, 12, 10).requires_grad_(True)
ops.max_pool_nhwc(xp, 3, 2, 1).float().sum().backward()
xq = nhwc(2, 8, 9, 9).requires_grad_(True)
ops.max_pool_nhwc(xq, 2, 2, 0).float().sum().backward()
logits = bf(33, 17).requires_grad_(True)
loss = ops.softmax_cross_entropy(logits, torch.randint(0, 17, (33,), device=dev))
(loss[0] if isinstance(loss, tuple) else loss).sum().backward()
ops.dropout(bf(1000), 0.3, training=True) if "training" in ops.dropout.__code__.co_varnames else ops.dropout(bf(1000), 0.3)
ops.normalize_u8_images(torch.randint(0, 256, (2, 6, 6, 3), dtype=torch.uint8, device=dev))
n = 4104
master = torch.randn(n, device=dev)
hyper = ops.make_hyper(dev)
ops.optim_tick(hyper)
ops.optim_flat(1, master=master, grad=bf(n), s1=torch.zeros(n, device=dev), s2=torch.zeros(n, device=dev),
               param_out=master.to(torch.bfloat16), hyper=hyper, zero_grad=True)
torch.cuda.synchronize()
print("sanitize targets done")
