"""A few launches of each hand-written single-GPU kernel on representative shapes, for
`ncu --set full` captures (one GPU; see /opt/skills/guides/B200_PROFILING.md):

    python bench/ncu_targets.py && ncu --set full --clock-control none --import-source on \
        -k regex:"gemm_tcgen05|bn_|softmax_xent|optim_flat|maxpool|u8_to" -c 24 -o gpurun_out/prof \
        python bench/ncu_targets.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)

# GEMM: large square (compute bound) and a ResNet 1x1-conv shape (output bound)
a = torch.randn(8192, 8192, device=dev).to(torch.bfloat16)
b = torch.randn(8192, 8192, device=dev).to(torch.bfloat16)
for _ in range(2):
    ops.gemm(a, b)
a2 = torch.randn(12544, 512, device=dev).to(torch.bfloat16)
b2 = torch.randn(2048, 512, device=dev).to(torch.bfloat16)
for _ in range(2):
    ops.gemm(a2, b2, bias=torch.zeros(2048, device=dev), relu=True)

# 1x1-convolution shapes: forward with the column-statistics epilogue, split-K weight gradient
a3 = torch.randn(200704, 64, device=dev).to(torch.bfloat16)
b3 = torch.randn(256, 64, device=dev).to(torch.bfloat16)
stats = torch.zeros(8, 2, 256, device=dev)
for _ in range(2):
    ops.gemm(a3, b3, colstats=stats)
dy4 = torch.randn(12544 * 4, 1024, device=dev).to(torch.bfloat16)
x4 = torch.randn(12544 * 4, 256, device=dev).to(torch.bfloat16)
ws4 = torch.zeros(1024, 256, device=dev)
for _ in range(2):
    ops.gemm(dy4, x4, a_mn=True, b_mn=True, out=ws4, accumulate=True, split_k=0)

# batch norm fwd+bwd on the largest ResNet-50 activation (256 x 56 x 56 x 256) and a small one
for (n, c, hw) in ((256, 256, 56), (64, 1024, 14)):
    x = torch.randn(n, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).requires_grad_(True)
    g = torch.ones(c, device=dev, requires_grad=True)
    bt = torch.zeros(c, device=dev, requires_grad=True)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    y = ops.batch_norm_nhwc(x, g, bt, rm, rv, True, 0.1, 1e-5, relu=True)
    y.backward(torch.ones_like(y))

# folded 1x1 convolution + batch-norm (ops/convbn.py): the fused forward GEMM with scale / shift /
# shortcut / ReLU / sign-bit epilogue and the dual-K input-gradient GEMM, ResNet-50 stage-2 shape
import math  # noqa: E402
os.environ.setdefault("TFC_BN_FOLD_MIN_MB", "0")
for (n, cin, hw) in ((256, 64, 56), (256, 128, 28)):
    cout = 4 * cin
    xx = torch.randn(n, cin, hw, hw, device=dev).relu().to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).requires_grad_(True)
    ww = (torch.randn(cout, cin, 1, 1, device=dev) / math.sqrt(cin)).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).requires_grad_(True)
    gg = torch.ones(cout, device=dev, requires_grad=True)
    bb = torch.zeros(cout, device=dev, requires_grad=True)
    rr = torch.randn(n, cout, hw, hw, device=dev).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last).requires_grad_(True)
    rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
    for _ in range(2):
        zz = ops.conv1x1_bn(xx, ww, gg, bb, rm, rv, 0.01, 1e-5, relu=True, residual=rr)
        zz.backward(torch.ones_like(zz))
    del xx, ww, rr, zz

# ResNet stem on the implicit GEMM (space-to-depth image through a 4-D tensor map), batch 256
from cloud_b200.ops import conv as conv_ops  # noqa: E402
img = torch.randint(0, 256, (256, 224, 224, 3), dtype=torch.uint8, device=dev)
wst = (torch.randn(64, 3, 7, 7, device=dev) / 12).to(torch.bfloat16).contiguous(
    memory_format=torch.channels_last).requires_grad_(True)
for _ in range(2):
    ys = ops.stem_conv_s2d(ops.normalize_u8_images_s2d(img), wst)
    ys.backward(torch.ones_like(ys))

# stem max-pool
xp = torch.rand(256, 64, 112, 112, device=dev).to(torch.bfloat16).contiguous(
    memory_format=torch.channels_last).requires_grad_(True)
yp = ops.max_pool_nhwc(xp, 3, 2, 1)
yp.backward(torch.ones_like(yp))

# loss and optimizer
logits = torch.randn(4096, 1000, device=dev).to(torch.bfloat16).requires_grad_(True)
labels = torch.randint(0, 1000, (4096,), device=dev)
ops.softmax_cross_entropy(logits, labels).backward()
n = 25_600_000
master = torch.randn(n, device=dev)
grad = torch.randn(n, device=dev).to(torch.bfloat16)
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
pout = master.to(torch.bfloat16)
hyper = ops.make_hyper(dev)
ops.optim_tick(hyper)
ops.optim_flat(1, master=master, grad=grad, s1=m, s2=v, param_out=pout, hyper=hyper, zero_grad=True)
torch.cuda.synchronize()
print("ncu targets done")
