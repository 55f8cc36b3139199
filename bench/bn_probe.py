"""Fused batch-norm kernels on ResNet-50 activation shapes: time and effective bandwidth per
kernel (CUDA events; tensors are rotated through a pool larger than L2 so every call is cold)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.ops import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.kernels()
import cloud_b200.ops.bn as bnmod  # noqa: E402,F401  (declares argtypes)
bnmod._k()
st = torch.cuda.current_stream().cuda_stream


def run(batch, hw, c, relu=1, with_res=False):
    m = batch * hw * hw
    nbuf = max(2, int(600e6 // (m * c * 2)) + 1)
    xs = [torch.randn(m, c, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    dys = [torch.randn(m, c, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    y = torch.empty(m, c, device=dev, dtype=torch.bfloat16)
    dx = torch.empty_like(y)
    g, b = torch.ones(c, device=dev), torch.zeros(c, device=dev)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    save = torch.zeros(2 * c, device=dev)
    save[c:] = 1.0
    sums = torch.zeros(16 * c, device=dev)
    dbg = torch.zeros(2 * c, device=dev)
    res = {}

    def timeit(name, fn, bytes_):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / iters * 1e3, bytes_

    sz = m * c * 2
    # forward = stats + apply (two launches); mode 2 = apply only
    timeit("fwd(stats+apply)", lambda i: lib.cb2_bn_forward(
        xs[i % nbuf].data_ptr(), None, y.data_ptr(), sums.data_ptr(), g.data_ptr(), b.data_ptr(),
        rm.data_ptr(), rv.data_ptr(), save.data_ptr(), save[c:].data_ptr(), None, m, c, 1e-5, 0.1, relu, 1, st),
        3 * sz)
    timeit("apply", lambda i: lib.cb2_bn_forward(
        xs[i % nbuf].data_ptr(), None, y.data_ptr(), sums.data_ptr(), g.data_ptr(), b.data_ptr(),
        rm.data_ptr(), rv.data_ptr(), save.data_ptr(), save[c:].data_ptr(), None, m, c, 1e-5, 0.1, relu, 2, st),
        2 * sz)
    timeit("bwd(reduce+apply)", lambda i: lib.cb2_bn_backward(
        dys[i % nbuf].data_ptr(), None, xs[i % nbuf].data_ptr(), save.data_ptr(), save[c:].data_ptr(),
        g.data_ptr(), b.data_ptr(), sums.data_ptr(), dbg.data_ptr(), dbg[c:].data_ptr(), dx.data_ptr(), None,
        m, c, relu, st), 5 * sz)
    out = f"N={batch} {hw}x{hw}x{c} ({sz / 1e6:.0f} MB): "
    for k, (us, by) in res.items():
        out += f"{k} {us:.1f} us ({by / us / 1e6:.2f} TB/s)  "
    print(out, flush=True)


for batch in (64, 256):
    for (hw, c) in ((56, 256), (56, 64), (28, 512), (14, 1024), (7, 2048)):
        run(batch, hw, c)
