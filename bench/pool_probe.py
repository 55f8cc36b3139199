"""Stem max-pool forward/backward timing at batch 256 (256 x 112 x 112 x 64 -> 56 x 56)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
x = torch.rand(256, 64, 112, 112, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
x.requires_grad_(True)


def timed(fn, iters=10):
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


y = ops.max_pool_nhwc(x, 3, 2, 1)
gy = torch.randn_like(y)
print("fwd us", round(timed(lambda: ops.max_pool_nhwc(x, 3, 2, 1))))
print("bwd us", round(timed(lambda: y.backward(gy, retain_graph=True))))
