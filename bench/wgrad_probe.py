"""1x1-convolution weight gradients (dW = dY^T @ X, reduction over N*H*W) on the tcgen05 GEMM:
sweep of tile width and split-K factor per ResNet-50 shape, against cuBLAS (torch.matmul)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402

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


for (cout, cin, k) in [(256, 64, 200704), (64, 256, 200704), (64, 64, 200704), (512, 128, 50176),
                       (128, 512, 50176), (1024, 256, 12544), (256, 1024, 12544), (2048, 512, 3136),
                       (512, 2048, 3136)]:
    dy = torch.randn(k, cout, device=dev).to(torch.bfloat16)
    x = torch.randn(k, cin, device=dev).to(torch.bfloat16)
    out = torch.zeros(cout, cin, device=dev)
    best = None
    res = []
    for bn in (64, 128, 256):
        if bn > max(cin, 64):
            continue
        tiles = ((cout + 127) // 128) * ((cin + bn - 1) // bn)
        for mult in (0.5, 1, 2, 4):
            sp = max(1, int(148 * mult) // tiles)
            sp = min(sp, (k // 64) // 2)
            us = timed(lambda: ops.gemm(dy, x, a_mn=True, b_mn=True, out=out, accumulate=True,
                                        block_n=bn, split_k=sp))
            res.append((us, bn, sp))
    res.sort()
    auto = timed(lambda: ops.gemm(dy, x, a_mn=True, b_mn=True, out=out, accumulate=True, split_k=0))
    cublas = timed(lambda: torch.matmul(dy.t(), x))
    print(f"dW[{cout}x{cin}] K={k}: auto {auto:.1f} us | best {res[0][0]:.1f} us (bn{res[0][1]} split{res[0][2]})"
          f" | cublas {cublas:.1f} us | all: " + " ".join(f"bn{b}/s{s}:{u:.1f}" for u, b, s in res), flush=True)
