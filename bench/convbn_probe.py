"""Times of the folded 1x1-convolution + batch-norm layer (ops/convbn.py) against the unfolded
path (convolution, fused NHWC batch-norm + shortcut + ReLU) on the ResNet-50 conv3 shapes at a
given batch, forward + backward, CUDA events, inputs larger than L2 for the big shapes.

    python bench/convbn_probe.py [--batch 256] [--out gpurun_out/convbn_probe.json]
"""
import argparse
import json
import math
import os
import sys

import torch

os.environ.setdefault("TFC_BN_FOLD_MIN_MB", "0")

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402
from cloud_b200.ops import bn as bn_ops  # noqa: E402


ITERS, WARM = 10, 3


def timed(fn, iters=None, warm=None):
    iters = iters or ITERS
    warm = WARM if warm is None else warm
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default="gpurun_out/convbn_probe.json")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warm", type=int, default=3)
    ap.add_argument("--only", default="", help="comma list of hw values")
    args = ap.parse_args()
    global ITERS, WARM
    ITERS, WARM = args.iters, args.warm
    dev = torch.device("cuda", 0)
    bn_ops.arm_scratch(dev)
    rows = []
    for hw, cin in ((56, 64), (28, 128), (14, 256), (7, 512)):
        if args.only and str(hw) not in args.only.split(","):
            continue
        cout = 4 * cin
        n = args.batch
        M = n * hw * hw
        x = (torch.randn(n, cin, hw, hw, device=dev).relu()).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last).requires_grad_(True)
        w = (torch.randn(cout, cin, 1, 1, device=dev) / math.sqrt(cin)).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last).requires_grad_(True)
        gamma = torch.ones(cout, device=dev, requires_grad=True)
        beta = torch.zeros(cout, device=dev, requires_grad=True)
        res = torch.randn(n, cout, hw, hw, device=dev).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last).requires_grad_(True)
        rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
        dz = torch.randn(n, cout, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        out = {}

        def folded_fwd():
            bn_ops.reset_scratch(dev)
            out["z"] = ops.conv1x1_bn(x, w, gamma, beta, rm, rv, 0.01, 1e-5, relu=True, residual=res)

        def folded_bwd():
            bn_ops.reset_scratch(dev)
            z = ops.conv1x1_bn(x, w, gamma, beta, rm, rv, 0.01, 1e-5, relu=True, residual=res)
            z.backward(dz)

        def plain_fwd():
            bn_ops.reset_scratch(dev)
            y = ops.conv1x1(x, w, None, bn_stats=True)
            out["z"] = ops.batch_norm_nhwc(y, gamma, beta, rm, rv, True, 0.01, 1e-5, relu=True, residual=res,
                                           sums=getattr(y, "_cb2_bn_sums", None))

        def plain_bwd():
            bn_ops.reset_scratch(dev)
            y = ops.conv1x1(x, w, None, bn_stats=True)
            z = ops.batch_norm_nhwc(y, gamma, beta, rm, rv, True, 0.01, 1e-5, relu=True, residual=res,
                                    sums=getattr(y, "_cb2_bn_sums", None))
            z.backward(dz)

        big_mb = M * cout * 2 / 1e6
        row = {"hw": hw, "cin": cin, "cout": cout, "M": M, "out_mb": round(big_mb, 1)}
        for name, fn in (("folded_fwd_us", folded_fwd), ("plain_fwd_us", plain_fwd),
                         ("folded_fwd_bwd_us", folded_bwd), ("plain_fwd_bwd_us", plain_bwd)):
            for p in (x, w, gamma, beta, res):
                p.grad = None
            row[name] = round(timed(fn), 1)
        # the fused forward GEMM alone: bytes = x + shortcut in, z + bits out
        z2 = torch.empty(M, cout, device=dev, dtype=torch.bfloat16)
        sc, sh = torch.ones(cout, device=dev), torch.zeros(cout, device=dev)
        mask = torch.empty(M * cout // 8, device=dev, dtype=torch.uint8)
        X = x.detach().permute(0, 2, 3, 1).reshape(M, cin)
        R = res.detach().permute(0, 2, 3, 1).reshape(M, cout)
        us = timed(lambda: ops.convbn.fused_gemm(X, w.detach().reshape(cout, cin), z2, scale=sc, shift=sh,
                                                 residual=R, relu=True, mask_out=mask))
        nbytes = M * cin * 2 + 2 * M * cout * 2 + M * cout / 8
        row["fused_gemm_us"] = round(us, 1)
        row["fused_gemm_tbs"] = round(nbytes / us / 1e6, 2)
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
