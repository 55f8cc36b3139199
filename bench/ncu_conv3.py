"""Two launches of the 3x3 implicit-GEMM convolution (forward with statistics, input gradient) on the
28 x 28 x 128 and 14 x 14 x 256 layers of ResNet-50 at batch 256, for an `ncu --set full` capture:

    python bench/ncu_conv3.py && ncu --set full --clock-control none --import-source on \
        -k regex:conv3x3 -c 8 -o gpurun_out/conv3 python bench/ncu_conv3.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.ops import conv as conv_mod  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
for hw, c in ((28, 128), (14, 256)):
    x = torch.randn(256, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(c, c, 3, 3, device=dev) / (3 * c ** 0.5)).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last)
    sums = torch.zeros(8, 2, c, device=dev)
    for _ in range(2):
        y = conv_mod._conv3_run(x, w, c, False, sums)
        dx = conv_mod._conv3_run(y, w, c, True)
torch.cuda.synchronize()
print("ncu conv3 targets done")
