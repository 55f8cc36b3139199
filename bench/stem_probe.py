"""ResNet stem convolution (7x7/2, 3->64) forward+backward through cuDNN with the input channels
padded to 3 / 4 / 8: which layout avoids cuDNN's slow small-C paths?"""
import torch
import torch.nn.functional as F

dev = torch.device("cuda:0")
B = 256


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


for c in (3, 4, 8):
    x = torch.randn(B, c, 224, 224, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(64, c, 7, 7, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w.requires_grad_(True)
    y = F.conv2d(x, w, None, stride=2, padding=3)
    gy = torch.randn_like(y)
    fwd = timed(lambda: F.conv2d(x, w, None, stride=2, padding=3))

    def wgrad():
        return torch.ops.aten.convolution_backward(gy, x, w, None, (2, 2), (3, 3), (1, 1), False, (0, 0), 1,
                                                   (False, True, False))
    wg = timed(wgrad)
    print(f"C={c}: fprop {fwd:.0f} us  wgrad {wg:.0f} us", flush=True)
# im2col alternative for wgrad: unfold + GEMM
x = torch.randn(B, 3, 224, 224, device=dev).to(torch.bfloat16)
print("unfold (im2col) time us:", timed(lambda: F.unfold(x, 7, padding=3, stride=2)))

# space-to-depth formulation: 7x7/2 on 224x224x3  ==  4x4/1 on 115x115x(12 -> 16)
for c in (12, 16):
    xs = torch.randn(B, c, 115, 115, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ws = torch.randn(64, c, 4, 4, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ys = F.conv2d(xs, ws)
    gys = torch.randn_like(ys)
    f2 = timed(lambda: F.conv2d(xs, ws))
    w2 = timed(lambda: torch.ops.aten.convolution_backward(gys, xs, ws, None, (1, 1), (0, 0), (1, 1), False, (0, 0), 1,
                                                           (False, True, False)))
    print(f"s2d C={c}: fprop {f2:.0f} us  wgrad {w2:.0f} us  out {tuple(ys.shape)}", flush=True)
x3 = torch.randn(B, 3, 224, 224, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)


def s2d(x):
    p = F.pad(x, (4, 2, 4, 2))                                   # [B,3,230,230]
    p = p.permute(0, 2, 3, 1).reshape(B, 115, 2, 115, 2, 3)      # NHWC view
    p = p.permute(0, 1, 3, 2, 4, 5).reshape(B, 115, 115, 12)
    p = F.pad(p, (0, 4))
    return p.permute(0, 3, 1, 2)


print("s2d transform us:", timed(lambda: s2d(x3)), s2d(x3).is_contiguous(memory_format=torch.channels_last))

# the implicit GEMM on the tcgen05 kernel (cloud_b200.ops.conv.stem_conv)
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops
from cloud_b200.ops import conv as convmod
w3 = (torch.randn(64, 3, 7, 7, device=dev) / 12).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w3.requires_grad_(True)
xs = convmod.space_to_depth_stem(x3)
wm = convmod.stem_weight_matrix(w3)
print("tcgen05 stem: s2d", round(timed(lambda: convmod.space_to_depth_stem(x3))), "us; fprop",
      round(timed(lambda: convmod._StemFn.apply(xs, wm.detach()))), "us")
yk = convmod._StemFn.apply(xs, wm)
gk = torch.randn_like(yk)
dwt = torch.zeros(256, 64, device=dev)
from cloud_b200.ops import _lib
import cloud_b200.ops.dense as dmod
gkc = gk.contiguous(memory_format=torch.channels_last)
print("tcgen05 stem wgrad", round(timed(lambda: _lib.kernels().cb2_conv_stem_wgrad(
    xs.data_ptr(), gkc.data_ptr(), dwt.data_ptr(), B, 112, 112, torch.cuda.current_stream().cuda_stream))), "us")
