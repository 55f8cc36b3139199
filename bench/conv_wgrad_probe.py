"""3x3 convolution weight gradients of ResNet-50 (batch 256): tcgen05 shifted-box kernel vs cuDNN."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.ops import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.kernels()
B = int(os.environ.get("BATCH", "256"))


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


for (hw, c) in ((56, 64), (28, 128), (14, 256), (7, 512)):
    x = torch.randn(B, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(B, c, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(c, c, 3, 3, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ws = torch.zeros(9 * c, c, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    mine = timed(lambda: lib.cb2_conv_wgrad(x.data_ptr(), dy.data_ptr(), ws.data_ptr(), B, hw, hw, c, c, 3, 1, 1, st))
    cudnn = timed(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (1, 1), (1, 1), (1, 1), False, (0, 0),
                                                              1, (False, True, False)))
    print(f"3x3 wgrad {hw}x{hw}x{c} batch {B}: tcgen05 {mine:.0f} us  cudnn {cudnn:.0f} us", flush=True)

# strided 1x1 convolutions (first conv and shortcut of the first block of stages 2-4)
for (hw, cin, cout) in ((56, 256, 128), (56, 256, 512), (28, 512, 256), (28, 512, 1024), (14, 1024, 512),
                        (14, 1024, 2048)):
    oh = hw // 2
    x = torch.randn(B, cin, hw, hw, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(B, cout, oh, oh, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(cout, cin, 1, 1, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ws = torch.zeros(cin, cout, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    mine = timed(lambda: lib.cb2_conv_wgrad(x.data_ptr(), dy.data_ptr(), ws.data_ptr(), B, oh, oh, cin, cout, 1, 0,
                                            2, st))
    cudnn = timed(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (2, 2), (0, 0), (1, 1), False, (0, 0),
                                                              1, (False, True, False)))
    ws.zero_()
    lib.cb2_conv_wgrad(x.data_ptr(), dy.data_ptr(), ws.data_ptr(), B, oh, oh, cin, cout, 1, 0, 2, st)
    ref = torch.ops.aten.convolution_backward(dy, x, w, None, (2, 2), (0, 0), (1, 1), False, (0, 0), 1,
                                              (False, True, False))[1].float().view(cout, cin).t()
    err = ((ws - ref).abs().max() / ref.abs().max()).item()
    print(f"1x1/2 wgrad {hw}x{hw}x{cin}->{cout} batch {B}: tcgen05 {mine:.0f} us  cudnn {cudnn:.0f} us  rel err {err:.1e}",
          flush=True)
