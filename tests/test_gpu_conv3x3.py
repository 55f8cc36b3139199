"""GPU: the 3x3 implicit-GEMM convolution kernel (csrc/kernels/conv3x3_tcgen05.cu, ops/conv.py)
against a plain fp32 PyTorch reference - forward, input gradient, the fused batch-norm statistics,
on shapes that exercise every scheduling case: images that are not a whole number of tiles, sub-tiles
past the last image row, the tail of the schedule split into single sub-tiles, 64- and 128-wide
channel tiles, non-square images.  (Reference workload: conv*_block*_2_conv of ResNet-50,
core/tests/examples/call_run_within_script_with_keras_fit.py:79-83.)"""
import pytest
import torch
import torch.nn.functional as F

from cloud_b200 import ops
from cloud_b200.ops import conv as conv_mod

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


def _make(b, h, w, cin, cout, seed=0):
    torch.manual_seed(seed)
    dev = _dev()
    x = torch.randn(b, cin, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    k = (torch.randn(cout, cin, 3, 3, device=dev) / (3 * cin ** 0.5)).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last)
    dy = torch.randn(b, cout, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    return x, k, dy


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


SHAPES = [
    (3, 14, 14, 64, 64),        # one tile per image, second sub-tile partly past the image
    (5, 28, 28, 128, 128),      # 3.5 tiles per image: the last sub-tile of every image is empty
    (2, 56, 56, 64, 128),       # two image rows per sub-tile, cin != cout
    (150, 14, 14, 128, 128),    # 150 tiles on 148 SMs: the last two run as single sub-tiles
    (3, 10, 12, 192, 64),       # not square, three channel blocks, 64-wide output tile
    (2, 9, 30, 64, 256),        # two output-channel tiles
    (1, 7, 7, 128, 128),        # an image smaller than one sub-tile
]


@pytest.mark.parametrize("b,h,w,cin,cout", SHAPES)
def test_forward_and_input_gradient_match_fp32_reference(b, h, w, cin, cout):
    x, k, dy = _make(b, h, w, cin, cout, seed=h * w + cin)
    want = F.conv2d(x.float(), k.float(), padding=1)
    got = conv_mod._conv3_run(x, k, cout, False)
    assert got.shape == want.shape and got.is_contiguous(memory_format=torch.channels_last)
    assert _rel(got, want) < 6e-3
    assert (got.float() - want).abs().max().item() < 0.06 * want.abs().max().item()
    want_dx = torch.ops.aten.convolution_backward(dy.float(), x.float(), k.float(), None, (1, 1), (1, 1), (1, 1),
                                                  False, (0, 0), 1, (True, False, False))[0]
    got_dx = conv_mod._conv3_run(dy, k, cin, True)
    assert _rel(got_dx, want_dx) < 6e-3
    assert (got_dx.float() - want_dx).abs().max().item() < 0.06 * want_dx.abs().max().item()


@pytest.mark.parametrize("b,h,w,cin,cout", [(5, 28, 28, 128, 128), (150, 14, 14, 128, 128), (3, 10, 12, 192, 64)])
def test_fused_statistics_cover_exactly_the_image_pixels(b, h, w, cin, cout):
    x, k, _ = _make(b, h, w, cin, cout, seed=7)
    sums = torch.zeros(8, 2, cout, device=_dev())
    y = conv_mod._conv3_run(x, k, cout, False, sums).float()
    s = sums.sum(0)
    want1, want2 = y.sum((0, 2, 3)), (y * y).sum((0, 2, 3))
    assert (s[0] - want1).abs().max().item() < 2e-3 * max(want1.abs().max().item(), 1.0)
    assert (s[1] - want2).abs().max().item() < 2e-3 * want2.abs().max().item()


def test_results_do_not_depend_on_the_run():
    x, k, dy = _make(37, 28, 28, 128, 256, seed=3)
    first = conv_mod._conv3_run(x, k, 256, False)
    for _ in range(4):
        assert torch.equal(conv_mod._conv3_run(x, k, 256, False), first)


def test_autograd_function_matches_cudnn_gradients(monkeypatch):
    monkeypatch.setattr(conv_mod, "_CONV3", "all")
    x, k, dy = _make(4, 14, 14, 128, 128, seed=11)
    x1 = x.clone().requires_grad_(True)
    k1 = k.clone().requires_grad_(True)
    assert conv_mod.conv3x3_own(x1, k1)
    y = ops.conv_kxk(x1, k1, None, bn_stats=True)
    assert hasattr(y, "_cb2_bn_sums")
    y.backward(dy)
    x2 = x.float().requires_grad_(True)
    k2 = k.float().requires_grad_(True)
    F.conv2d(x2, k2, padding=1).backward(dy.float())
    assert _rel(x1.grad, x2.grad) < 6e-3
    assert _rel(k1.grad, k2.grad) < 1e-2


def test_unsupported_geometry_is_an_error_not_a_wrong_answer():
    x, k, _ = _make(1, 8, 4, 64, 64)           # 4 columns: below the kernel's minimum
    with pytest.raises(RuntimeError, match="conv3x3"):
        conv_mod._conv3_run(x, k, 64, False)
