"""The algorithm of the 3x3 implicit-GEMM kernel (csrc/kernels/conv3x3_tcgen05.cu) restated in numpy
and checked against torch on the CPU: the padded-raster strip, the nine taps as constant row shifts
into it, sub-tiles that over-read into rows nobody stores, the clipped store, the work-unit schedule
(whole tiles, then the tail as single sub-tiles; sub-tiles past the image skipped; whole waves
rotated), and the input gradient as the same computation with mirrored taps and the weight tile read
transposed.  The GPU tests check the kernel; this checks that what the kernel is meant to compute is
a convolution - for every geometry, without a GPU."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F


def _schedule(B, H, W, n_tiles, sms):
    """(image, first row, channel tile, active sub-tiles) of every work unit, as c3_unit() yields them."""
    Wp = W + 2
    rsub = min(128 // Wp, H)
    rt = 2 * rsub
    tiles_per_image = -(-H // rt)
    m_tiles = B * tiles_per_image
    tiles = m_tiles * n_tiles
    rem = tiles % sms if tiles > sms else 0
    full = tiles - rem if (rem > 0 and 2 * rem <= sms) else tiles
    total = full + 2 * (tiles - full)
    grid = min(total, sms)
    units = []
    for u in range(total):
        tile, half = u, -1
        if u >= full:
            x = u - full
            tile, half = full + (x >> 1), x & 1
        else:
            wave = u // grid
            if (wave + 1) * grid <= full:
                tile = wave * grid + (u - wave * grid + wave) % grid
        mt = tile % m_tiles
        h0 = (mt % tiles_per_image) * rt + (rsub if half == 1 else 0)
        on = (1 if h0 < H else 0) | (2 if (half < 0 and h0 + rsub < H) else 0)
        units.append((mt // tiles_per_image, h0, tile // m_tiles, on))
    return units, rsub, rt


def _conv3x3_as_the_kernel_does(x, w, input_gradient, sms=148, rng=None):
    """x [B,H,W,K] float, w [Cout,3,3,Cin] -> y [B,H,W,N].  Rows the kernel never stores are filled
    from `rng` garbage, as uninitialised shared memory would."""
    B, H, W, K = x.shape
    cout, _, _, cin = w.shape
    N = cin if input_gradient else cout
    bn = 128 if N % 128 == 0 else 64
    Wp = W + 2
    units, rsub, rt = _schedule(B, H, W, N // bn, sms)
    y = np.full((B, H, W, N), np.nan, dtype=np.float64)
    written = np.zeros((B, H, W, N), dtype=np.int32)
    rows_in_stage = (rsub + 2) * Wp + 2 + 128
    for (b, h0, nt, on) in units:
        n0 = nt * bn
        acc = [np.zeros((128, bn)), np.zeros((128, bn))]
        for cb in range(K // 64):
            # the TMA box: rows h0-1 .. h0+rt, columns -1 .. W: zeros outside the image; the stage is
            # larger than the box and holds garbage behind it
            stage = rng.standard_normal((rows_in_stage, 64)) * 1e3
            strip = np.zeros((rt + 2, Wp, 64))
            for r in range(rt + 2):
                hh = h0 - 1 + r
                if 0 <= hh < H:
                    strip[r, 1:W + 1] = x[b, hh, :, cb * 64:(cb + 1) * 64]
            stage[:(rt + 2) * Wp] = strip.reshape(-1, 64)
            for tap in range(9):
                dy, dx = divmod(tap, 3)
                wt = 8 - tap if input_gradient else tap
                if not input_gradient:      # [bn output channels, 64 input channels] of this tap
                    wtile = w[n0:n0 + bn, wt // 3, wt % 3, cb * 64:(cb + 1) * 64]
                else:                       # [bn input channels, 64 output channels]: the tile transposed
                    wtile = w[cb * 64:(cb + 1) * 64, wt // 3, wt % 3, n0:n0 + bn].T
                for s in range(2):
                    if not (on >> s) & 1:
                        continue
                    off = (s * rsub + dy) * Wp + dx
                    acc[s] += stage[off:off + 128] @ wtile.T
        for s in range(2):
            if not (on >> s) & 1:
                continue
            tile = acc[s][:rsub * Wp].reshape(rsub, Wp, bn)
            for r in range(rsub):                     # the store box: W + 2 wide over a tensor W wide
                hh = h0 + s * rsub + r
                if hh < H:
                    y[b, hh, :, n0:n0 + bn] = tile[r, :W]
                    written[b, hh, :, n0:n0 + bn] += 1
    return y, written


GEOMETRIES = [(2, 14, 14, 64, 64), (1, 28, 28, 64, 128), (1, 9, 30, 128, 64), (3, 7, 7, 64, 64),
              (1, 56, 56, 64, 64), (2, 10, 12, 64, 192)]


@pytest.mark.parametrize("B,H,W,cin,cout", GEOMETRIES)
@pytest.mark.parametrize("sms", [148, 3])
def test_strip_and_tap_shifts_compute_the_convolution(B, H, W, cin, cout, sms):
    rng = np.random.default_rng(H * W + cin + sms)
    x = rng.standard_normal((B, H, W, cin))
    w = rng.standard_normal((cout, 3, 3, cin)) / (3 * np.sqrt(cin))
    y, written = _conv3x3_as_the_kernel_does(x, w, False, sms, rng)
    assert (written == 1).all()                       # every pixel stored exactly once, by one unit
    want = F.conv2d(torch.tensor(x).permute(0, 3, 1, 2), torch.tensor(w).permute(0, 3, 1, 2), padding=1)
    np.testing.assert_allclose(y, want.permute(0, 2, 3, 1).numpy(), rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("B,H,W,cin,cout", GEOMETRIES[:4])
def test_mirrored_taps_and_transposed_weight_tiles_compute_the_input_gradient(B, H, W, cin, cout):
    rng = np.random.default_rng(7 * H + W)
    dy = rng.standard_normal((B, H, W, cout))
    w = rng.standard_normal((cout, 3, 3, cin)) / (3 * np.sqrt(cin))
    dx, written = _conv3x3_as_the_kernel_does(dy, w, True, 148, rng)
    assert (written == 1).all()
    x = torch.zeros(B, cin, H, W, dtype=torch.float64, requires_grad=True)
    out = F.conv2d(x, torch.tensor(w).permute(0, 3, 1, 2), padding=1)
    out.backward(torch.tensor(dy).permute(0, 3, 1, 2))
    np.testing.assert_allclose(dx, x.grad.permute(0, 2, 3, 1).numpy(), rtol=1e-9, atol=1e-9)


def test_tail_of_the_schedule_is_split_only_when_it_pays():
    # 150 tiles on 148 SMs: the last two tiles run as four single sub-tiles
    units, rsub, _ = _schedule(150, 14, 14, 1, 148)
    assert len(units) == 148 + 4 and [u[3] for u in units[-4:]] == [1, 1, 1, 1]
    assert [u[1] for u in units[-4:]] == [0, rsub, 0, rsub]
    # 1024 tiles: the remainder (136) fills more than half of the SMs - no split
    units, _, _ = _schedule(256, 28, 28, 1, 148)
    assert len(units) == 1024
    # 28 x 28: the fourth tile of every image has nothing in its second sub-tile
    assert sum(1 for u in units if u[3] == 1) == 256
    # whole waves are rotated: the CTA that takes the cheap tile changes from wave to wave
    first_cta = [units[wave * 148][1] for wave in range(6)]
    assert len(set(first_cta)) > 1
