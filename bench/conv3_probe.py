"""3x3 convolution kernel (csrc/kernels/conv3x3_tcgen05.cu) against cuDNN on the ResNet-50 shapes:
numerics against an fp32 reference and device time of forward / input gradient.

    python bench/conv3_probe.py [--batch 256] [--check-only]
"""
import argparse
import ctypes
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.ops import _lib  # noqa: E402


def lib():
    k = _lib.kernels()
    vp, i = ctypes.c_void_p, ctypes.c_int
    k.cb2_conv3x3.restype = i
    k.cb2_conv3x3.argtypes = [vp, vp, vp, i, i, i, i, i, i, vp, vp]
    k.cb2_conv3x3_debug.argtypes = [i] * 5
    return k


def ours(k, x_nhwc, w_ohwi, cout, grad, stats=None):
    b, h, w, c = x_nhwc.shape
    cin = w_ohwi.shape[3]
    n = cin if grad else cout
    y = torch.empty(b, h, w, n, device=x_nhwc.device, dtype=torch.bfloat16)
    rc = k.cb2_conv3x3(x_nhwc.data_ptr(), w_ohwi.data_ptr(), y.data_ptr(), b, h, w, cin, cout, int(grad),
                       0 if stats is None else stats.data_ptr(), torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(k.cb2_last_error().decode())
    return y


def timed(fn, iters=5, rot=1):
    """Median device time of one call, in us.  ``fn(i)`` is called with i = 0..rot-1 between ONE pair
    of events (the event clock ticks in ~2 us steps - too coarse for one 50 us kernel); the rot
    operand sets together are larger than L2, so every call starts cold."""
    for i in range(rot):
        fn(i)
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(rot):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / rot)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--check-only", action="store_true")
    ap.add_argument("--base-off", type=int, default=None)
    ap.add_argument("--sweep", action="store_true", help="pipeline depths / skipped traffic, forward only")
    args = ap.parse_args()
    k = lib()
    torch.manual_seed(0)
    dev = torch.device("cuda")
    out = {"batch": args.batch, "rows": []}
    modes = [args.base_off] if args.base_off is not None else [0, 1]
    # which descriptor mode is right: small problem, both modes
    xb = torch.randn(2, 14, 14, 64, device=dev).to(torch.bfloat16)
    wb = (torch.randn(64, 3, 3, 64, device=dev) * 0.05).to(torch.bfloat16)
    ref = F.conv2d(xb.float().permute(0, 3, 1, 2), wb.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    good = None
    for m in modes:
        k.cb2_conv3x3_debug(m, 0, 0, 0, 0)
        y = ours(k, xb, wb, 64, False).float()
        torch.cuda.synchronize()
        err = ((y - ref).norm() / ref.norm()).item()
        out[f"mode{m}_rel_err"] = err
        if err < 2e-2 and good is None:
            good = m
    if good is None:
        print(json.dumps(out))
        sys.exit(1)
    k.cb2_conv3x3_debug(good, 0, 0, 0, 0)
    out["base_off_mode"] = good
    shapes = [(56, 64), (28, 128), (14, 256), (7, 512)]
    for hw, c in shapes:
        b = args.batch
        x = torch.randn(b, hw, hw, c, device=dev).to(torch.bfloat16)
        w = (torch.randn(c, 3, 3, c, device=dev) * (1.0 / (3 * c ** 0.5))).to(torch.bfloat16)
        dy = torch.randn(b, hw, hw, c, device=dev).to(torch.bfloat16)
        xc = x.permute(0, 3, 1, 2)            # channels_last views
        wc = w.permute(0, 3, 1, 2)
        dyc = dy.permute(0, 3, 1, 2)
        row = {"hw": hw, "c": c}
        # numerics on a slice of the batch (fp32 reference is expensive)
        nb = min(b, 8)
        ref = F.conv2d(xc[:nb].float(), wc.float(), padding=1).permute(0, 2, 3, 1)
        stats = torch.zeros(8, 2, c, device=dev)
        y = ours(k, x, w, c, False, stats)
        row["fwd_rel_err"] = ((y[:nb].float() - ref).norm() / ref.norm()).item()
        yf = y.float()
        s = stats.sum(0)
        row["stats_sum_err"] = ((s[0] - yf.sum((0, 1, 2))).abs().max() / yf.sum((0, 1, 2)).abs().max()).item()
        row["stats_sq_err"] = ((s[1] - (yf * yf).sum((0, 1, 2))).abs().max() / (yf * yf).sum((0, 1, 2)).max()).item()
        refg = torch.ops.aten.convolution_backward(
            dyc[:nb].float(), xc[:nb].float(), wc.float(), None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
            (True, False, False))[0].permute(0, 2, 3, 1)
        dx = ours(k, dy, w, c, True)
        row["dgrad_rel_err"] = ((dx[:nb].float() - refg).norm() / refg.norm()).item()
        # the whole batch against cuDNN's bf16 result, several runs: a race shows up as a handful of
        # elements far off, or as two runs that differ
        yc = F.conv2d(xc, wc, padding=1).permute(0, 2, 3, 1).float()
        dxc = torch.ops.aten.convolution_backward(dyc, xc, wc, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                                  (True, False, False))[0].permute(0, 2, 3, 1).float()
        tol_f, tol_g = 0.05 * yc.abs().max().item(), 0.05 * dxc.abs().max().item()
        bad_f = bad_g = 0
        first_f = first_g = None
        for it in range(6):
            yy = ours(k, x, w, c, False, stats if it % 2 else None)
            gg = ours(k, dy, w, c, True)
            bad_f = max(bad_f, int(((yy.float() - yc).abs() > tol_f).sum().item()))
            bad_g = max(bad_g, int(((gg.float() - dxc).abs() > tol_g).sum().item()))
            if first_f is None:
                first_f, first_g = yy.clone(), gg.clone()
            else:
                row["nondeterministic"] = row.get("nondeterministic", 0) + int((yy != first_f).sum().item()) + \
                    int((gg != first_g).sum().item())
        row["bad_elements_fwd"], row["bad_elements_dgrad"] = bad_f, bad_g
        del yc, dxc, first_f, first_g
        # rotating operand sets: together larger than L2 (126 MB)
        per = 2 * x.numel() * 2
        rot = max(2, min(16, (300 << 20) // per + 1))
        xs = [x] + [torch.randn_like(x) for _ in range(rot - 1)]
        dys = [dy] + [torch.randn_like(dy) for _ in range(rot - 1)]
        row["rot"] = rot
        if args.sweep:
            sw = {}
            for (a_st, b_st, ob, skip) in [(2, 0, 2, 0), (2, 0, 1, 0), (2, 0, 2, 16), (2, 0, 2, 7), (2, 0, 2, 8),
                                           (2, 0, 2, 15), (2, 0, 2, 0)]:
                k.cb2_conv3x3_debug(good, a_st, b_st, ob, skip)
                key = f"a{a_st}_b{b_st}_o{ob}_skip{skip}"
                while key in sw:
                    key += "'"
                try:
                    sw[key] = round(timed(lambda i: ours(k, xs[i], w, c, False), 5, rot), 2)
                except RuntimeError as e:
                    sw[key] = str(e)[:60]
            k.cb2_conv3x3_debug(good, 0, 0, 0, 0)
            row["sweep_us"] = sw
        if not args.check_only:
            flops = 2.0 * b * hw * hw * c * c * 9
            t = timed(lambda i: ours(k, xs[i], w, c, False), 5, rot)
            row["fwd_us"], row["fwd_tflops"] = round(t, 2), round(flops / t / 1e6, 1)
            row["fwd_stats_us"] = round(timed(lambda i: ours(k, xs[i], w, c, False, stats), 5, rot), 2)
            t = timed(lambda i: ours(k, dys[i], w, c, True), 5, rot)
            row["dgrad_us"], row["dgrad_tflops"] = round(t, 2), round(flops / t / 1e6, 1)
            torch.backends.cudnn.benchmark = True
            xcs = [t_.permute(0, 3, 1, 2) for t_ in xs]
            dycs = [t_.permute(0, 3, 1, 2) for t_ in dys]
            row["cudnn_fwd_us"] = round(timed(lambda i: F.conv2d(xcs[i], wc, padding=1), 5, rot), 2)
            row["cudnn_dgrad_us"] = round(timed(lambda i: torch.ops.aten.convolution_backward(
                dycs[i], xcs[i], wc, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1, (True, False, False)), 5, rot), 2)
        del xs, dys
        out["rows"].append(row)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
