"""GPU diagnostic for the tcgen05 GEMM: correctness of the three operand-major combinations, a
descriptor-field sweep if the defaults are wrong (so one gpurun call answers the question), and
throughput against torch.matmul (cuBLAS) on large shapes.  Writes gpurun_out/gemm_selftest.json.
"""
import itertools
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402
from cloud_b200.ops import _lib  # noqa: E402

dev = torch.device("cuda:0")
OUT = {"correctness": {}, "sweeps": {}, "perf": []}


def ref(a, b, a_mn, b_mn):
    af = a.float().t() if a_mn else a.float()
    bf = b.float() if b_mn else b.float().t()
    return af @ bf


def run_case(M, N, K, a_mn, b_mn, block_n=0):
    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).to(torch.bfloat16)
    try:
        got = ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32, block_n=block_n)
        torch.cuda.synchronize()
    except Exception as e:  # launch failure / trap
        return {"error": str(e)[:200]}
    want = ref(a, b, a_mn, b_mn)
    err = (got - want).abs().max().item()
    rel = err / (want.abs().max().item() + 1e-6)
    return {"max_abs_err": err, "rel": rel, "ok": bool(rel < 2e-2)}


def main():
    print("device:", torch.cuda.get_device_name(0), flush=True)
    modes = {"kk": (False, False), "k_mn": (False, True), "mn_mn": (True, True), "mn_k": (True, False)}
    all_ok = True
    for name, (a_mn, b_mn) in modes.items():
        for (M, N, K) in [(128, 128, 64), (128, 128, 256), (256, 256, 512), (1000, 264, 520)]:
            for bn in (64, 128, 256):
                r = run_case(M, N, K, a_mn, b_mn, bn)
                OUT["correctness"][f"{name}_{M}x{N}x{K}_bn{bn}"] = r
                ok = r.get("ok", False)
                all_ok &= ok
                print(name, M, N, K, "bn", bn, r, flush=True)
                if "error" in r:
                    break

    # descriptor sweep for failing modes (fields in 16-byte units)
    lib = _lib.kernels()
    for name, (a_mn, b_mn) in modes.items():
        key = f"{name}_128x128x256_bn128"
        if OUT["correctness"].get(key, {}).get("ok"):
            continue
        print("sweeping descriptors for", name, flush=True)
        found = []
        k_opts = {"lbo": [1, 0, 64], "sbo": [64, 128], "kadv": [2]}
        mn_opts = {"lbo": [512, 64, 128, 1024], "sbo": [64, 512, 128, 1024], "kadv": [128, 256, 64]}
        ao = mn_opts if a_mn else k_opts
        bo = mn_opts if b_mn else k_opts
        for al, as_, ak, bl, bs, bk in itertools.product(ao["lbo"], ao["sbo"], ao["kadv"],
                                                         bo["lbo"], bo["sbo"], bo["kadv"]):
            lib.cb2_gemm_debug_descriptors(1, al, as_, bl, bs, ak, bk)
            r = run_case(128, 128, 256, a_mn, b_mn, 128)
            if "error" in r:
                found.append({"cfg": [al, as_, ak, bl, bs, bk], "error": r["error"]})
                break
            if r["ok"]:
                found.append({"cfg": [al, as_, ak, bl, bs, bk], "rel": r["rel"]})
                print("  OK with a_lbo,a_sbo,a_kadv,b_lbo,b_sbo,b_kadv =", al, as_, ak, bl, bs, bk, flush=True)
        lib.cb2_gemm_debug_descriptors(0, 0, 0, 0, 0, 0, 0)
        OUT["sweeps"][name] = found

    # throughput
    if all_ok:
        for (M, N, K) in [(8192, 8192, 8192), (4096, 4096, 4096), (200704, 256, 64), (200704, 64, 256),
                          (12544, 2048, 512), (128, 512, 784)]:
            a = torch.randn(M, K, device=dev).to(torch.bfloat16)
            b = torch.randn(N, K, device=dev).to(torch.bfloat16)
            for label, fn in (("tcgen05", lambda: ops.gemm(a, b)),
                              ("cublas", lambda: torch.matmul(a, b.t()))):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                iters = 20
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / iters
                tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
                OUT["perf"].append({"shape": [M, N, K], "impl": label, "ms": ms, "tflops": tf})
                print(f"{label:8s} {M}x{N}x{K}: {ms:.4f} ms  {tf:.1f} TFLOP/s", flush=True)
    # ResNet-50 1x1 convolutions (NHWC activations are plain [N*H*W, C] matrices): our GEMM vs
    # cuDNN through torch.conv2d on channels_last tensors; these are bandwidth-bound shapes
    if all_ok:
        OUT["conv1x1"] = []
        import torch.nn.functional as F
        for (n, hw, cin, cout) in [(64, 56, 64, 64), (64, 56, 64, 256), (64, 56, 256, 64), (64, 28, 128, 512),
                                   (64, 28, 512, 128), (64, 14, 256, 1024), (64, 14, 1024, 256),
                                   (64, 7, 512, 2048), (64, 7, 2048, 512)]:
            M = n * hw * hw
            x = torch.randn(n, cin, hw, hw, device=dev).to(torch.bfloat16).contiguous(
                memory_format=torch.channels_last)
            w = torch.randn(cout, cin, 1, 1, device=dev).to(torch.bfloat16).contiguous(
                memory_format=torch.channels_last)
            xm = x.permute(0, 2, 3, 1).reshape(M, cin)
            wm = w.reshape(cout, cin)
            want = F.conv2d(x, w).permute(0, 2, 3, 1).reshape(M, cout).float()
            got = ops.gemm(xm, wm).float()
            rel = ((got - want).abs().max() / want.abs().max()).item()
            row = {"shape": [M, cout, cin], "rel_err": rel}
            for label, fn in (("tcgen05", lambda: ops.gemm(xm, wm)), ("cudnn", lambda: F.conv2d(x, w))):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                iters = 20
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / iters
                gbs = 2.0 * (M * cin + M * cout + cin * cout) / (ms * 1e-3) / 1e9
                row[label + "_us"] = ms * 1e3
                row[label + "_gbs"] = gbs
            OUT["conv1x1"].append(row)
            print(f"conv1x1 M={M} cin={cin} cout={cout}: rel {rel:.2e}  tcgen05 {row['tcgen05_us']:.1f} us "
                  f"({row['tcgen05_gbs']:.0f} GB/s)  cudnn {row['cudnn_us']:.1f} us ({row['cudnn_gbs']:.0f} GB/s)",
                  flush=True)
    OUT["all_ok"] = all_ok
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/gemm_selftest.json", "w") as f:
        json.dump(OUT, f, indent=1)
    print("ALL_OK" if all_ok else "SOME_FAILED")


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"selftest took {time.time() - t0:.1f}s")
