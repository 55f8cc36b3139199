"""Where does the GEMM epilogue spend its time?  Times output-bound shapes with parts of the
epilogue switched off (results are wrong on purpose in those runs):
bit 1 = no tcgen05.ld, 2 = no st.shared, 4 = no TMA store, 8 = no proxy fence, 16 = no epilogue.
With --one it launches a single shape a few times (target for an ncu source-level capture).
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402
from cloud_b200.ops import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.kernels()


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


def main():
    shapes = [(200704, 256, 64), (50176, 512, 128), (12544, 2048, 512)]
    if "--one" in sys.argv:
        M, N, K = shapes[0]
        a = torch.randn(M, K, device=dev).to(torch.bfloat16)
        b = torch.randn(N, K, device=dev).to(torch.bfloat16)
        for _ in range(3):
            ops.gemm(a, b)
        torch.cuda.synchronize()
        print("one done")
        return
    for (M, N, K) in shapes:
        a = torch.randn(M, K, device=dev).to(torch.bfloat16)
        b = torch.randn(N, K, device=dev).to(torch.bfloat16)
        bias = torch.zeros(N, device=dev)
        for bn in (64, 128, 256):
            row = []
            for bits in (0, 1, 2, 4, 8, 4 | 8, 1 | 2 | 4 | 8, 16):
                lib.cb2_gemm_debug_epilogue(bits)
                row.append(f"{bits}:{timed(lambda: ops.gemm(a, b, block_n=bn)):.1f}")
            lib.cb2_gemm_debug_epilogue(0)
            wb = timed(lambda: ops.gemm(a, b, bias=bias, relu=True, block_n=bn))
            print(f"{M}x{N}x{K} bn{bn} us by skipped bits: {' '.join(row)}  | bias+relu {wb:.1f}"
                  f"  | cublas {timed(lambda: torch.matmul(a, b.t())):.1f}", flush=True)


if __name__ == "__main__":
    main()
