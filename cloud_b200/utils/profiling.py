"""Tracing hooks (the reference has none beyond user-level TensorBoard
callbacks): NVTX ranges around forward / loss / backward / optimizer when ``TFC_NVTX=1`` so an
Nsight timeline shows the step structure, and a helper that wraps a command in ``ncu`` for the
GPU box (``python -m cloud_b200.utils.profiling --ncu -- python bench.py ...``)."""
from __future__ import annotations

import contextlib
import os
import shutil
import subprocess
import sys

_ENABLED = os.environ.get("TFC_NVTX") == "1"


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _ENABLED:
        yield
        return
    import torch
    if not torch.cuda.is_available():
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def ncu_command(cmd, out="gpurun_out/prof", kernel_regex=None, skip=0, count=3, full=True):
    """The ncu invocation of /opt/skills/guides/B200_PROFILING.md for `cmd` (a list)."""
    ncu = shutil.which("ncu") or "/usr/local/cuda/bin/ncu"
    args = [ncu, "--clock-control", "none"]
    if full:
        args += ["--set", "full", "--import-source", "on", "-o", out]
    else:
        args += ["--metrics", "gpu__time_duration.sum", "--csv", "--log-file", out + ".csv"]
    if kernel_regex:
        args += ["-k", "regex:" + kernel_regex]
    args += ["-s", str(skip), "-c", str(count)]
    return args + list(cmd)


if __name__ == "__main__":
    argv = sys.argv[1:]
    if "--" not in argv:
        print("usage: python -m cloud_b200.utils.profiling [--ncu|--launches] [-k REGEX] -- <cmd>")
        sys.exit(2)
    i = argv.index("--")
    opts, cmd = argv[:i], argv[i + 1:]
    regex = opts[opts.index("-k") + 1] if "-k" in opts else None
    sys.exit(subprocess.call(ncu_command(cmd, kernel_regex=regex, full="--launches" not in opts)))
