"""ctypes loader for the in-tree native libraries.

The kernel library is plain C ABI (``include/cloud_b200_kernels.h``), so no torch extension build is
needed: tensors are passed as raw device pointers plus the current CUDA stream.  On a GPU box a
missing library is a hard error (no silent PyTorch fallback); on a CPU-only box the ops raise
``KernelsUnavailable`` when called.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path

# CLOUD_B200_LIB_DIR: load the two libraries from another directory (A/B runs of two builds on one box)
_LIB_DIR = Path(os.environ.get("CLOUD_B200_LIB_DIR") or Path(__file__).resolve().parent.parent / "_lib")
_lock = threading.Lock()
_kernels = None
_runtime = None


class KernelsUnavailable(RuntimeError):
    pass


class CommCtx(ctypes.Structure):
    """Mirror of ``cb2_comm_ctx`` (include/cloud_b200_kernels.h)."""
    MAX_RANKS = 8
    _fields_ = [
        ("rank", ctypes.c_int),
        ("world", ctypes.c_int),
        ("grad_ptrs", ctypes.c_void_p * 8),
        ("param_ptrs", ctypes.c_void_p * 8),
        ("pad_ptrs", ctypes.c_void_p * 8),
        ("mc_grad", ctypes.c_void_p),
        ("mc_param", ctypes.c_void_p),
        ("epochs", ctypes.c_void_p),
        ("ll_ptrs", ctypes.c_void_p * 8),
        ("ll_bytes", ctypes.c_size_t),
        ("ll_state", ctypes.c_void_p),
    ]


def _maybe_build(which: str) -> None:
    if os.environ.get("CLOUD_B200_NO_AUTOBUILD"):
        return
    from .. import build
    if which == "kernels":
        build.build_kernels()
    else:
        build.build_runtime()


def _declare_kernels(lib: ctypes.CDLL) -> None:
    vp, i, u, f, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_float, ctypes.c_size_t
    u64 = ctypes.c_uint64
    lib.cb2_last_error.restype = ctypes.c_char_p
    lib.cb2_last_error.argtypes = []
    lib.cb2_version.restype = i
    lib.cb2_gemm_bf16.restype = i
    lib.cb2_gemm_bf16.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, u, i, vp]
    lib.cb2_gemm_bf16_ex.restype = i
    lib.cb2_gemm_bf16_ex.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, u, i, i, vp, vp]
    lib.cb2_gemm_colstat_copies.restype = i
    lib.cb2_gemm_splitk_bf16.restype = i
    lib.cb2_gemm_splitk_bf16.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, i, i, i, vp]
    lib.cb2_conv_wgrad.restype = i
    lib.cb2_conv_wgrad.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, vp]
    lib.cb2_conv_stem_forward.restype = i
    lib.cb2_conv_stem_forward.argtypes = [vp, vp, vp, i, i, i, vp]
    lib.cb2_conv_stem_wgrad.restype = i
    lib.cb2_conv_stem_wgrad.argtypes = [vp, vp, vp, i, i, i, vp]
    lib.cb2_gemm_debug_epilogue.restype = i
    lib.cb2_gemm_debug_epilogue.argtypes = [i]
    lib.cb2_gemm_debug_descriptors.restype = i
    lib.cb2_gemm_debug_descriptors.argtypes = [i, u, u, u, u, u, u]
    lib.cb2_softmax_xent.restype = i
    lib.cb2_softmax_xent.argtypes = [vp, i, vp, vp, vp, vp, i, i, i, f, vp]
    lib.cb2_relu_bwd_bias_grad.restype = i
    lib.cb2_relu_bwd_bias_grad.argtypes = [vp, vp, vp, vp, i, i, i, i, i, vp]
    lib.cb2_dropout.restype = i
    lib.cb2_dropout.argtypes = [vp, vp, sz, f, u64, u64, vp, i, vp]
    lib.cb2_u8_to_s2d_bf16.restype = i
    lib.cb2_u8_to_s2d_bf16.argtypes = [vp, vp, i, i, i, ctypes.POINTER(f), ctypes.POINTER(f), vp]
    lib.cb2_u8_to_bf16_normalize.restype = i
    lib.cb2_u8_to_bf16_normalize.argtypes = [vp, vp, sz, i, ctypes.POINTER(f), ctypes.POINTER(f), vp]
    lib.cb2_optim_tick.restype = i
    lib.cb2_optim_tick.argtypes = [vp, vp]
    lib.cb2_optim_flat.restype = i
    lib.cb2_optim_flat.argtypes = [i, vp, vp, i, vp, vp, vp, i, sz, vp, i, vp]
    lib.cb2_comm_fused.restype = i
    lib.cb2_comm_fused.argtypes = [ctypes.POINTER(CommCtx), i, i, i, i, sz, sz, vp, vp, vp, sz, vp,
                                   i, i, vp, i, vp]
    lib.cb2_comm_broadcast.restype = i
    lib.cb2_comm_broadcast.argtypes = [ctypes.POINTER(CommCtx), i, i, sz, sz, i, i, vp]
    lib.cb2_comm_barrier.restype = i
    lib.cb2_comm_barrier.argtypes = [ctypes.POINTER(CommCtx), vp]
    for name, argtypes in _OPTIONAL_KERNEL_SYMBOLS.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = i
            fn.argtypes = argtypes


# entry points added after the first release are declared only if the loaded library has them
_OPTIONAL_KERNEL_SYMBOLS: dict = {}


def kernels() -> ctypes.CDLL:
    """The sm_100a kernel library (built on first use if the toolchain is present)."""
    global _kernels
    if _kernels is not None:
        return _kernels
    with _lock:
        if _kernels is not None:
            return _kernels
        path = _LIB_DIR / "libcloud_b200_kernels.so"
        if not path.exists():
            try:
                _maybe_build("kernels")
            except Exception as e:  # pragma: no cover - toolchain missing
                raise KernelsUnavailable(f"cannot build {path.name}: {e}") from e
        if not path.exists():
            raise KernelsUnavailable(
                f"{path} is missing; run `python -c 'import __graft_entry__ as g; g.build()'`")
        try:
            lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        except OSError as e:
            raise KernelsUnavailable(f"cannot load {path}: {e}") from e
        _declare_kernels(lib)
        _kernels = lib
        return lib


def runtime() -> ctypes.CDLL:
    """The C++ runtime library (arena / HDF5 / monitoring)."""
    global _runtime
    if _runtime is not None:
        return _runtime
    with _lock:
        if _runtime is not None:
            return _runtime
        path = _LIB_DIR / "libcloud_b200_rt.so"
        stale = False
        if path.exists() and not os.environ.get("CLOUD_B200_NO_AUTOBUILD"):
            src_dir = Path(__file__).resolve().parent.parent.parent / "csrc" / "runtime"
            if src_dir.exists():
                newest = max((p.stat().st_mtime_ns for p in src_dir.glob("*")), default=0)
                stale = newest > path.stat().st_mtime_ns
        if not path.exists() or stale:
            try:
                _maybe_build("runtime")
            except Exception as e:
                if not path.exists():
                    raise KernelsUnavailable(f"cannot build {path.name}: {e}") from e
        lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        _runtime = lib
        return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = kernels().cb2_last_error().decode(errors="replace")
        raise RuntimeError(f"cloud_b200 kernel call failed ({what}, rc={rc}): {msg}")
