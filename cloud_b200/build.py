"""In-tree native build for cloud_b200.

Two shared libraries are produced under ``cloud_b200/_lib``:

* ``libcloud_b200_kernels.so`` - every CUDA kernel (``csrc/kernels/*.cu``), compiled by nvcc for
  ``sm_100a`` only (``-gencode arch=compute_100a,code=sm_100a -lineinfo``).  nvcc cross-compiles,
  so this works on a box without a GPU.
* ``libcloud_b200_rt.so`` - the C++ runtime (``csrc/runtime/*.cc``): symmetric-memory arena
  (CUDA VMM / multicast through a dlopen'ed driver), HDF5 writer/reader, metrics registry and
  exporter, job launcher helpers.

The reference builds its only native piece with Bazel (src/cpp/monitoring/BUILD:25-92); Bazel is
not available here, so this is a small dependency-free driver around nvcc / g++ with mtime caching
and parallel compilation.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = Path(__file__).resolve().parent / "_lib"
OBJ_DIR = ROOT / "build" / "obj"

KERNEL_LIB = LIB_DIR / "libcloud_b200_kernels.so"
RUNTIME_LIB = LIB_DIR / "libcloud_b200_rt.so"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
              "-Xcompiler", "-fvisibility=default", "--expt-relaxed-constexpr"]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-Wno-unused-function", "-pthread"]


def _find(tool: str, extra: list[str]) -> str | None:
    path = shutil.which(tool)
    if path:
        return path
    for cand in extra:
        if os.path.exists(cand):
            return cand
    return None


def nvcc_path() -> str | None:
    return _find("nvcc", ["/usr/local/cuda/bin/nvcc"])


def cuda_home() -> str:
    nv = nvcc_path()
    if nv:
        return str(Path(nv).resolve().parent.parent)
    return os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _stamp(src: Path, flags: list[str]) -> str:
    h = hashlib.sha1()
    h.update(" ".join(flags).encode())
    deps = [src] + sorted(src.parent.glob("*.cuh")) + sorted(src.parent.glob("*.h")) + \
        sorted(INCLUDE.glob("*.h"))
    for d in deps:
        st = d.stat()
        h.update(f"{d}:{st.st_mtime_ns}:{st.st_size}".encode())
    return h.hexdigest()


def _compile_one(cmd: list[str], src: Path, obj: Path, flags: list[str], verbose: bool) -> Path:
    stamp_file = obj.with_suffix(obj.suffix + ".stamp")
    stamp = _stamp(src, flags)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    obj.parent.mkdir(parents=True, exist_ok=True)
    full = cmd + flags + ["-c", str(src), "-o", str(obj)]
    if verbose:
        print("[cloud_b200.build]", " ".join(full), flush=True)
    res = subprocess.run(full, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"compile failed: {src}\n{res.stdout}\n{res.stderr}")
    if verbose and res.stderr.strip():
        print(res.stderr, file=sys.stderr)
    stamp_file.write_text(stamp)
    return obj


def build_kernels(verbose: bool = False, ptxas_verbose: bool = False) -> Path:
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build the sm_100a kernel library")
    srcs = sorted((CSRC / "kernels").glob("*.cu"))
    flags = NVCC_ARCH + NVCC_FLAGS + [f"-I{INCLUDE}", f"-I{cuda_home()}/include"]
    if ptxas_verbose:
        flags = flags + ["-Xptxas", "-v"]
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        objs = list(ex.map(
            lambda s: _compile_one([nvcc], s, OBJ_DIR / "kernels" / (s.stem + ".o"), flags,
                                   verbose), srcs))
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime_ns for o in objs)
    if not KERNEL_LIB.exists() or KERNEL_LIB.stat().st_mtime_ns < newest:
        cmd = [nvcc, "-shared", "-o", str(KERNEL_LIB)] + [str(o) for o in objs] + \
            NVCC_ARCH + ["-lcudart"]
        if verbose:
            print("[cloud_b200.build]", " ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return KERNEL_LIB


def build_runtime(verbose: bool = False) -> Path:
    cxx = _find("g++", ["/usr/bin/g++"])
    if cxx is None:
        raise RuntimeError("g++ not found; cannot build the runtime library")
    srcs = sorted((CSRC / "runtime").glob("*.cc"))
    srcs = [s for s in srcs if not s.name.startswith("test_")]
    flags = CXX_FLAGS + [f"-I{INCLUDE}", f"-I{cuda_home()}/include"]
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        objs = list(ex.map(
            lambda s: _compile_one([cxx], s, OBJ_DIR / "runtime" / (s.stem + ".o"), flags,
                                   verbose), srcs))
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    newest = max(o.stat().st_mtime_ns for o in objs)
    if not RUNTIME_LIB.exists() or RUNTIME_LIB.stat().st_mtime_ns < newest:
        cmd = [cxx, "-shared", "-o", str(RUNTIME_LIB)] + [str(o) for o in objs] + \
            ["-ldl", "-pthread"]
        if verbose:
            print("[cloud_b200.build]", " ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return RUNTIME_LIB


def build_runtime_tests(verbose: bool = False) -> list[Path]:
    """Build the stand-alone C++ unit tests (csrc/runtime/test_*.cc) against the runtime lib."""
    cxx = _find("g++", ["/usr/bin/g++"])
    build_runtime(verbose)
    outs = []
    for src in sorted((CSRC / "runtime").glob("test_*.cc")):
        out = ROOT / "build" / "bin" / src.stem
        out.parent.mkdir(parents=True, exist_ok=True)
        if out.exists() and out.stat().st_mtime_ns > max(
                src.stat().st_mtime_ns, RUNTIME_LIB.stat().st_mtime_ns):
            outs.append(out)
            continue
        cmd = [cxx] + CXX_FLAGS + [f"-I{INCLUDE}", f"-I{CSRC / 'runtime'}", str(src), "-o", str(out),
                                   f"-L{LIB_DIR}", "-lcloud_b200_rt", f"-Wl,-rpath,{LIB_DIR}",
                                   "-ldl", "-pthread"]
        if verbose:
            print("[cloud_b200.build]", " ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"test build failed: {src}\n{res.stdout}\n{res.stderr}")
        outs.append(out)
    return outs


def build_runtime_tests_sanitized(sanitizer: str = "thread", verbose: bool = False) -> list[Path]:
    """The C++ unit tests and the whole runtime library compiled into one binary with
    ``-fsanitize=<sanitizer>`` (thread | address | undefined): the race / memory checks of
    SURVEY.md section 5.2 for the exporter thread, the registry locks and the HDF5 code."""
    cxx = _find("g++", ["/usr/bin/g++"])
    if cxx is None:
        raise RuntimeError("g++ not found")
    lib_srcs = [s for s in sorted((CSRC / "runtime").glob("*.cc")) if not s.name.startswith("test_")]
    outs = []
    for src in sorted((CSRC / "runtime").glob("test_*.cc")):
        out = ROOT / "build" / "bin" / f"{src.stem}_{sanitizer}san"
        out.parent.mkdir(parents=True, exist_ok=True)
        newest = max(p.stat().st_mtime_ns for p in lib_srcs + [src])
        if out.exists() and out.stat().st_mtime_ns > newest:
            outs.append(out)
            continue
        flags = [f for f in CXX_FLAGS if f not in ("-O3", "-O2")] + ["-O1", "-g", f"-fsanitize={sanitizer}",
                                                                      "-fno-omit-frame-pointer"]
        cmd = [cxx] + flags + [f"-I{INCLUDE}", f"-I{CSRC / 'runtime'}", f"-I{cuda_home()}/include", str(src)] + \
            [str(s) for s in lib_srcs] + ["-o", str(out), "-ldl", "-pthread"]
        if verbose:
            print("[cloud_b200.build]", " ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"sanitized test build failed: {src}\n{res.stdout}\n{res.stderr}")
        outs.append(out)
    return outs


def build_all(verbose: bool = False) -> dict[str, Path]:
    return {"kernels": build_kernels(verbose), "runtime": build_runtime(verbose)}


if __name__ == "__main__":
    out = build_all(verbose="-q" not in sys.argv)
    for k, v in out.items():
        print(f"{k}: {v}")
