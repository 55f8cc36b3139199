"""Static checks of the GPU-only code paths that the CPU test run cannot execute: a refactor on a
GPU-less box must not be able to break them silently.

* every method / attribute the data-parallel engine uses on its communicator exists;
* every kernel entry point the Python layer calls (``cb2_*`` through ctypes) is exported by the
  built kernel library, and so is every function declared in ``include/cloud_b200_kernels.h``;
* the same for the C++ runtime library and ``include/cloud_b200_runtime.h``.
"""
import ast
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "cloud_b200")


def _exported(lib):
    path = os.path.join(PKG, "_lib", lib)
    if not os.path.exists(path):
        from cloud_b200 import build
        build.build_all()
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def _python_sources():
    for base, _dirs, files in os.walk(PKG):
        for f in files:
            if f.endswith(".py"):
                yield os.path.join(base, f)


def test_engine_only_uses_what_the_communicator_has():
    from cloud_b200.parallel import comm, engine
    tree = ast.parse(open(engine.__file__).read())
    used = set()
    for node in ast.walk(tree):
        # self.comm.<attr>
        if isinstance(node, ast.Attribute) and isinstance(node.value, ast.Attribute) \
                and node.value.attr == "comm" and isinstance(node.value.value, ast.Name) \
                and node.value.value.id == "self":
            used.add(node.attr)
    assert {"alloc_pair", "alloc_flat", "make_ctx", "fused", "all_reduce_sum", "broadcast"} <= used
    instance_attrs = set(re.findall(r"self\.(\w+)\s*[:=]", open(comm.__file__).read()))
    missing = {a for a in used if not hasattr(comm.Communicator, a) and a not in instance_attrs}
    assert not missing, "engine.py uses Communicator members that do not exist: {}".format(sorted(missing))
    module_used = {n.attr for n in ast.walk(tree) if isinstance(n, ast.Attribute)
                   and isinstance(n.value, ast.Name) and n.value.id == "comm_mod"}
    assert not {a for a in module_used if not hasattr(comm, a)}


def test_front_end_only_uses_what_the_engine_has():
    from cloud_b200.parallel import engine
    members = set(dir(engine.DPEngine)) | set(re.findall(r"self\.(\w+)\s*[:=]", open(engine.__file__).read()))
    for rel in ("keras/models.py", "keras/optimizers.py", "keras/saving.py", "keras/tape.py"):
        src = open(os.path.join(PKG, rel)).read()
        for attr in set(re.findall(r"\b(?:eng|self\._engine|model\._engine)\.(\w+)", src)):
            assert attr in members, "{} uses DPEngine.{} which does not exist".format(rel, attr)


def test_every_kernel_entry_point_python_calls_is_exported():
    exported = _exported("libcloud_b200_kernels.so")
    called = {}
    for path in _python_sources():
        # attribute access on the ctypes library handle: `lib.cb2_xyz` / `_k().cb2_xyz(...)`
        for name in re.findall(r"\.(cb2_[a-z0-9_]*[a-z0-9])\b", open(path).read()):
            called.setdefault(name, path)
    assert len(called) >= 25
    missing = {n: p for n, p in called.items() if n not in exported}
    assert not missing, "not exported by the kernel library: {}".format(missing)


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return set(re.findall(r"\b(cb2[a-z0-9]*_[a-z0-9_]+)\s*\(", text))


@pytest.mark.parametrize("header,lib", [("cloud_b200_kernels.h", "libcloud_b200_kernels.so"),
                                        ("cloud_b200_runtime.h", "libcloud_b200_rt.so")])
def test_every_declared_entry_point_is_defined(header, lib):
    declared = _declared(header)
    assert len(declared) > 20
    missing = declared - _exported(lib)
    assert not missing, "{} declares functions {} does not define: {}".format(header, lib, sorted(missing))


def test_runtime_entry_points_python_calls_are_exported():
    exported = _exported("libcloud_b200_rt.so")
    called = set()
    for path in _python_sources():
        called |= set(re.findall(r"\.(cb2(?:rt|h5|mon)_[a-z0-9_]*[a-z0-9])\b", open(path).read()))
    assert len(called) > 20
    assert not (called - exported), sorted(called - exported)


def test_every_module_of_the_package_imports():
    """Import errors in modules that only a GPU run touches (ops/*, parallel/symm.py, bench helpers)
    show up here."""
    import importlib
    import pkgutil
    import cloud_b200
    failed = {}
    for info in pkgutil.walk_packages(cloud_b200.__path__, "cloud_b200."):
        if info.name.endswith("__main__"):
            continue
        try:
            importlib.import_module(info.name)
        except Exception as e:  # noqa: BLE001
            failed[info.name] = repr(e)
    assert not failed, failed


def test_bench_and_tool_scripts_compile():
    import py_compile
    for folder in ("bench", "tools", "examples"):
        for name in sorted(os.listdir(os.path.join(ROOT, folder))):
            if name.endswith(".py"):
                py_compile.compile(os.path.join(ROOT, folder, name), doraise=True)
    py_compile.compile(os.path.join(ROOT, "bench.py"), doraise=True)
    py_compile.compile(os.path.join(ROOT, "__graft_entry__.py"), doraise=True)


def test_no_name_is_used_that_is_defined_nowhere_in_its_module():
    """A coarse undefined-name check (every name loaded in a module is imported, assigned, or a
    parameter somewhere in that module): typos in branches only a GPU run reaches fail here."""
    import builtins
    known = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__path__"}
    files = list(_python_sources()) + [os.path.join(ROOT, "bench.py"), os.path.join(ROOT, "__graft_entry__.py")]
    files += [os.path.join(ROOT, "bench", f) for f in os.listdir(os.path.join(ROOT, "bench")) if f.endswith(".py")]
    problems = {}
    for path in files:
        tree = ast.parse(open(path).read())
        defined = set(known)
        for n in ast.walk(tree):
            if isinstance(n, (ast.FunctionDef, ast.ClassDef, ast.AsyncFunctionDef)):
                defined.add(n.name)
            elif isinstance(n, ast.Import):
                defined.update((a.asname or a.name).split(".")[0] for a in n.names)
            elif isinstance(n, ast.ImportFrom):
                defined.update(a.asname or a.name for a in n.names)
            elif isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
                defined.add(n.id)
            elif isinstance(n, ast.arg):
                defined.add(n.arg)
            elif isinstance(n, ast.ExceptHandler) and n.name:
                defined.add(n.name)
            elif isinstance(n, (ast.Global, ast.Nonlocal)):
                defined.update(n.names)
        bad = sorted({(n.id, n.lineno) for n in ast.walk(tree)
                      if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load) and n.id not in defined})
        if bad:
            problems[os.path.relpath(path, ROOT)] = bad
    assert not problems, problems
