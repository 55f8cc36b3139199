"""Race / memory checks of the C++ runtime: the unit tests and the whole
runtime library compiled into one binary with ThreadSanitizer and with AddressSanitizer.  The
thread build covers the exporter's background thread and the registry locks; the address build
covers the HDF5 reader against truncated and corrupted files (csrc/runtime/test_hdf5.cc); the
thread build also checks the release/acquire flag protocol of the shared-memory collectives, whose
ranks run as threads on one mapping there (csrc/runtime/test_shm_comm.cc)."""
import subprocess

import pytest


@pytest.mark.parametrize("sanitizer", ["thread", "address"])
def test_cpp_unit_tests_pass_under_sanitizer(sanitizer):
    from cloud_b200 import build
    try:
        bins = build.build_runtime_tests_sanitized(sanitizer)
    except RuntimeError as e:
        if "sanitize" in str(e) and ("cannot find" in str(e) or "unrecognized" in str(e)):
            pytest.skip(f"toolchain has no -fsanitize={sanitizer} runtime")
        raise
    assert {b.name.split("_" + sanitizer)[0] for b in bins} >= {"test_hdf5", "test_monitoring", "test_shm_comm"}
    for b in bins:
        res = subprocess.run([str(b)], capture_output=True, text=True, timeout=600)
        assert res.returncode == 0, (res.stdout + res.stderr)[-3000:]
        assert "PASSED" in res.stdout
        assert "WARNING: ThreadSanitizer" not in res.stderr and "ERROR: AddressSanitizer" not in res.stderr
