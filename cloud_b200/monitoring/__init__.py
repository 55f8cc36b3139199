"""Python binding of the C++ metrics runtime (``csrc/runtime/mon_*.cc``): registry of counters /
gauges / histograms, allow-list, periodic exporter to a local sink.

The reference's exporter is a TensorFlow plug-in nobody calls from Python
(src/cpp/monitoring/README.md:1-6); here the training engine feeds the registry directly:

    /cloud_b200/train/step_time_usecs   histogram   device time of one training step
    /cloud_b200/comm/allreduce_usecs    histogram   fused all-reduce + optimizer kernels of a step
    /cloud_b200/comm/bytes_reduced      counter     gradient bytes reduced
    /cloud_b200/comm/bucket_launches    counter     fused bucket kernels launched
    /cloud_b200/data/wait_usecs         histogram   host waited for the input pipeline
    /cloud_b200/train/images_per_sec    gauge

Environment (the reference's names are accepted too): ``TFC_MONITORING_ENABLED`` (default false;
legacy ``TF_MONITORING_STACKDRIVER_EXPORTER_ENABLED``), ``TFC_MONITORING_ALLOWLIST`` (comma list;
legacy ``TF_MONITORING_STACKDRIVER_METRICS_WHITELIST``), ``TFC_MONITORING_PROJECT_ID``,
``TFC_MONITORING_SINK`` = ``jsonl:<path>`` | ``prometheus:<path>`` | ``memory``,
``TFC_MONITORING_INTERVAL_MS`` (default 10000, the reference's 10 s).
"""
from __future__ import annotations

import ctypes

from ..ops import _lib

STEP_TIME = "/cloud_b200/train/step_time_usecs"
ALLREDUCE_TIME = "/cloud_b200/comm/allreduce_usecs"
BYTES_REDUCED = "/cloud_b200/comm/bytes_reduced"
BUCKET_LAUNCHES = "/cloud_b200/comm/bucket_launches"
DATA_WAIT = "/cloud_b200/data/wait_usecs"
IMAGES_PER_SEC = "/cloud_b200/train/images_per_sec"

SINK_JSONL, SINK_PROMETHEUS, SINK_MEMORY = 0, 1, 2


def _rt():
    lib = _lib.runtime()
    if not getattr(lib, "_mon_declared", False):
        cp, ll, d, i = ctypes.c_char_p, ctypes.c_longlong, ctypes.c_double, ctypes.c_int
        lib.cb2mon_define_counter.argtypes = [cp, cp]
        lib.cb2mon_define_gauge.argtypes = [cp, cp]
        lib.cb2mon_define_histogram.argtypes = [cp, cp, ctypes.POINTER(d), i]
        lib.cb2mon_counter_inc.argtypes = [cp, ll]
        lib.cb2mon_gauge_set.argtypes = [cp, ll]
        lib.cb2mon_observe.argtypes = [cp, d]
        lib.cb2mon_is_allowlisted.argtypes = [cp]
        lib.cb2mon_config_debug_string.argtypes = [cp, ctypes.c_size_t]
        lib.cb2mon_set_sink.argtypes = [i, cp]
        lib.cb2mon_memory_sink_dump.argtypes = [i, cp, ctypes.c_size_t]
        lib._mon_declared = True
    return lib


def define_counter(name: str, description: str = "") -> None:
    _rt().cb2mon_define_counter(name.encode(), description.encode())


def define_gauge(name: str, description: str = "") -> None:
    _rt().cb2mon_define_gauge(name.encode(), description.encode())


def define_histogram(name: str, description: str = "", bucket_limits=None) -> None:
    limits = list(bucket_limits or [])
    arr = (ctypes.c_double * max(len(limits), 1))(*limits)
    _rt().cb2mon_define_histogram(name.encode(), description.encode(), arr, len(limits))


def increment(name: str, delta: int = 1) -> None:
    _rt().cb2mon_counter_inc(name.encode(), int(delta))


def set_gauge(name: str, value: int) -> None:
    _rt().cb2mon_gauge_set(name.encode(), int(value))


def observe(name: str, value: float) -> None:
    _rt().cb2mon_observe(name.encode(), float(value))


def start() -> bool:
    """Start the periodic exporter if the environment enables it (idempotent)."""
    return bool(_rt().cb2mon_start())


def stop() -> None:
    _rt().cb2mon_stop()


def export_now() -> None:
    _rt().cb2mon_export_now()


def export_count() -> int:
    return int(_rt().cb2mon_export_count())


def reset() -> None:
    _rt().cb2mon_reset()


def is_allowlisted(name: str) -> bool:
    return bool(_rt().cb2mon_is_allowlisted(name.encode()))


def config_debug_string() -> str:
    buf = ctypes.create_string_buffer(4096)
    _rt().cb2mon_config_debug_string(buf, 4096)
    return buf.value.decode()


def set_sink(kind: int, path: str | None = None) -> None:
    _rt().cb2mon_set_sink(int(kind), None if path is None else path.encode())


def memory_sink_requests(descriptors: bool = False) -> list[str]:
    buf = ctypes.create_string_buffer(1 << 20)
    n = _rt().cb2mon_memory_sink_dump(1 if descriptors else 0, buf, 1 << 20)
    if n < 0:
        raise RuntimeError("the active sink is not the in-memory sink")
    return [l for l in buf.value.decode().split("\n") if l]


def define_defaults() -> None:
    usec = [50, 100, 200, 500, 1e3, 2e3, 5e3, 1e4, 2e4, 5e4, 1e5, 2e5, 5e5, 1e6, 1e7]
    define_histogram(STEP_TIME, "Device time of one training step (us)", usec)
    define_histogram(ALLREDUCE_TIME, "Fused all-reduce + optimizer time per step (us)", usec)
    define_histogram(DATA_WAIT, "Host wait for the input pipeline per step (us)", usec)
    define_counter(BYTES_REDUCED, "Gradient bytes reduced across replicas")
    define_counter(BUCKET_LAUNCHES, "Fused gradient-bucket kernels launched")
    define_gauge(IMAGES_PER_SEC, "Training throughput of the last epoch (samples/s)")
