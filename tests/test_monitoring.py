"""Monitoring: the C++ unit test binary (port of src/cpp/monitoring/stackdriver_client_test.cc) and
the Python binding end to end in a subprocess (env switches are read once per process)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_unit_tests_pass():
    from cloud_b200 import build
    bins = build.build_runtime_tests()
    assert bins, "no C++ test binaries were built"
    for b in bins:
        res = subprocess.run([str(b)], capture_output=True, text=True, timeout=120)
        assert res.returncode == 0, res.stdout + res.stderr
        assert "PASSED" in res.stdout


def _run(code, env):
    e = dict(os.environ, PYTHONPATH=ROOT, **env)
    return subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=120)


def test_python_binding_exports_allowlisted_nonempty_metrics(tmp_path):
    out = tmp_path / "m.jsonl"
    code = """
from cloud_b200 import monitoring as m
m.define_defaults()
assert m.start() is False                      # disabled unless the environment enables it
m.observe(m.STEP_TIME, 120.0); m.observe(m.STEP_TIME, 9000.0)
m.increment(m.BYTES_REDUCED, 1 << 20)
m.increment("/not/allowlisted", 5)
m.export_now(); m.export_now()
print(m.config_debug_string()); print(m.is_allowlisted(m.STEP_TIME), m.is_allowlisted("/not/allowlisted"))
"""
    res = _run(code, {"TFC_MONITORING_SINK": "jsonl:%s" % out, "TFC_MONITORING_PROJECT_ID": "proj"})
    assert res.returncode == 0, res.stderr
    assert "True False" in res.stdout and "allowlist: [" in res.stdout
    lines = [json.loads(l) for l in out.read_text().splitlines()]
    calls = [l["call"] for l in lines]
    assert calls.count("CreateMetricDescriptor") == 2          # once per metric
    assert calls.count("CreateTimeSeries") == 2
    ts = [l["request"] for l in lines if l["call"] == "CreateTimeSeries"][0]
    assert ts["name"] == "projects/proj"
    types = {s["metric"]["type"] for s in ts["time_series"]}
    assert types == {"custom.googleapis.com/cloud_b200/train/step_time_usecs",
                     "custom.googleapis.com/cloud_b200/comm/bytes_reduced"}
    hist = [s for s in ts["time_series"] if "step_time" in s["metric"]["type"]][0]
    dist = hist["points"]["value"]["distribution_value"]
    assert dist["count"] == 2 and abs(dist["mean"] - 4560.0) < 1e-6
    assert sum(dist["bucket_counts"]) == 2
    assert len(dist["bucket_options"]["explicit_buckets"]["bounds"]) == len(dist["bucket_counts"]) - 1


def test_periodic_exporter_with_legacy_env_names_and_prometheus_sink(tmp_path):
    out = tmp_path / "m.prom"
    code = """
import time
from cloud_b200 import monitoring as m
m.increment("/my/counter", 3)
assert m.start() is True and m.start() is True
t0 = time.time()
while m.export_count() < 2 and time.time() - t0 < 10: time.sleep(0.02)
m.stop()
print("exports", m.export_count())
"""
    res = _run(code, {"TF_MONITORING_STACKDRIVER_EXPORTER_ENABLED": "true",
                      "TF_MONITORING_STACKDRIVER_METRICS_WHITELIST": "/my/counter",
                      "TFC_MONITORING_INTERVAL_MS": "30", "TFC_MONITORING_SINK": "prometheus:%s" % out})
    assert res.returncode == 0, res.stderr
    assert "exports" in res.stdout
    assert "my_counter 3" in out.read_text()


def test_fit_feeds_the_hot_path_metrics(tmp_path):
    """The engine's own metrics (step time, data wait, throughput, bucket
    launches) reach the sink after a fit; the defaults of the allow-list are exactly these."""
    out = tmp_path / "fit.jsonl"
    code = """
import numpy as np
import cloud_b200 as tfc
from cloud_b200 import monitoring as m
tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
model = tfc.models.compiled("mnist_mlp")
model.fit(np.random.rand(256, 784).astype("float32"), np.random.randint(0, 10, 256), batch_size=64, epochs=2, verbose=0)
m.export_now()
"""
    res = _run(code, {"TFC_MONITORING_SINK": "jsonl:%s" % out, "CUDA_VISIBLE_DEVICES": ""})
    assert res.returncode == 0, res.stderr
    series = [s for l in out.read_text().splitlines() for s in json.loads(l).get("request", {}).get("time_series", [])]
    by_type = {s["metric"]["type"].split("cloud_b200/")[1]: s for s in series}
    assert {"train/step_time_usecs", "data/wait_usecs", "train/images_per_sec", "comm/bucket_launches"} <= set(by_type)
    assert by_type["data/wait_usecs"]["points"]["value"]["distribution_value"]["count"] == 2       # one per epoch
    assert int(by_type["comm/bucket_launches"]["points"]["value"]["int64_value"]) == 8             # 4 steps x 2 epochs
