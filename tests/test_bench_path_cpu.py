"""The Python calls of bench.py's two arms, on the CPU at toy size: ``bench.build_model`` (the
flagship ResNet-50 classifier), the device-timed arm (``_ensure_engine`` / ``_to_internal`` /
``_run_step`` on a uint8 batch) and the end-to-end arm (``fit`` over uint8 torch tensors with a
per-batch callback reading the loss).  The numbers mean nothing here; the point is that a change
made on a GPU-less box cannot break the headline benchmark's code path without a test failing."""
import importlib.util
import os

import numpy as np
import torch

import cloud_b200 as tfc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench_module():
    spec = importlib.util.spec_from_file_location("bench_main", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_both_arms_of_the_benchmark_run_on_cpu(monkeypatch):
    bench = _bench_module()
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    torch.manual_seed(0)
    # same builder, smaller picture: patch the input size the builder hard-codes
    real_input = keras.Input
    monkeypatch.setattr(keras, "Input", lambda shape=None, **kw: real_input(shape=(32, 32, 3), **kw))
    model = bench.build_model(keras, classes=10)
    assert model.count_params() == 23608202                       # ResNet-50 trunk + 10-class head
    batch = 4
    x_dev = torch.randint(0, 256, (batch, 32, 32, 3), dtype=torch.uint8)
    y_dev = torch.randint(0, 10, (batch,))
    model._ensure_engine(x_dev)
    model.train(True)
    xi = model._to_internal(x_dev)
    assert xi.dtype == torch.uint8 and tuple(xi.shape) == (batch, 3, 32, 32)      # raw bytes reach the stem
    red = None
    for _ in range(2):
        red = model._run_step(xi, y_dev)
    loss = float(red[0]) / max(float(red[2]), 1.0)
    assert np.isfinite(loss) and float(red[2]) == batch
    # end-to-end arm
    host_x = torch.randint(0, 256, (batch * 3, 32, 32, 3), dtype=torch.uint8)
    host_y = torch.randint(0, 10, (batch * 3,), dtype=torch.int64)
    seen = []

    class ReadLoss(keras.callbacks.Callback):
        def on_train_batch_end(self, batch, logs=None):
            seen.append(logs["loss"])

    model.fit(host_x, host_y, batch_size=batch, epochs=1, shuffle=False, verbose=0, callbacks=[ReadLoss()])
    assert len(seen) == 3 and all(np.isfinite(v) for v in seen)
    assert model._h2d_bytes >= 0 and model._d2h_bytes > 0 and set(model._fit_profile) >= {"wait_data_s", "launch_s", "steps"}
    assert isinstance(model._engine.kernel_launches(), int)


def test_reference_arm_answers_without_a_gpu(capsys):
    import json
    import subprocess
    import sys
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and ("unavailable" in line or "value" in line)
