"""GPU: the front-end paths added around the training engine - custom training loops
(GradientTape + apply_gradients / minimize on bf16 arenas), frozen layers next to arena-resident
ones, weight penalties - against the same maths in fp32 PyTorch on the CPU path.
(Reference usage: core/tests/testdata/mnist_example_using_ctl.py:103-131,
core/tests/examples/call_run_within_script_with_keras_fit.py:79-87,
experimental/cloud_fit/tests/unit/client_test.py:69-90.)"""
import numpy as np
import pytest
import torch

import cloud_b200 as tfc

pytestmark = pytest.mark.gpu
keras = tfc.keras
L = keras.layers


def _blobs(n, d=32, classes=4, seed=0):
    rng = np.random.RandomState(seed)
    centers = rng.randn(classes, d) * 2
    y = rng.randint(0, classes, n)
    return (centers[y] + rng.randn(n, d)).astype("float32"), y.astype("int64")


def _mlp(seed=3):
    torch.manual_seed(seed)
    return keras.Sequential([L.Dense(64, activation="relu", input_shape=(32,)), L.Dense(4)])


def _on(device):
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy(device))


def test_custom_training_loop_on_the_engine_matches_the_cpu_path():
    x, y = _blobs(256)
    loss_object = keras.losses.SparseCategoricalCrossentropy(from_logits=True,
                                                             reduction=keras.losses.Reduction.NONE)
    weights = {}
    for device in ("cuda:0", "/cpu:0"):
        _on(device)
        model, opt = _mlp(), keras.optimizers.Adam(1e-2)
        losses = []
        for step in range(8):
            xb, yb = x[step * 32:(step + 1) * 32], y[step * 32:(step + 1) * 32]
            with tfc.GradientTape() as tape:
                logits = model(xb, training=True)
                loss = tfc.nn.compute_average_loss(loss_object(yb, logits), global_batch_size=32)
            if step % 2:            # both spellings of the update
                opt.minimize(loss, model.trainable_variables, tape=tape)
            else:
                grads = tape.gradient(loss, model.trainable_variables)
                opt.apply_gradients(zip(grads, model.trainable_variables))
            losses.append(float(loss.detach()))
        weights[device] = (model.get_weights(), losses)
        if device.startswith("cuda"):
            assert all(p.is_cuda for p in model.trainable_variables)
            assert model.trainable_variables[0].dtype == torch.bfloat16     # adopted into the bf16 arena
    (wg, lg), (wc, lc) = weights["cuda:0"], weights["/cpu:0"]
    assert lg[-1] < lg[0] and abs(lg[-1] - lc[-1]) < 0.1
    for a, b in zip(wg, wc):
        assert np.abs(a - b).max() < 0.05                                  # bf16 compute vs fp32


def test_frozen_base_on_gpu_keeps_weights_and_uses_moving_statistics():
    _on("cuda:0")
    torch.manual_seed(0)
    inp = L.Input(shape=(16, 16, 8))
    h = L.Conv2D(16, 3, padding="same", name="c1")(inp)
    h = L.ReLU()(L.BatchNormalization(name="bn1")(h))
    base = keras.Model(inp, h)
    head = L.Dense(3, name="head")
    model = keras.Model(inp, head(L.GlobalAveragePooling2D()(base.output)))
    base.trainable = False
    conv, bn = model.get_layer("c1"), model.get_layer("bn1")
    rng = np.random.default_rng(0)
    xs, ys = rng.normal(size=(32, 16, 16, 8)).astype("float32"), rng.integers(0, 3, 32)
    model.predict(xs[:2])                                   # builds + places
    with torch.no_grad():
        conv.bias.fill_(0.5)
        bn.moving_mean.fill_(0.1)
        bn.moving_variance.fill_(2.0)
    model.compile(keras.optimizers.SGD(0.05), keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    fp32_before = [w.copy() for w in base.get_weights()]
    model.train_on_batch(xs[:16], ys[:16])       # the engine exists now: frozen kernels are bf16 copies
    before = [w.copy() for w in base.get_weights()]
    for a, b in zip(fp32_before, before):
        np.testing.assert_allclose(a, b, rtol=1e-2, atol=1e-3)            # only the rounding to bf16
    head_before = [w.copy() for w in head.get_weights()]
    hist = model.fit(xs, ys, batch_size=16, epochs=3, verbose=0)
    assert np.isfinite(hist.history["loss"]).all()
    assert all(np.array_equal(a, b) for a, b in zip(before, base.get_weights()))
    assert not np.array_equal(head_before[0], head.get_weights()[0])
    assert conv.kernel.dtype == torch.bfloat16 and not conv.kernel.requires_grad   # frozen, compute dtype
    ref = torch.nn.functional.conv2d(torch.tensor(xs).permute(0, 3, 1, 2), conv.kernel.float().cpu(),
                                     conv.bias.float().cpu(), padding=1)
    ref = torch.nn.functional.batch_norm(ref, bn.moving_mean.cpu(), bn.moving_variance.cpu(), bn.gamma.cpu().float(),
                                         bn.beta.cpu().float(), False, 0.0, bn.epsilon).relu().mean((2, 3))
    ref = ref @ head.kernel.float().cpu().t() + head.bias.float().cpu()
    np.testing.assert_allclose(model.predict(xs), ref.detach().numpy(), atol=0.05, rtol=0.05)


def test_weight_penalty_on_gpu():
    _on("cuda:0")
    model = tfc.models.compiled("linear_probe")
    x = np.array([[9.], [10.], [11.]] * 10, dtype="float32")
    y = x / 2 + 6
    hist = model.fit(x, y, batch_size=len(x), epochs=1, verbose=0)
    assert hist.history["loss"][0] == pytest.approx(25.0025, rel=2e-2)
    w = model.get_weights()[0][0, 0]
    assert w == pytest.approx(0.5 + 0.002 * 99.99, rel=2e-2)


def test_recompile_with_steps_per_execution_trains_the_live_weights(tmp_path):
    """compile -> fit -> compile -> fit with a K-step CUDA graph: the second fit must not replay the
    graph recorded against the first engine's arenas, and a save/load round trip
    keeps the fp32 master values, not their bf16 rounding."""
    _on("cuda:0")
    x, y = _blobs(512)
    model = _mlp()
    for _round in range(2):
        model.compile(optimizer=keras.optimizers.Adam(1e-2),
                      loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                      steps_per_execution=4)
        before = [w.copy() for w in model.get_weights()]
        h = model.fit(x, y, batch_size=32, epochs=2, shuffle=False, verbose=0)
        after = model.get_weights()
        assert all(not np.array_equal(a, b) for a, b in zip(before, after)), "weights did not move"
        assert h.history["loss"][-1] < h.history["loss"][0]
    # get_weights answers with fp32 masters: finer than bf16
    w = model.get_weights()[0]
    assert not np.array_equal(w, torch.from_numpy(w).to(torch.bfloat16).float().numpy())
    path = str(tmp_path / "w.h5")
    model.save_weights(path)
    model.set_weights([np.zeros_like(v) for v in model.get_weights()])
    model.load_weights(path)
    for a, b in zip(model.get_weights(), [w] + model.get_weights()[1:]):
        assert np.array_equal(a, b)
    assert np.array_equal(model.get_weights()[0], w)


def test_gpu_ranks_feed_the_collective_metrics(tmp_path):
    """The all-reduce time histogram, the bytes-reduced counter and the data-wait histogram reach the
    exporter's sink from a GPU rank (the reference's default metric set, stackdriver_config.cc:35-46,
    is fed by TensorFlow's runtime; here by the engine: CUDA-event pairs around the bucket kernels)."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "gpu.jsonl"
    code = '''
import numpy as np
import cloud_b200 as tfc
from cloud_b200 import monitoring as m
tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
model = tfc.models.compiled("mnist_mlp")
model.fit(np.random.rand(512, 784).astype("float32"), np.random.randint(0, 10, 512), batch_size=64, epochs=2, verbose=0)
m.export_now()
'''
    env = dict(os.environ, PYTHONPATH=root, TFC_MONITORING_SINK="jsonl:%s" % out)
    res = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-3000:]
    series = [s for l in out.read_text().splitlines() for s in json.loads(l).get("request", {}).get("time_series", [])]
    by_type = {s["metric"]["type"].split("cloud_b200/")[1]: s for s in series}
    assert {"comm/allreduce_usecs", "comm/bytes_reduced", "data/wait_usecs", "comm/bucket_launches"} <= set(by_type)
    dist = by_type["comm/allreduce_usecs"]["points"]["value"]["distribution_value"]
    assert dist["count"] >= 2 and 0.5 < dist["mean"] < 1e5          # eager steps were timed on the device
    assert int(by_type["comm/bytes_reduced"]["points"]["value"]["int64_value"]) > 0
