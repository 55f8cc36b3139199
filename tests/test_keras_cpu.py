"""Keras-shaped front end on the CPU path: training loop, functional graphs, callbacks, optimizers
against torch.optim, HDF5 weight round trips.  (New tier: the reference leaves all of this to
tf.keras - SURVEY.md appendix B.)"""
import json
import os

import numpy as np
import pytest
import torch

import cloud_b200 as tfc
from cloud_b200.io import hdf5

keras = tfc.keras
L = keras.layers


@pytest.fixture(autouse=True)
def _cpu_strategy():
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    torch.manual_seed(0)
    np.random.seed(0)
    yield


def _blobs(n=512, d=20, classes=4, seed=0):
    rng = np.random.RandomState(seed)
    centers = rng.randn(classes, d) * 3
    y = rng.randint(0, classes, n)
    x = centers[y] + rng.randn(n, d)
    return x.astype("float32"), y.astype("int64")


def _mlp(d=20, classes=4, dropout=0.1):
    m = keras.Sequential([L.Dense(32, activation="relu", input_shape=(d,)), L.Dropout(dropout),
                          L.Dense(classes, activation="softmax")])
    m.compile(optimizer=keras.optimizers.Adam(1e-2), loss="sparse_categorical_crossentropy",
              metrics=["accuracy"])
    return m


def test_sequential_fit_learns_and_reports_history():
    x, y = _blobs()
    m = _mlp()
    h = m.fit(x, y, batch_size=64, epochs=6, verbose=0, shuffle=True)
    assert set(h.history) >= {"loss", "accuracy"}
    assert len(h.history["loss"]) == 6
    assert h.history["loss"][-1] < 0.5 * h.history["loss"][0]
    assert h.history["accuracy"][-1] > 0.9
    ev = m.evaluate(x, y, batch_size=128, verbose=0, return_dict=True)
    assert ev["accuracy"] > 0.9 and abs(ev["loss"] - h.history["loss"][-1]) < 0.5
    p = m.predict(x[:10])
    assert p.shape == (10, 4) and np.allclose(p.sum(1), 1.0, atol=1e-3)


def test_functional_conv_graph_with_residual_and_pooling():
    inp = keras.Input(shape=(16, 16, 3))
    x = L.Conv2D(8, 3, padding="same", name="c1")(inp)
    x = L.BatchNormalization(activation="relu", name="bn1")(x)
    s = x
    x = L.Conv2D(8, 3, padding="same", use_bias=False, name="c2")(x)
    x = L.BatchNormalization(name="bn2")([x, s])                      # fused shortcut input
    x = L.MaxPooling2D(2)(x)
    x = L.Concatenate()([L.GlobalAveragePooling2D()(x), L.GlobalMaxPooling2D()(x)])
    out = L.Dense(3, activation="softmax", name="head")(x)
    m = keras.Model(inp, out)
    assert m.get_layer("c1").bias_absorbed_by_bn and m.get_layer("c2").feeds_bn
    m.compile(optimizer="sgd", loss="sparse_categorical_crossentropy", metrics=["acc"])
    xs = np.random.rand(24, 16, 16, 3).astype("float32")
    ys = np.random.randint(0, 3, 24)
    before = m.evaluate(xs, ys, verbose=0, return_dict=True)["loss"]
    m.fit(xs, ys, batch_size=8, epochs=5, verbose=0)
    after = m.evaluate(xs, ys, verbose=0, return_dict=True)["loss"]
    assert after < before
    assert m.output_shape == (None, 3) and m.predict(xs[:2]).shape == (2, 3)
    cfg = m.to_json()
    m2 = keras.model_from_json(cfg)
    assert [l.name for l in m2.layers] == [l.name for l in m.layers]


def test_callbacks_early_stopping_checkpoint_and_lr_schedule(tmp_path):
    x, y = _blobs(256)
    m = _mlp()
    seen = []
    ckpt = str(tmp_path / "w-{epoch:02d}.h5")
    cbs = [keras.callbacks.EarlyStopping(monitor="loss", patience=0, min_delta=10.0),   # stops after epoch 2
           keras.callbacks.ModelCheckpoint(ckpt, save_weights_only=True),
           keras.callbacks.LearningRateScheduler(lambda epoch, lr: lr * 0.5),
           keras.callbacks.LambdaCallback(on_epoch_end=lambda e, logs: seen.append((e, logs["loss"])))]
    h = m.fit(x, y, batch_size=64, epochs=10, verbose=0, callbacks=cbs)
    assert len(h.history["loss"]) == 2 and [e for e, _ in seen] == [0, 1]
    assert os.path.exists(str(tmp_path / "w-01.h5")) and os.path.exists(str(tmp_path / "w-02.h5"))
    assert abs(m.optimizer.current_lr() - 1e-2 * 0.25) < 1e-9


@pytest.mark.parametrize("name,make,ref", [
    ("adam", lambda: keras.optimizers.Adam(1e-2), lambda p: torch.optim.Adam(p, 1e-2, eps=1e-7)),
    ("sgd_momentum", lambda: keras.optimizers.SGD(0.05, momentum=0.9),
     lambda p: torch.optim.SGD(p, 0.05, momentum=0.9)),
    ("rmsprop", lambda: keras.optimizers.RMSprop(1e-2), None),
    ("adamw", lambda: keras.optimizers.AdamW(1e-2, weight_decay=0.1),
     lambda p: torch.optim.AdamW(p, 1e-2, eps=1e-7, weight_decay=0.1)),
])
def test_optimizers_follow_their_reference_update_rule(name, make, ref):
    x, y = _blobs(128, d=6, classes=2)
    m = keras.Sequential([L.Dense(2, input_shape=(6,), name="d")])
    m.compile(optimizer=make(), loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    w0 = [w.copy() for w in m.get_weights()]
    for _ in range(5):
        m.train_on_batch(x, y)
    w1 = m.get_weights()
    assert not np.allclose(w0[0], w1[0])
    if ref is None:
        return
    k = torch.tensor(w0[0], requires_grad=True)
    b = torch.tensor(w0[1], requires_grad=True)
    opt = ref([k, b])
    xt, yt = torch.tensor(x), torch.tensor(y)
    for _ in range(5):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(xt @ k + b, yt).backward()
        opt.step()
    assert np.allclose(w1[0], k.detach().numpy(), atol=2e-4), name
    assert np.allclose(w1[1], b.detach().numpy(), atol=2e-4), name


def test_weights_round_trip_in_keras_hdf5_layout(tmp_path):
    x, y = _blobs(64)
    m = _mlp()
    m.fit(x, y, batch_size=32, epochs=1, verbose=0)
    path = str(tmp_path / "weights.h5")
    m.save_weights(path)
    with hdf5.File(path, "r") as f:
        names = [n.decode() if isinstance(n, bytes) else n for n in f.attrs["layer_names"]]
        assert names == [l.name for l in m.layers]
        dense = names[0]
        wnames = [n.decode() if isinstance(n, bytes) else n for n in f[dense].attrs["weight_names"]]
        assert wnames == [f"{dense}/kernel:0", f"{dense}/bias:0"]
        assert f[dense][wnames[0]].shape == (20, 32)            # Keras Dense layout [in, out]
    m2 = _mlp()
    m2.build((None, 20)) if hasattr(m2, "build") else None
    m2.predict(x[:1])
    m2.load_weights(path)
    for a, b in zip(m.get_weights(), m2.get_weights()):
        assert np.array_equal(a, b)
    assert np.allclose(m.predict(x[:8]), m2.predict(x[:8]), atol=1e-6)


def test_whole_model_save_and_load(tmp_path):
    x, y = _blobs(64)
    m = _mlp(dropout=0.0)          # no dropout: the two copies must take identical steps
    m.fit(x, y, batch_size=32, epochs=1, verbose=0)
    d = str(tmp_path / "saved")
    m.save(d)
    assert os.path.exists(os.path.join(d, "config.json"))
    json.load(open(os.path.join(d, "config.json")))
    m2 = keras.load_model(d)
    assert np.allclose(m.predict(x[:8]), m2.predict(x[:8]), atol=1e-6)
    # the optimizer state came along: one more identical step gives identical weights
    m.train_on_batch(x[:32], y[:32])
    m2.train_on_batch(x[:32], y[:32])
    for a, b in zip(m.get_weights(), m2.get_weights()):
        assert np.allclose(a, b, atol=1e-6)


def test_fit_on_dataset_and_validation():
    x, y = _blobs(256)
    ds = tfc.data.Dataset.from_tensor_slices((x, y)).shuffle(256, seed=1).batch(32)
    m = _mlp()
    h = m.fit(ds, epochs=3, verbose=0, validation_data=(x[:64], y[:64]))
    assert "val_loss" in h.history and "val_accuracy" in h.history
    assert h.history["val_accuracy"][-1] > 0.8


def test_image_normalization_defers_only_to_a_stem_convolution():
    inp = keras.Input(shape=(32, 32, 3))
    n = L.ImageNormalization(name="n")(inp)
    out = L.GlobalAveragePooling2D()(L.Conv2D(64, 7, strides=2, padding=3, name="stem")(n))
    m = keras.Model(inp, out)
    assert m.get_layer("n").defer_to_stem
    inp2 = keras.Input(shape=(32, 32, 3))
    n2 = L.ImageNormalization(name="n2")(inp2)
    out2 = L.GlobalAveragePooling2D()(L.Conv2D(8, 3, name="c")(n2))
    m2 = keras.Model(inp2, out2)
    assert not m2.get_layer("n2").defer_to_stem
    img = np.random.randint(0, 256, (2, 32, 32, 3)).astype("uint8")
    assert m.predict(img).shape == (2, 64) and m2.predict(img).shape == (2, 8)    # CPU: never deferred


def test_steps_per_execution_groups_steps_without_changing_the_training():
    """compile(steps_per_execution=K): K steps per call (one CUDA graph on a GPU); same result."""
    x, y = _blobs(10 * 32 + 7)          # 11 batches of 32, the last one short
    results = {}
    for spe in (1, 4):
        torch.manual_seed(0)
        m = keras.Sequential([L.Dense(16, activation="relu", input_shape=(20,)),
                              L.Dense(4, activation="softmax")])
        m.compile(optimizer=keras.optimizers.SGD(0.05), loss="sparse_categorical_crossentropy",
                  metrics=["accuracy"], steps_per_execution=spe)
        ends = []
        h = m.fit(x, y, batch_size=32, epochs=2, shuffle=False, verbose=0,
                  callbacks=[keras.callbacks.LambdaCallback(on_batch_end=lambda b, logs: ends.append(b))])
        results[spe] = (m.get_weights(), h.history, ends)
        assert m.optimizer.iterations == 22
    w1, h1, e1 = results[1]
    w4, h4, e4 = results[4]
    for a, b in zip(w1, w4):
        assert np.allclose(a, b, atol=1e-6)
    assert np.allclose(h1["loss"], h4["loss"], rtol=1e-5) and np.allclose(h1["accuracy"], h4["accuracy"])
    assert e1[:11] == list(range(11))
    assert e4[:5] == [3, 7, 8, 9, 10]      # two full groups, then the leftovers one by one


@pytest.mark.parametrize("shuffle", [False, True])
def test_array_source_groups_partition_each_global_batch_across_replicas(shuffle):
    from cloud_b200.keras.batch_sources import _ArraySource
    n, B, K, world = 100, 8, 3, 2
    x = torch.arange(n).float().view(n, 1)
    y = torch.arange(n)
    plain = list(_ArraySource(x, y, B, shuffle, 1, 0, seed=5))                 # the global batches
    per_rank = [list(_ArraySource(x, y, B, shuffle, world, r, seed=5, group=K)) for r in range(world)]
    assert len(per_rank[0]) == len(per_rank[1])
    step = 0
    for items in zip(*per_rank):
        k = getattr(items[0][2], "k", 1)
        for i in range(k):
            rows = torch.cat([(it[1][i] if k > 1 else it[1]) for it in items])
            want = plain[step][1]
            assert sorted(rows.tolist()) == sorted(want.tolist())
            step += 1
        if k > 1:
            assert items[0][0].shape == (K, B // world, 1) and int(items[0][2]) == B
    assert step == len(plain) == 13


def test_tensorboard_jsonl_progbar_and_terminate_on_nan(tmp_path, capsys):
    x, y = _blobs(128)
    m = _mlp(dropout=0.0)
    tb = keras.callbacks.TensorBoard(log_dir=str(tmp_path / "logs"))
    m.fit(x, y, batch_size=32, epochs=2, verbose=2, validation_data=(x[:32], y[:32]), callbacks=[tb])
    out = capsys.readouterr().out
    assert "Epoch 1/2" in out and "Epoch 2/2" in out and "loss:" in out and "val_loss:" in out
    recs = [json.loads(l) for l in open(tmp_path / "logs" / "train" / "scalars.jsonl")]
    assert [r["step"] for r in recs] == [0, 1] and {"loss", "accuracy"} <= set(recs[0])
    vrecs = [json.loads(l) for l in open(tmp_path / "logs" / "validation" / "scalars.jsonl")]
    assert len(vrecs) == 2 and "loss" in vrecs[0]
    # a diverging run stops at the first non-finite loss
    m2 = keras.Sequential([L.Dense(4, input_shape=(20,))])
    m2.compile(optimizer="sgd", loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    x_bad = x.copy()
    x_bad[5, 3] = np.nan
    h = m2.fit(x_bad, y, batch_size=32, epochs=5, verbose=0, shuffle=False,
               callbacks=[keras.callbacks.TerminateOnNaN()])
    assert len(h.history["loss"]) == 1 and m2.stop_training


def test_learning_rate_schedules_and_metrics_objects():
    sched = keras.optimizers.ExponentialDecay(0.1, decay_steps=10, decay_rate=0.5, staircase=True)
    assert sched(0) == pytest.approx(0.1) and sched(10) == pytest.approx(0.05) and sched(25) == pytest.approx(0.025)
    pw = keras.optimizers.PiecewiseConstantDecay([5, 10], [1.0, 0.5, 0.1])
    assert [pw(s) for s in (0, 5, 6, 11)] == [1.0, 1.0, 0.5, 0.1]
    opt = keras.optimizers.SGD(learning_rate=sched)
    x, y = _blobs(64)
    m = keras.Sequential([L.Dense(4, activation="softmax", input_shape=(20,))])
    m.compile(optimizer=opt, loss="sparse_categorical_crossentropy")
    m.fit(x, y, batch_size=8, epochs=2, verbose=0)             # 16 steps: the rate has halved once
    assert m.optimizer.current_lr() == pytest.approx(0.05)
    acc = keras.metrics.SparseCategoricalAccuracy()
    acc.update_state(np.array([0, 1, 2]), np.array([[0.9, 0.1, 0], [0.2, 0.7, 0.1], [0.5, 0.4, 0.1]]))
    assert float(acc.result()) == pytest.approx(2 / 3)
    mean = keras.metrics.Mean()
    mean.update_state(2.0)
    mean.update_state(4.0)
    assert float(mean.result()) == pytest.approx(3.0)
    mean.reset_states() if hasattr(mean, "reset_states") else mean.reset_state()


# -- regularizers, initializers, frozen layers (the cloud_fit test model of the reference uses all of
# them: experimental/cloud_fit/tests/unit/client_test.py:69-90) ------------------------------------
def test_weight_penalty_is_part_of_the_loss_and_of_the_gradient():
    inp = L.Input(shape=(1,))
    times_w = L.Dense(1, kernel_initializer=keras.initializers.Constant([[0.5]]),
                      kernel_regularizer=keras.regularizers.l2(0.01), use_bias=False)
    plus_1 = L.Dense(1, kernel_initializer=keras.initializers.Constant([[1.0]]),
                     bias_initializer=keras.initializers.Constant([1.0]), trainable=False)
    model = keras.Model(inp, plus_1(times_w(inp)))
    model.compile(keras.optimizers.SGD(0.002), "mean_squared_error", run_eagerly=True)
    x = np.array([[9.], [10.], [11.]] * 10, dtype="float32")
    y = x / 2 + 6
    hist = model.fit(x, y, batch_size=len(x), epochs=1, verbose=0)
    # loss = mean((0.5 x + 1 - y)^2) + 0.01 * 0.5^2 ; dloss/dw = mean(2 (0.5 x + 1 - y) x) + 0.02 * 0.5
    assert hist.history["loss"][0] == pytest.approx(25.0 + 0.0025, rel=1e-6)
    w, frozen_w, frozen_b = model.get_weights()
    assert w[0, 0] == pytest.approx(0.5 + 0.002 * (100.0 - 0.01), rel=1e-6)
    assert frozen_w[0, 0] == 1.0 and frozen_b[0] == 1.0
    assert len(model.losses) == 1 and float(model.losses[0].detach()) == pytest.approx(0.01 * w[0, 0] ** 2, rel=1e-5)
    assert model.evaluate(x, y, verbose=0) == pytest.approx(
        float(np.mean((w[0, 0] * x + 1 - y) ** 2)) + 0.01 * w[0, 0] ** 2, rel=1e-5)
    # the penalty survives a save / load round trip through the layer config
    clone = keras.Model.from_config(model.get_config())
    assert clone.layers[0].kernel_regularizer.l2 == pytest.approx(0.01)
    assert keras.regularizers.get("l1_l2").l1 == 0.01 and keras.regularizers.get(None) is None
    assert float(keras.regularizers.l1_l2(0.5, 0.25)(torch.tensor([1.0, -2.0]))) == pytest.approx(0.5 * 3 + 0.25 * 5)


def test_initializers_have_keras_statistics_and_configs():
    init = keras.initializers
    t = init.get("glorot_uniform").fill(torch.empty(400, 600), 400, 600)
    limit = (6.0 / 1000) ** 0.5
    assert t.abs().max() <= limit and t.std() == pytest.approx(limit / 3 ** 0.5, rel=0.05)
    t = init.HeNormal(seed=1).fill(torch.empty(512, 512), 512, 512)
    assert t.std() == pytest.approx((2.0 / 512) ** 0.5, rel=0.05)
    assert torch.equal(t, init.HeNormal(seed=1).fill(torch.empty(512, 512), 512, 512))     # seeded
    assert init.Constant([[1.0, 2.0]]).fill(torch.empty(1, 2), 1, 2).tolist() == [[1.0, 2.0]]
    assert init.get("zeros").fill(torch.empty(3), 1, 1).sum() == 0
    cfg = init.serialize(init.RandomNormal(stddev=0.2, seed=3))
    assert init.get(cfg).stddev == 0.2 and cfg["class_name"] == "RandomNormal"
    with pytest.raises(ValueError, match="Unknown initializer"):
        init.get("no_such_init")
    dense = L.Dense(4, kernel_initializer=init.RandomUniform(-0.01, 0.01), bias_initializer="ones")
    dense(torch.zeros(2, 8))
    assert dense.kernel.abs().max() <= 0.01 and dense.bias.tolist() == [1.0] * 4


def test_frozen_base_keeps_its_weights_and_runs_batch_norm_on_moving_statistics():
    inp = L.Input(shape=(8, 8, 3))
    x = L.Conv2D(4, 3, padding="same", name="c1")(inp)          # its bias is absorbed while BN trains
    x = L.ReLU()(L.BatchNormalization(name="bn1")(x))
    base = keras.Model(inp, x)
    head = L.Dense(3, name="head")
    model = keras.Model(inp, head(L.GlobalAveragePooling2D()(base.output)))
    base.trainable = False
    conv, bn = model.get_layer("c1"), model.get_layer("bn1")
    assert not base.trainable and not conv.trainable and conv.bias_absorbed_by_bn is False
    with torch.no_grad():
        conv.bias.fill_(0.7)
        bn.moving_mean.fill_(0.1)
        bn.moving_variance.fill_(2.0)
    model.compile(keras.optimizers.SGD(0.1), keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    assert len(model.trainable_weights) == 2
    rng = np.random.default_rng(0)
    xs, ys = rng.normal(size=(16, 8, 8, 3)).astype("float32"), rng.integers(0, 3, 16)
    before = [w.copy() for w in base.get_weights()]
    head_before = None
    first = model.train_on_batch(xs, ys)
    head_before = [w.copy() for w in head.get_weights()]
    model.fit(xs, ys, batch_size=8, epochs=2, verbose=0)
    assert all(np.array_equal(a, b) for a, b in zip(before, base.get_weights()))      # incl. moving stats
    assert not np.array_equal(head_before[0], head.get_weights()[0])
    ref = torch.nn.functional.conv2d(torch.tensor(xs).permute(0, 3, 1, 2), conv.kernel.float(),
                                     conv.bias.float(), padding=1)
    ref = torch.nn.functional.batch_norm(ref, bn.moving_mean, bn.moving_variance, bn.gamma, bn.beta,
                                         False, 0.0, bn.epsilon).relu().mean((2, 3))
    ref = ref @ head.kernel.float().t() + head.bias
    np.testing.assert_allclose(model.predict(xs), ref.detach().numpy(), atol=1e-5)
    assert np.isfinite(first)
    base.trainable = True
    assert conv.bias_absorbed_by_bn is True and bn.gamma.requires_grad


def test_optimizer_learning_rate_reads_like_a_variable():
    opt = keras.optimizers.Adam(learning_rate=3e-4)
    assert opt.lr.numpy() == pytest.approx(3e-4) and opt.learning_rate == pytest.approx(3e-4)
    opt.lr = 1e-2
    assert opt.learning_rate.numpy() == pytest.approx(1e-2) and opt.current_lr() == pytest.approx(1e-2)
    opt.learning_rate.assign(5e-3)
    assert float(opt.lr) == pytest.approx(5e-3)
    assert json.loads(json.dumps(opt.get_config()))["learning_rate"] == pytest.approx(5e-3)


def test_layer_configs_round_trip():
    cases = [L.Rescaling(1 / 255.0, offset=-1), L.RandomFlip("horizontal"), L.RandomRotation(0.1),
             L.Concatenate(axis=1), L.Reshape((2, 2)), L.ZeroPadding2D(((1, 2), (3, 4))),
             L.Dense(4, activation="relu", kernel_regularizer=keras.regularizers.l2(0.5))]
    for layer in cases:
        cfg = layer.get_config()
        clone = type(layer).from_config(dict(cfg))
        assert clone.get_config() == cfg, type(layer).__name__


# -- custom training loops (reference: core/tests/testdata/mnist_example_using_ctl.py:103-131) -----
def _ctl_model():
    torch.manual_seed(3)
    return keras.Sequential([L.Dense(16, activation="relu", input_shape=(20,)), L.Dense(4)])


def test_gradient_tape_and_apply_gradients_match_fit():
    x, y = _blobs(128)
    loss_object = keras.losses.SparseCategoricalCrossentropy(from_logits=True,
                                                             reduction=keras.losses.Reduction.NONE)
    ctl, opt = _ctl_model(), keras.optimizers.Adam(1e-2)
    for step in range(6):
        xb, yb = x[step * 16:(step + 1) * 16], y[step * 16:(step + 1) * 16]
        with tfc.GradientTape() as tape:
            logits = ctl(xb, training=True)
            loss = tfc.nn.compute_average_loss(loss_object(yb, logits), global_batch_size=16)
        grads = tape.gradient(loss, ctl.trainable_variables)
        assert len(grads) == 4 and all(g is not None for g in grads)
        opt.apply_gradients(zip(grads, ctl.trainable_variables))
    assert opt.iterations == 6
    fitted = _ctl_model()
    fitted.compile(keras.optimizers.Adam(1e-2), keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    fitted.fit(x[:96], y[:96], batch_size=16, epochs=1, shuffle=False, verbose=0)
    for a, b in zip(ctl.get_weights(), fitted.get_weights()):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)


def test_minimize_overlaps_the_update_with_backward_and_matches_apply_gradients():
    x, y = _blobs(64)
    xent = keras.losses.SparseCategoricalCrossentropy(from_logits=True)
    a, opt_a = _ctl_model(), keras.optimizers.SGD(0.1, momentum=0.9)
    b, opt_b = _ctl_model(), keras.optimizers.SGD(0.1, momentum=0.9)
    for step in range(4):
        xb, yb = x[step * 16:(step + 1) * 16], y[step * 16:(step + 1) * 16]
        with tfc.GradientTape() as tape:
            loss = xent(yb, a(xb, training=True))
        opt_a.minimize(loss, a.trainable_variables, tape=tape)          # step 0: engine is created here
        with tfc.GradientTape() as tape:
            loss = xent(yb, b(xb, training=True))
        opt_b.apply_gradients(zip(tape.gradient(loss, b.trainable_variables), b.trainable_variables))
    for wa, wb in zip(a.get_weights(), b.get_weights()):
        np.testing.assert_allclose(wa, wb, rtol=1e-5, atol=1e-6)
    # Keras also accepts a callable loss
    before = a.get_weights()[0].copy()
    opt_a.minimize(lambda: xent(y[:16], a(x[:16], training=True)), a.trainable_variables)
    assert not np.array_equal(before, a.get_weights()[0]) and opt_a.iterations == 5


def test_tape_details():
    w = torch.nn.Parameter(torch.tensor([1.0, 2.0]))
    unused = torch.nn.Parameter(torch.tensor([5.0]))
    x = torch.tensor([3.0, 4.0])
    with tfc.GradientTape(persistent=True) as tape:
        tape.watch(x)
        out = (w * x).sum()
    assert tape.gradient(out, w).tolist() == [3.0, 4.0]
    assert tape.gradient(out, x).tolist() == [1.0, 2.0]                   # persistent: a second call
    g = tape.gradient(out, [w, unused])
    assert g[1] is None
    assert tape.gradient(out, [w, unused], unconnected_gradients="zero")[1].tolist() == [0.0]
    with tfc.GradientTape() as once:
        out = (w * w).sum()
    once.gradient(out, w)
    with pytest.raises(RuntimeError, match="non-persistent"):
        once.gradient(out, w)
    per_example = torch.tensor([1.0, 2.0, 3.0])
    assert float(tfc.nn.compute_average_loss(per_example, global_batch_size=6)) == 1.0
    assert float(tfc.nn.compute_average_loss(per_example)) == 2.0          # one replica: plain mean


def test_reduce_lr_on_plateau_and_csv_logger(tmp_path):
    x, y = _blobs(128)
    m = _mlp()
    m.compile(optimizer=keras.optimizers.SGD(0.5), loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    log = tmp_path / "logs" / "train.csv"
    # min_delta so large that no epoch counts as an improvement: the rate halves every second epoch
    plateau = keras.callbacks.ReduceLROnPlateau(monitor="loss", factor=0.5, patience=2, min_delta=100.0,
                                                min_lr=0.1)
    hist = m.fit(x, y, batch_size=32, epochs=7, verbose=0,
                 callbacks=[plateau, keras.callbacks.CSVLogger(str(log))])
    assert hist.history["lr"] == [0.5, 0.5, 0.5, 0.25, 0.25, 0.125, 0.125]
    assert float(m.optimizer.lr) == 0.1                                   # clipped at min_lr
    rows = log.read_text().strip().splitlines()
    assert rows[0] == "epoch,accuracy,loss,lr" and len(rows) == 8
    assert float(rows[-1].split(",")[2]) == pytest.approx(hist.history["loss"][-1])
    m.fit(x, y, batch_size=32, epochs=1, verbose=0, callbacks=[keras.callbacks.CSVLogger(str(log), append=True)])
    assert len(log.read_text().strip().splitlines()) == 9                 # appended, one header
    with pytest.raises(ValueError):
        keras.callbacks.ReduceLROnPlateau(factor=1.5)


def test_fit_keyword_contract():
    x, y = _blobs(64)
    m = _mlp()
    hist = m.fit(x, y, batch_size=32, epochs=4, verbose=0, validation_data=(x, y), validation_freq=2)
    assert len(hist.history["loss"]) == 4 and len(hist.history["val_loss"]) == 2      # epochs 2 and 4
    with pytest.raises(NotImplementedError, match="array inputs"):
        m.fit(tfc.data.Dataset.from_tensor_slices((x, y)).batch(16), sample_weight=np.ones(len(x)), verbose=0)
    with pytest.raises(TypeError, match="unexpected keyword"):
        m.fit(x, y, epochz=3)
    m.fit(x, y, batch_size=32, epochs=1, verbose=0, workers=4, use_multiprocessing=False)   # accepted, meaningless here


def test_weights_from_a_dataset_match_weights_from_arrays():
    """Keras: a dataset may yield (x, y, sample_weight); class_weight applies to every batch's labels."""
    x, y = _blobs(64)
    w = (1.0 + (np.arange(64) % 3)).astype("float32")
    cw = {0: 2.0, 3: 0.5}

    def run(data, **fit_kwargs):
        torch.manual_seed(9)
        m = keras.Sequential([L.Dense(16, activation="relu", input_shape=(20,)), L.Dense(4, activation="softmax")])
        m.compile(keras.optimizers.SGD(0.1), "sparse_categorical_crossentropy", metrics=["accuracy"])
        if isinstance(data, tuple):
            m.fit(*data, batch_size=16, epochs=2, shuffle=False, verbose=0, **fit_kwargs)
        else:
            m.fit(data, epochs=2, verbose=0, **fit_kwargs)
        assert m._weighted_targets is None                      # the flag does not outlive the call
        return m.get_weights()

    want = run((x, y), sample_weight=w)
    got = run(tfc.data.Dataset.from_tensor_slices((x, y, w)).batch(16))
    for a, b in zip(got, want):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)
    want = run((x, y), class_weight=cw)
    got = run(tfc.data.Dataset.from_tensor_slices((x, y)).batch(16), class_weight=cw)
    for a, b in zip(got, want):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)
    want = run((x, y), sample_weight=w, class_weight=cw)
    got = run(tfc.data.Dataset.from_tensor_slices((x, y, w)).batch(16), class_weight=cw)
    for a, b in zip(got, want):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)
    # nothing to weight: the dataset goes through untouched (and the fused loss path stays available)
    plain = run(tfc.data.Dataset.from_tensor_slices((x, y)).batch(16))
    for a, b in zip(plain, run((x, y))):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)
    with pytest.raises(NotImplementedError, match="sample_weight"):
        run(tfc.data.Dataset.from_tensor_slices((x, y)).batch(16), sample_weight=w)


def test_sample_and_class_weights_scale_each_sample_loss():
    x, y = _blobs(64)
    # weight 0 for half of the samples == training on the other half with a batch twice as "large"
    w = (np.arange(64) % 2).astype("float32")

    def run(**fit_kwargs):
        torch.manual_seed(5)
        m = keras.Sequential([L.Dense(16, activation="relu", input_shape=(20,)), L.Dense(4, activation="softmax")])
        m.compile(keras.optimizers.SGD(0.1), "sparse_categorical_crossentropy", metrics=["accuracy"])
        hist = m.fit(x, y, batch_size=16, epochs=2, shuffle=False, verbose=0, **fit_kwargs)
        return m.get_weights(), hist.history

    weighted, hist_w = run(sample_weight=w)
    # oracle: plain torch, the same SGD steps with sum(w * loss) / batch
    torch.manual_seed(5)
    ref = keras.Sequential([L.Dense(16, activation="relu", input_shape=(20,)), L.Dense(4, activation="softmax")])
    ref(x[:1])
    params = [p for p in ref.parameters()]
    for _epoch in range(2):
        for i in range(0, 64, 16):
            out = ref(x[i:i + 16], training=True)
            per = torch.nn.functional.nll_loss(torch.log(out.float()), torch.tensor(y[i:i + 16]), reduction="none")
            loss = (per * torch.tensor(w[i:i + 16])).sum() / 16
            grads = torch.autograd.grad(loss, params)
            with torch.no_grad():
                for p, g in zip(params, grads):
                    p -= 0.1 * g
    for a, b in zip(weighted, ref.get_weights()):
        np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-5)
    plain, hist_p = run()
    assert not np.allclose(weighted[0], plain[0])
    assert hist_w["loss"][0] < hist_p["loss"][0]                       # half of the terms are zeroed
    # class weights are per-sample weights looked up from the labels; both kinds multiply
    by_class, _ = run(class_weight={0: 2.0, 3: 0.0})
    lookup = np.ones(4, dtype="float32")
    lookup[0], lookup[3] = 2.0, 0.0
    by_sample, _ = run(sample_weight=lookup[y])
    for a, b in zip(by_class, by_sample):
        np.testing.assert_allclose(a, b, rtol=1e-6)
    both, _ = run(sample_weight=w, class_weight={0: 2.0, 3: 0.0})
    product, _ = run(sample_weight=w * lookup[y])
    for a, b in zip(both, product):
        np.testing.assert_allclose(a, b, rtol=1e-6)
    with pytest.raises(ValueError, match="sample_weight has"):
        run(sample_weight=np.ones(3))
    # the flag does not leak into later calls
    m = _mlp()
    m.fit(x, y, sample_weight=np.ones(64), epochs=1, verbose=0)
    assert m._weighted_targets is None and np.isfinite(m.evaluate(x, y, verbose=0)[0])
    m.fit(x, y, epochs=1, verbose=0)


def test_more_learning_rate_schedules():
    sch = keras.optimizers.schedules
    cos = sch.CosineDecay(0.1, decay_steps=100, alpha=0.1)
    assert cos(0) == pytest.approx(0.1) and cos(50) == pytest.approx(0.1 * (0.9 * 0.5 + 0.1)) and cos(100) == cos(500) == pytest.approx(0.01)
    warm = sch.CosineDecay(0.0, decay_steps=10, warmup_target=1.0, warmup_steps=4)
    assert [round(warm(s), 3) for s in (0, 2, 4, 9, 14)] == [0.0, 0.5, 1.0, 0.5, 0.0]
    poly = sch.PolynomialDecay(1.0, decay_steps=10, end_learning_rate=0.1, power=2.0)
    assert poly(0) == 1.0 and poly(5) == pytest.approx(0.9 * 0.25 + 0.1) and poly(50) == pytest.approx(0.1)
    assert sch.PolynomialDecay(1.0, 10, 0.0, cycle=True)(15) == pytest.approx(0.25)
    inv = sch.InverseTimeDecay(1.0, decay_steps=10, decay_rate=0.5, staircase=True)
    assert inv(9) == 1.0 and inv(25) == pytest.approx(1 / 2.0)
    # a schedule drives the optimizer during fit and survives serialisation
    x, y = _blobs(64)
    m = _mlp()
    m.compile(keras.optimizers.SGD(sch.CosineDecay(0.5, decay_steps=8)), "sparse_categorical_crossentropy")
    m.fit(x, y, batch_size=16, epochs=2, verbose=0)
    assert m.optimizer.current_lr() == pytest.approx(0.0, abs=1e-9) and m.optimizer.iterations == 8
    again = keras.optimizers.deserialize(keras.optimizers.serialize(m.optimizer))
    assert isinstance(again._schedule, sch.CosineDecay) and again._schedule.decay_steps == 8


def test_precision_recall_and_top_k_metrics():
    m = keras.metrics
    y_true = np.array([1, 0, 1, 1, 0, 0])
    scores = np.array([0.9, 0.8, 0.4, 0.7, 0.2, 0.1], dtype="float32")       # predicted positive: 0, 1, 3
    p, r = m.Precision(), m.Recall()
    p.update_state(y_true, scores), r.update_state(y_true, scores)
    assert p.result() == pytest.approx(2 / 3) and r.result() == pytest.approx(2 / 3)
    p.update_state(np.array([1, 1]), np.array([0.9, 0.1], dtype="float32"))
    assert p.result() == pytest.approx(3 / 4)                                 # sums, not a mean of ratios
    probs = np.array([[0.1, 0.7, 0.2], [0.6, 0.3, 0.1], [0.1, 0.3, 0.6]], dtype="float32")
    labels = np.array([1, 1, 0])
    per_class = m.Precision(class_id=1)
    per_class.update_state(labels, probs)
    assert per_class.result() == 1.0                                          # only sample 0 crosses 0.5 for class 1
    top2 = m.SparseTopKCategoricalAccuracy(k=2)
    top2.update_state(labels, probs)
    assert top2.result() == pytest.approx(2 / 3)
    assert m.TopKCategoricalAccuracy(k=1)(np.eye(3)[labels], probs) == pytest.approx(1 / 3)
    # inside fit / evaluate: two metrics with denominators of their own next to plain ones
    x, y = _blobs(128, classes=2)
    model = keras.Sequential([L.Dense(8, activation="relu", input_shape=(20,)), L.Dense(1, activation="sigmoid")])
    model.compile(keras.optimizers.Adam(1e-2), "binary_crossentropy",
                  metrics=["accuracy", m.Precision(), "recall", "mae"])
    hist = model.fit(x, y.astype("float32"), batch_size=32, epochs=6, verbose=0)
    assert set(hist.history) == {"loss", "accuracy", "precision", "recall", "mae"}
    logs = model.evaluate(x, y.astype("float32"), verbose=0, return_dict=True)
    pred = model.predict(x).reshape(-1) > 0.5
    tp = float((pred & (y == 1)).sum())
    assert logs["precision"] == pytest.approx(tp / max(pred.sum(), 1), abs=1e-6)
    assert logs["recall"] == pytest.approx(tp / (y == 1).sum(), abs=1e-6)
    assert logs["accuracy"] == pytest.approx(float((pred == (y == 1)).mean()), abs=1e-6)


def test_keras_sequence_feeds_fit_evaluate_and_predict_sources():
    class Batches(keras.utils.Sequence):
        def __init__(self):
            self.x, self.y = _blobs(96)
            self.epochs_seen = 0

        def __len__(self):
            return 6

        def __getitem__(self, i):
            return self.x[i * 16:(i + 1) * 16], self.y[i * 16:(i + 1) * 16]

        def on_epoch_end(self):
            self.epochs_seen += 1

    seq = Batches()
    m = _mlp()
    hist = m.fit(seq, epochs=3, verbose=0)
    assert len(hist.history["loss"]) == 3 and seq.epochs_seen == 3
    assert hist.history["accuracy"][-1] > 0.8
    loss, acc = m.evaluate(seq, verbose=0)
    assert acc > 0.8 and np.isfinite(loss)

    class Duck:                                    # not derived from Sequence, same protocol
        def __len__(self):
            return 2

        def __getitem__(self, i):
            return seq[i]
    assert len(m.fit(Duck(), epochs=1, verbose=0).history["loss"]) == 1


def test_auc_metric_matches_a_rank_based_reference():
    rng = np.random.RandomState(0)
    y = rng.randint(0, 2, 400)
    scores = np.clip(0.35 * y + rng.rand(400) * 0.65, 0, 1).astype("float32")
    auc = keras.metrics.AUC(num_thresholds=100)
    for i in range(0, 400, 100):                       # accumulated over batches
        auc.update_state(y[i:i + 100], scores[i:i + 100])
    pos, neg = scores[y == 1], scores[y == 0]
    exact = float(((pos[:, None] > neg[None, :]).sum() + 0.5 * (pos[:, None] == neg[None, :]).sum()) / (len(pos) * len(neg)))
    assert auc.result() == pytest.approx(exact, abs=5e-3)
    perfect = keras.metrics.AUC()
    perfect.update_state([0, 0, 1, 1], [0.1, 0.2, 0.8, 0.9])
    assert perfect.result() == pytest.approx(1.0) and keras.metrics.AUC().result() == 0.0
    # through fit / evaluate next to other metrics
    x, yb = _blobs(128, classes=2)
    model = keras.Sequential([L.Dense(8, activation="relu", input_shape=(20,)), L.Dense(1, activation="sigmoid")])
    model.compile(keras.optimizers.Adam(1e-2), "binary_crossentropy", metrics=["accuracy", "auc", "precision"])
    hist = model.fit(x, yb.astype("float32"), batch_size=32, epochs=5, verbose=0)
    assert hist.history["auc"][-1] > 0.9
    logs = model.evaluate(x, yb.astype("float32"), verbose=0, return_dict=True)
    p = model.predict(x).reshape(-1)
    pos, neg = p[yb == 1], p[yb == 0]
    exact = float((pos[:, None] > neg[None, :]).mean())
    assert logs["auc"] == pytest.approx(exact, abs=1e-2) and 0 <= logs["precision"] <= 1
    with pytest.raises(ValueError, match="scalars per step"):
        model.compile("adam", "binary_crossentropy", metrics=[keras.metrics.AUC(name="a"), keras.metrics.AUC(name="b")])
