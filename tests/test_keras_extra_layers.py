"""Layers of keras/extra_layers.py against plain PyTorch references, their Keras weight layouts, and
a model that uses them end to end (build, train, save, load)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import cloud_b200 as tfc

keras = tfc.keras
L = keras.layers


@pytest.fixture(autouse=True)
def _cpu():
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    torch.manual_seed(0)
    keras.backend.clear_session()


def _nhwc(x):           # a layer inside a model sees NCHW-logical tensors
    return torch.tensor(x).permute(0, 3, 1, 2)


def test_separable_and_depthwise_convolutions_match_torch_and_keep_keras_layouts():
    x = np.random.rand(2, 9, 8, 3).astype("float32")
    sep = L.SeparableConv2D(5, 3, strides=2, padding="same", depth_multiplier=2, activation="relu")
    y = sep(_nhwc(x))
    dk, pk, b = sep.get_weights()
    assert dk.shape == (3, 3, 3, 2) and pk.shape == (1, 1, 6, 5) and b.shape == (5,)
    # reference from the Keras-layout weights
    dw = torch.tensor(dk).permute(2, 3, 0, 1).reshape(6, 1, 3, 3)
    xi = F.pad(_nhwc(x), (0, 1, 1, 1))                      # 'same' for 9x8, k 3, s 2: rows (1,1), cols (0,1)
    ref = F.conv2d(xi, dw, None, stride=2, groups=3)
    ref = F.conv2d(ref, torch.tensor(pk).permute(3, 2, 0, 1), torch.tensor(b)).relu()
    np.testing.assert_allclose(y.detach().numpy(), ref.numpy(), atol=1e-5)
    assert tuple(y.shape) == (2, 5, 5, 4) and sep.compute_output_shape((None, 9, 8, 3)) == (None, 5, 4, 5)
    # weights written in the Keras layout come back unchanged
    new = [np.random.rand(*w.shape).astype("float32") for w in (dk, pk, b)]
    sep.set_weights(new)
    for a, c in zip(sep.get_weights(), new):
        np.testing.assert_array_equal(a, c)
    dwl = L.DepthwiseConv2D(3, padding="valid", use_bias=False)
    out = dwl(_nhwc(x))
    (k,) = dwl.get_weights()
    ref = F.conv2d(_nhwc(x), torch.tensor(k).permute(2, 3, 0, 1).reshape(3, 1, 3, 3), groups=3)
    np.testing.assert_allclose(out.detach().numpy(), ref.numpy(), atol=1e-5)


def test_conv1d_pooling_and_causal_padding():
    x = np.random.rand(2, 11, 4).astype("float32")
    conv = L.Conv1D(6, 3, strides=2, padding="same", activation="tanh")
    y = conv(torch.tensor(x))
    k, b = conv.get_weights()
    assert k.shape == (3, 4, 6)
    ref = F.conv1d(F.pad(torch.tensor(x).transpose(1, 2), (1, 1)), torch.tensor(k).permute(2, 1, 0),
                   torch.tensor(b), stride=2).transpose(1, 2).tanh()
    np.testing.assert_allclose(y.detach().numpy(), ref.numpy(), atol=1e-5)
    assert conv.compute_output_shape((None, 11, 4)) == (None, 6, 6)
    causal = L.Conv1D(2, 3, padding="causal", dilation_rate=2, use_bias=False)
    yc = causal(torch.tensor(x))
    assert yc.shape == (2, 11, 2)
    x2 = x.copy()
    x2[:, 7:] += 1.0                                         # the future must not leak into the past
    np.testing.assert_allclose(causal(torch.tensor(x2))[:, :7].detach().numpy(), yc[:, :7].detach().numpy(), atol=1e-6)
    assert L.MaxPooling1D(2)(torch.tensor(x)).shape == (2, 5, 4)
    assert L.MaxPooling1D(3, strides=2, padding="same")(torch.tensor(x)).shape == (2, 6, 4)
    np.testing.assert_allclose(L.GlobalAveragePooling1D()(torch.tensor(x)).numpy(), x.mean(1), atol=1e-6)
    np.testing.assert_allclose(L.GlobalMaxPooling1D()(torch.tensor(x)).numpy(), x.max(1), atol=1e-6)
    np.testing.assert_allclose(L.AveragePooling1D(2)(torch.tensor(x)).numpy(),
                               x[:, :10].reshape(2, 5, 2, 4).mean(2), atol=1e-6)


def test_normalisation_embedding_activations_merges_and_resampling():
    x = np.random.randn(3, 5, 8).astype("float32")
    ln = L.LayerNormalization(epsilon=1e-5)
    y = ln(torch.tensor(x))
    np.testing.assert_allclose(y.detach().numpy(), (x - x.mean(-1, keepdims=True)) / np.sqrt(x.var(-1, keepdims=True) + 1e-5),
                               atol=1e-4)
    img = np.random.randn(2, 4, 4, 6).astype("float32")
    yi = ln_img = L.LayerNormalization(epsilon=1e-5)(_nhwc(img))
    ref = (img - img.mean(-1, keepdims=True)) / np.sqrt(img.var(-1, keepdims=True) + 1e-5)
    np.testing.assert_allclose(yi.permute(0, 2, 3, 1).detach().numpy(), ref, atol=1e-4)
    emb = L.Embedding(1000, 4, input_length=7)
    ids = np.array([[0, 999, 257, 5, 5, 3, 1]])
    out = emb(torch.tensor(ids))
    (table,) = emb.get_weights()
    assert table.shape == (1000, 4)
    np.testing.assert_allclose(out.detach().numpy(), table[ids], atol=0)
    t = torch.tensor([[-2.0, 0.5]])
    assert L.LeakyReLU(0.1)(t)[0].tolist() == pytest.approx([-0.2, 0.5])
    assert L.ELU()(t)[0, 0] == pytest.approx(np.exp(-2.0) - 1)
    pre = L.PReLU()
    assert pre(t).tolist() == [[0.0, 0.5]]                   # alpha starts at zero
    a, b = torch.tensor([[1.0, 4.0]]), torch.tensor([[3.0, 2.0]])
    assert L.Subtract()([a, b]).tolist() == [[-2.0, 2.0]] and L.Multiply()([a, b]).tolist() == [[3.0, 8.0]]
    assert L.Average()([a, b]).tolist() == [[2.0, 3.0]] and L.Maximum()([a, b]).tolist() == [[3.0, 4.0]]
    up = L.UpSampling2D(2)(_nhwc(img))
    assert up.shape == (2, 6, 8, 8) and torch.equal(up[:, :, ::2, ::2], _nhwc(img))
    crop = L.Cropping2D(((1, 0), (0, 2)))(_nhwc(img))
    assert crop.shape == (2, 6, 3, 2) and L.Cropping2D(1).compute_output_shape((None, 4, 4, 6)) == (None, 2, 2, 6)


def test_models_built_from_these_layers_train_save_and_load(tmp_path):
    # a small separable-convolution image model
    inp = L.Input(shape=(12, 12, 3))
    h = L.SeparableConv2D(8, 3, padding="same")(inp)
    h = L.LeakyReLU(0.1)(L.BatchNormalization()(h))
    h = L.UpSampling2D(2)(L.MaxPooling2D()(h))
    h = L.Multiply()([h, L.Conv2D(8, 1, activation="sigmoid")(h)])
    out = L.Dense(3, activation="softmax")(L.GlobalAveragePooling2D()(L.LayerNormalization()(h)))
    model = keras.Model(inp, out)
    model.compile("adam", "sparse_categorical_crossentropy", metrics=["accuracy"])
    x = np.random.rand(32, 12, 12, 3).astype("float32")
    y = np.random.randint(0, 3, 32)
    hist = model.fit(x, y, batch_size=8, epochs=3, verbose=0)
    assert hist.history["loss"][-1] < hist.history["loss"][0]
    model.save(str(tmp_path / "img"))
    again = keras.models.load_model(str(tmp_path / "img"))
    np.testing.assert_allclose(again.predict(x), model.predict(x), atol=1e-5)
    # a text model: token ids stay integers on their way in
    text = keras.Sequential([L.Embedding(500, 8, input_length=10), L.Conv1D(16, 3, padding="same", activation="relu"),
                             L.GlobalMaxPooling1D(), L.Dense(2, activation="softmax")])
    text.compile(keras.optimizers.Adam(1e-2), "sparse_categorical_crossentropy", metrics=["accuracy"])
    ids = np.random.randint(0, 500, (64, 10))
    labels = (ids[:, 0] > 250).astype("int64")
    hist = text.fit(ids, labels, batch_size=16, epochs=12, verbose=0)
    assert hist.history["accuracy"][-1] > 0.8
    text.save_weights(str(tmp_path / "text.h5"))
    from cloud_b200.io import hdf5
    with hdf5.File(str(tmp_path / "text.h5"), "r") as f:
        assert f["embedding"]["embedding"]["embeddings:0"].shape == (500, 8)
        assert f["conv1d"]["conv1d"]["kernel:0"].shape == (3, 8, 16)
