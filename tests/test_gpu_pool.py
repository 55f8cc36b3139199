"""GPU numerics: NHWC max-pool kernels (zero padding folded in) against a PyTorch fp32 oracle, and
the ResNet stem / conv-bias-absorption / direct-gradient paths of the Keras front end."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from cloud_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,C,H,k,s,pad", [(4, 64, 112, 3, 2, 1), (2, 32, 17, 2, 2, 0),
                                            (3, 128, 30, 3, 1, 1), (2, 8, 9, 3, 2, 0)])
def test_max_pool_nhwc(N, C, H, k, s, pad):
    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    # strictly positive, distinct values: no ties and zero padding never wins
    x = (torch.rand(N, C, H, H, device=dev) + 0.5).to(torch.bfloat16)
    x = x.contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = ops.max_pool_nhwc(x, k, s, pad)
    xr = x.detach().float().requires_grad_(True)
    yr = F.max_pool2d(F.pad(xr, (pad, pad, pad, pad)), k, s)
    assert y.shape == yr.shape
    assert torch.equal(y.float(), yr.detach())
    gy = torch.randn_like(yr)
    y.backward(gy.to(torch.bfloat16))
    yr.backward(gy.to(torch.bfloat16).float())
    # bf16 ties inside a window (equal after rounding) may route the gradient differently:
    # compare where the window maxima are unique
    diff = (x.grad.float() - xr.grad).abs()
    assert (diff > 1e-2).float().mean().item() < 0.02
    assert abs(x.grad.float().sum().item() - xr.grad.sum().item()) < 1e-2 * gy.abs().sum().item()


def test_resnet_stem_training_step_matches_torch():
    """conv(pad=3)+BN+ReLU+maxpool(pad=1) block trained one step on the fused kernels vs the same
    maths in fp32 torch ops."""
    import cloud_b200 as tfc
    from cloud_b200.keras import layers
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
    torch.manual_seed(0)
    inp = keras.Input(shape=(32, 32, 8))
    x = layers.Conv2D(16, 3, strides=1, padding=1, name="c")(inp)
    x = layers.BatchNormalization(activation="relu", name="bn")(x)
    x = layers.MaxPooling2D(3, strides=2, padding=1, name="p")(x)
    x = layers.GlobalAveragePooling2D()(x)
    out = layers.Dense(4, activation="softmax", name="d")(x)
    m = keras.Model(inp, out)
    m.compile(optimizer=keras.optimizers.SGD(0.1), loss="sparse_categorical_crossentropy",
              run_eagerly=True)
    xs = np.random.RandomState(0).rand(16, 32, 32, 8).astype("float32")
    ys = np.random.RandomState(1).randint(0, 4, 16)
    w0 = [w.copy() for w in m.get_weights()]
    logs = m.train_on_batch(xs, ys, return_dict=True)
    w1 = m.get_weights()
    # fp32 oracle with the same initial weights
    k, b, gamma, beta, mm, mv, dk, db = [torch.tensor(w, device="cuda", requires_grad=True) for w in w0]
    xt = torch.tensor(xs, device="cuda").permute(0, 3, 1, 2)
    h = F.conv2d(xt, k.permute(3, 2, 0, 1), b, padding=1)
    h = F.batch_norm(h, None, None, gamma, beta, True, 0.0, 1e-3)
    h = F.max_pool2d(F.pad(torch.relu(h), (1, 1, 1, 1)), 3, 2)
    logits = h.mean((2, 3)) @ dk + db
    loss = F.cross_entropy(logits, torch.tensor(ys, device="cuda"))
    loss.backward()
    assert abs(logs["loss"] - loss.item()) < 5e-2
    for name, before, after, ref in (("kernel", w0[0], w1[0], k), ("gamma", w0[2], w1[2], gamma),
                                     ("beta", w0[3], w1[3], beta), ("dense", w0[6], w1[6], dk)):
        got = (after - before) / -0.1
        want = ref.grad.detach().cpu().numpy()
        # the gradient is inferred from a bf16 weight delta (8 mantissa bits), so compare the
        # direction and the overall size rather than element-wise
        cos = float((got * want).sum() / (np.linalg.norm(got) * np.linalg.norm(want) + 1e-12))
        assert cos > 0.97, (name, cos)
        assert 0.8 < np.linalg.norm(got) / (np.linalg.norm(want) + 1e-12) < 1.25, name
    # the absorbed conv bias gets a zero gradient, exactly
    assert np.array_equal(w0[1], w1[1])


def test_steps_per_execution_graph_matches_single_step_training():
    """K steps captured in one CUDA graph give the weights K single-step replays give."""
    import cloud_b200 as tfc
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
    rng = np.random.RandomState(0)
    x = rng.rand(24 * 64, 48).astype("float32")
    y = rng.randint(0, 5, 24 * 64)
    out = {}
    for spe in (1, 6):
        torch.manual_seed(0)
        m = keras.Sequential([keras.layers.Dense(64, activation="relu", input_shape=(48,)),
                              keras.layers.Dense(5, activation="softmax")])
        m.compile(optimizer=keras.optimizers.Adam(1e-3), loss="sparse_categorical_crossentropy",
                  metrics=["accuracy"], steps_per_execution=spe)
        h = m.fit(x, y, batch_size=64, epochs=3, shuffle=False, verbose=0)
        out[spe] = (m.get_weights(), h.history["loss"])
        if spe > 1:
            assert m._graph_k is not None and m._graph_k.get("graph") is not None
    for a, b in zip(out[1][0], out[6][0]):
        assert np.allclose(a, b, atol=2e-2, rtol=2e-2)
    assert np.allclose(out[1][1], out[6][1], rtol=2e-2)
