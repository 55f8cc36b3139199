"""Worker script (run under torch.distributed.run): a custom training loop - GradientTape +
optimizer.apply_gradients / minimize, per-example losses averaged over the GLOBAL batch - on
WORLD_SIZE replicas must end with the weights of a single replica fed the same global batches
(reference semantics: core/tests/testdata/mnist_example_using_ctl.py:97-131).
Works on CPU ranks (shared-memory collectives) and on GPUs (peer-memory kernels)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cloud_b200 as tfc  # noqa: E402

keras = tfc.keras
L = keras.layers
GLOBAL_BATCH, STEPS = 64, 6


def build(seed):
    torch.manual_seed(seed)
    keras.backend.clear_session()
    return keras.Sequential([L.Dense(64, activation="relu", input_shape=(32,)), L.Dense(4)])


def train(strategy, model, x, y, lo, hi):
    loss_object = keras.losses.SparseCategoricalCrossentropy(from_logits=True,
                                                             reduction=keras.losses.Reduction.NONE)
    optimizer = keras.optimizers.Adam(1e-2)
    total = 0.0
    for step in range(STEPS):
        xb = x[step * GLOBAL_BATCH:(step + 1) * GLOBAL_BATCH][lo:hi]
        yb = y[step * GLOBAL_BATCH:(step + 1) * GLOBAL_BATCH][lo:hi]
        with tfc.GradientTape() as tape:
            logits = model(xb, training=True)
            loss = tfc.nn.compute_average_loss(loss_object(yb, logits), global_batch_size=GLOBAL_BATCH)
        if step % 2:
            optimizer.minimize(loss, model.trainable_variables, tape=tape)
        else:
            optimizer.apply_gradients(zip(tape.gradient(loss, model.trainable_variables),
                                          model.trainable_variables))
        total = float(strategy.reduce(tfc.distribute.ReduceOp.SUM, loss.detach(), axis=None))
    return total, optimizer


def main():
    use_gpu = torch.cuda.is_available() and os.environ.get("CHECK_DEVICE", "auto") != "cpu"
    rng = np.random.RandomState(0)
    x = rng.rand(GLOBAL_BATCH * STEPS, 32).astype("float32")
    y = x[:, :4].argmax(1).astype("int64")
    strategy = tfc.distribute.MirroredStrategy()
    world, rank = strategy.num_replicas_in_sync, strategy.replica_id
    per = GLOBAL_BATCH // world
    with strategy.scope():
        model = build(1 + rank)                 # replicas start apart: the engine's broadcast aligns them
        last, opt = train(strategy, model, x, y, rank * per, (rank + 1) * per)
    backend = opt._engine.comm.backend
    w_dist = model.get_weights()

    single = tfc.distribute.OneDeviceStrategy(strategy.device)
    with single.scope():
        ref = build(1)
        ref._strategy = single
        last_ref, _ = train(single, ref, x, y, 0, GLOBAL_BATCH)
    tol = 5e-2 if use_gpu else 1e-5
    worst = max(float(np.abs(a - b).max()) for a, b in zip(w_dist, ref.get_weights()))
    assert worst < tol, "rank {}: weights differ from the single-replica loop by {}".format(rank, worst)
    assert abs(last - last_ref) < tol, (last, last_ref)
    if rank == 0:
        print("CTL OK world={} backend={} worst_diff={:.3g}".format(world, backend, worst), flush=True)


if __name__ == "__main__":
    main()
