"""Custom training loop under a user-managed strategy (reference:
core/tests/testdata/mnist_example_using_ctl.py: strategy.scope, experimental_distribute_dataset,
per-replica batch = 64, loss divided by the GLOBAL batch, strategy.reduce(SUM)).  The per-step
work is plain torch autograd; the engine's begin_step/end_step pair replaces
`optimizer.apply_gradients` (fused all-reduce + Adam on the hand-written kernels)."""
import numpy as np
import torch

import cloud_b200 as tfc
from cloud_b200 import ops
from cloud_b200.parallel.engine import DPEngine

keras = tfc.keras

strategy = tfc.distribute.MultiWorkerMirroredStrategy()
BATCH_PER_REPLICA = 64
GLOBAL_BATCH = BATCH_PER_REPLICA * strategy.num_replicas_in_sync

(x, y), _ = keras.datasets.mnist.load_data()
x = (x[:4096, ..., None] / 255.0).astype("float32")
y = y[:4096].astype("int64")
dataset = tfc.data.Dataset.from_tensor_slices((x, y)).shuffle(4096, seed=0).batch(GLOBAL_BATCH)

with strategy.scope():
    model = keras.Sequential([
        keras.layers.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
        keras.layers.MaxPooling2D(),
        keras.layers.Conv2D(64, 3, activation="relu"),
        keras.layers.MaxPooling2D(),
        keras.layers.Flatten(),
        keras.layers.Dense(64, activation="relu"),
        keras.layers.Dense(10),
    ])
    model.compile(optimizer=keras.optimizers.Adam(),
                  loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    engine = model._ensure_engine(x[:1])           # flat arenas + fused optimizer state
    dist_dataset = strategy.experimental_distribute_dataset(dataset)
    accuracy = keras.metrics.SparseCategoricalAccuracy()


def train_step(inputs):
    images, labels = inputs
    images = model._to_internal(images)
    labels = torch.as_tensor(labels, device=strategy.device)
    engine.begin_step()
    logits = model.call(images, training=True)
    # sum of per-example losses divided by the GLOBAL batch (mnist_example_using_ctl.py:97-101)
    loss = ops.softmax_cross_entropy(logits, labels, grad_scale=1.0 / GLOBAL_BATCH)
    loss.backward()
    engine.end_step()
    accuracy.update_state(labels.cpu(), logits.detach().float().cpu())
    return loss.detach()


for epoch in range(2):
    total, batches = 0.0, 0
    model.train(True)
    for batch in dist_dataset:
        per_replica = strategy.run(train_step, args=(batch,))
        total += float(strategy.reduce(tfc.distribute.ReduceOp.SUM, per_replica, axis=None))
        batches += 1
    if strategy.is_chief:
        print("Epoch {}: loss {:.4f} accuracy {:.4f}".format(epoch, total / max(batches, 1), accuracy.result()))
    accuracy.reset_states()
