"""Checkpoint / restore across strategies (reference: core/tests/testdata/save_and_load.py):
ModelCheckpoint(save_weights_only=True) under MirroredStrategy, load_weights under
OneDeviceStrategy, full-model save + load_model, then continue training."""
import argparse
import os

import numpy as np

import cloud_b200 as tfc

keras = tfc.keras
parser = argparse.ArgumentParser()
parser.add_argument("--path", default="./save_and_load_out")
args, _ = parser.parse_known_args()

strategy = tfc.distribute.MirroredStrategy()
BATCH = 64 * strategy.num_replicas_in_sync
rng = np.random.RandomState(0)
x = rng.rand(1024, 28, 28, 1).astype("float32")
y = rng.randint(0, 10, 1024)


def create_model():
    m = keras.Sequential([
        keras.layers.Conv2D(16, 3, activation="relu", input_shape=(28, 28, 1)),
        keras.layers.MaxPooling2D(),
        keras.layers.Flatten(),
        keras.layers.Dense(32, activation="relu"),
        keras.layers.Dense(10)])
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True),
              optimizer=keras.optimizers.Adam(), metrics=["accuracy"])
    return m


with strategy.scope():
    model = create_model()
    ckpt = os.path.join(args.path, "ckpt", "cp.ckpt")
    model.fit(x, y, batch_size=BATCH, epochs=2, verbose=0,
              callbacks=[keras.callbacks.ModelCheckpoint(filepath=ckpt, save_weights_only=True)])
    base = model.evaluate(x, y, batch_size=BATCH, verbose=0)

# restore on a single device
with tfc.distribute.OneDeviceStrategy("/cpu:0").scope():
    restored = create_model()
    restored.build((None, 28, 28, 1))
    restored.load_weights(keras.saving.latest_checkpoint(os.path.dirname(ckpt)))
    again = restored.evaluate(x, y, batch_size=64, verbose=0)
assert abs(again[1] - base[1]) < 0.05, (again, base)

# full model: save outside any scope, load with and without a strategy, keep training
saved = os.path.join(args.path, "saved_model")
model.save(saved)
strategy.barrier()
plain = keras.models.load_model(saved)
plain.fit(x, y, batch_size=64, epochs=1, verbose=0)
with strategy.scope():
    replicated = keras.models.load_model(saved)
    replicated.fit(x, y, batch_size=BATCH, epochs=1, verbose=0)
if strategy.is_chief:
    print("save_and_load OK", base, again)
