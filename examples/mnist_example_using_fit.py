"""MNIST CNN trained with model.fit - run it directly or through tfc.run (see call_run_on_script.py).
Counterpart of the reference's core/tests/testdata/mnist_example_using_fit.py (callbacks:
LearningRateScheduler + ModelCheckpoint, tf.data-style pipeline)."""
import os

import cloud_b200 as tfc

keras = tfc.keras

(x_train, y_train), (x_test, y_test) = keras.datasets.mnist.load_data()   # synthetic on this box
x_train = x_train[..., None][:8192]
y_train = y_train[:8192]

BATCH = 64


def scale(image, label):
    return image.astype("float32") / 255.0, label


train = tfc.data.Dataset.from_tensor_slices((x_train, y_train)).map(scale).cache() \
    .shuffle(10000).batch(BATCH)

model = keras.Sequential([
    keras.layers.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
    keras.layers.MaxPooling2D(),
    keras.layers.Flatten(),
    keras.layers.Dense(64, activation="relu"),
    keras.layers.Dense(10, activation="softmax"),
])
model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(),
              metrics=["accuracy"])


def decay(epoch):
    return 1e-3 if epoch < 3 else (1e-4 if epoch < 7 else 1e-5)


ckpt_dir = os.environ.get("CHECKPOINT_DIR", "./training_checkpoints")
callbacks = [keras.callbacks.LearningRateScheduler(decay),
             keras.callbacks.ModelCheckpoint(os.path.join(ckpt_dir, "ckpt_{epoch}"),
                                             save_weights_only=True)]
model.fit(train, epochs=2, callbacks=callbacks, verbose=2)
