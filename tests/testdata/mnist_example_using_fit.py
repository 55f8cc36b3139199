"""Test entry point: a small MNIST-shaped model trained with model.fit (counterpart of the
reference's core/tests/testdata/mnist_example_using_fit.py, written against cloud_b200.keras)."""
import numpy as np
import cloud_b200 as tfc

keras = tfc.keras

rng = np.random.RandomState(0)
x = rng.rand(256, 28, 28, 1).astype("float32")
y = rng.randint(0, 10, 256)

model = keras.Sequential([
    keras.layers.Conv2D(8, 3, activation="relu", input_shape=(28, 28, 1)),
    keras.layers.MaxPooling2D(),
    keras.layers.Flatten(),
    keras.layers.Dense(16, activation="relu"),
    keras.layers.Dense(10, activation="softmax"),
])
model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(),
              metrics=["accuracy"])
model.fit(x, y, epochs=1, batch_size=64, verbose=0)
print("TRAINED", tfc.remote())
