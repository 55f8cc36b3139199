"""Entry point of the multi-file example (reference:
core/tests/examples/multi_file_example/train_model.py): builds the model through a sibling module
and trains it on an MNIST-shaped tf.data-style pipeline."""
import numpy as np

import cloud_b200 as tfc

from create_model import create_keras_model

keras = tfc.keras
(x_train, y_train), (x_test, y_test) = keras.datasets.mnist.load_data()
x_train, y_train = x_train[:2048, ..., None], y_train[:2048]
x_test, y_test = x_test[:512, ..., None], y_test[:512]

BUFFER_SIZE = 10000
BATCH_SIZE = 64


def scale(image, label):
    return np.asarray(image, dtype="float32") / 255, label


train_dataset = tfc.data.Dataset.from_tensor_slices((x_train, y_train)).map(scale).cache()
train_dataset = train_dataset.shuffle(BUFFER_SIZE).batch(BATCH_SIZE)
eval_dataset = tfc.data.Dataset.from_tensor_slices((x_test, y_test)).map(scale).batch(BATCH_SIZE)

model = create_keras_model()
epochs = 2 if tfc.remote() else 1
model.fit(train_dataset, epochs=epochs, verbose=2)
print("eval:", model.evaluate(eval_dataset, verbose=0))
