"""Entry point of the multi-file example (reference:
core/tests/examples/multi_file_example/train_model.py): builds the model through a sibling module
and trains it on an MNIST-shaped tf.data-style pipeline."""
import numpy as np

import cloud_b200 as tfc

from create_model import create_keras_model      # sibling module: the directory travels with the job

keras = tfc.keras
BATCH, SHUFFLE_BUFFER = 64, 10000


def pipeline(images, labels, training):
    """uint8 images -> float32 in [0, 1], cached, (shuffled,) batched."""
    ds = tfc.data.Dataset.from_tensor_slices((images[..., None], labels))
    ds = ds.map(lambda im, lb: (np.asarray(im, dtype="float32") / 255, lb))
    if training:
        ds = ds.cache().shuffle(SHUFFLE_BUFFER)
    return ds.batch(BATCH)


(train_x, train_y), (test_x, test_y) = keras.datasets.mnist.load_data()
model = create_keras_model()
model.fit(pipeline(train_x[:2048], train_y[:2048], True), epochs=2 if tfc.remote() else 1, verbose=2)
print("eval:", model.evaluate(pipeline(test_x[:512], test_y[:512], False), verbose=0))
