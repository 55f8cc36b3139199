"""Model definition of the multi-file example (reference:
core/tests/examples/multi_file_example/create_model.py): imported by train_model.py, so the whole
directory has to travel with the job."""
import cloud_b200 as tfc

keras = tfc.keras


def create_keras_model(num_classes=10, image_shape=(28, 28, 1)):
    """Small conv net for MNIST-shaped images, compiled and ready for ``fit``."""
    L = keras.layers
    net = keras.Sequential(name="mnist_cnn")
    net.add(L.Conv2D(32, 3, activation="relu", input_shape=image_shape, name="features"))
    net.add(L.MaxPooling2D(name="pool"))
    net.add(L.Flatten(name="flatten"))
    for units, act, name in ((64, "relu", "hidden"), (num_classes, "softmax", "probabilities")):
        net.add(L.Dense(units, activation=act, name=name))
    net.compile(optimizer=keras.optimizers.Adam(), loss="sparse_categorical_crossentropy",
                metrics=["accuracy"])
    return net
