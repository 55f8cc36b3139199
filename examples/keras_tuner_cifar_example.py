"""Random search over a small CIFAR-10 conv net with the (re-created) KerasTuner engine
(reference: core/tests/testdata/keras_tuner_cifar_example.py): hyperparameters declared inside
``build_model`` (including conditional ones), augmentation layers inside a nested Sequential,
RMSprop on an ExponentialDecay schedule, tf.data-shaped input pipeline, early stopping, best model
saved at the end.  Synthetic CIFAR-shaped data (no network on this box); sizes kept small so the
script also finishes on the CPU preset:

    python keras_tuner_cifar_example.py --path /tmp/best_cifar_model [--max-trials 3]
"""
import argparse

import numpy as np

import cloud_b200 as tfc
from cloud_b200 import tuner as tuners

keras = tfc.keras
preprocessing = keras.layers.experimental.preprocessing


ACTIVATIONS = ["relu", "elu"]


def _augmentation():
    """Random flip / rotation / width / height, as a nested Sequential used like a layer."""
    return keras.Sequential([preprocessing.RandomFlip(), preprocessing.RandomRotation(0.1),
                             preprocessing.RandomWidth(0.1), preprocessing.RandomHeight(0.1)])


def _conv(hp, tag, filters):
    act = hp.Choice("conv_activation_" + tag, values=ACTIVATIONS, default="relu")
    return keras.layers.Conv2D(filters, (3, 3), padding="same", activation=act)


def build_model(hp):
    """The hypermodel: every hyperparameter is declared here, the deeper ones only when the block
    they belong to exists (conditional search space)."""
    image = keras.Input(shape=(32, 32, 3))
    net = _conv(hp, "0", 32)(_augmentation()(image))
    n_blocks = hp.Int("num_blocks", 1, 3)
    for blk in range(n_blocks):
        for cv in range(hp.Int("num_conv_%d" % blk, 1, 3)):
            width = hp.Int("filter_%d_%d" % (blk, cv), 16, 256, step=16)
            net = _conv(hp, "%d_%d" % (blk, cv), width)(net)
        net = keras.layers.MaxPooling2D(pool_size=(2, 2))(net)
    drop = hp.Float("dropout_1", min_value=0.0, max_value=0.9, default=0.25, step=0.05)
    net = keras.layers.GlobalMaxPooling2D()(keras.layers.Dropout(rate=drop)(net))
    net = keras.layers.Dense(512, activation="relu")(net)
    model = keras.Model(image, keras.layers.Dense(10, activation="softmax")(net))

    schedule = keras.optimizers.schedules.ExponentialDecay(
        initial_learning_rate=hp.Choice("initial_learning_rate", [1e-1, 1e-2, 1e-3]),
        decay_steps=100000, decay_rate=hp.Choice("decay_rate", [0.5, 0.75, 0.95]), staircase=True)
    model.compile(optimizer=keras.optimizers.RMSprop(learning_rate=schedule),
                  loss="sparse_categorical_crossentropy", metrics=["sparse_categorical_accuracy"])
    return model


def main():
    parser = argparse.ArgumentParser(description="Keras model save path")
    parser.add_argument("--path", required=True, type=str, help="Keras model save path")
    parser.add_argument("--max-trials", type=int, default=3)
    parser.add_argument("--samples", type=int, default=512)
    parser.add_argument("--directory", type=str, default="test_dir")
    args = parser.parse_args()

    tuner = tuners.RandomSearch(build_model, objective="val_sparse_categorical_accuracy",
                                max_trials=args.max_trials, executions_per_trial=1,
                                directory=args.directory)
    tuner.search_space_summary()

    (x_train, y_train), (x_test, y_test) = keras.datasets.cifar10.load_data()
    x_train, y_train = x_train[:args.samples], y_train[:args.samples]
    x_test, y_test = x_test[:args.samples // 4], y_test[:args.samples // 4]
    train_dataset = tfc.data.Dataset.from_tensor_slices((x_train, y_train))
    test_dataset = tfc.data.Dataset.from_tensor_slices((x_test, y_test))
    BUFFER_SIZE, BATCH_SIZE = 10000, 64

    def scale(image, label):
        return np.asarray(image, dtype="float32") / 255, label

    train_dataset = train_dataset.map(scale).cache().shuffle(BUFFER_SIZE).batch(BATCH_SIZE)
    test_dataset = test_dataset.map(scale).batch(BATCH_SIZE)

    tuner.search(train_dataset, epochs=2, validation_data=test_dataset,
                 callbacks=[keras.callbacks.EarlyStopping(monitor="val_loss", patience=3, mode="min")])
    print("Tuner results summary")
    tuner.results_summary()

    best_model = tuner.get_best_models(num_models=1)[0]
    scores = best_model.evaluate(x_test.astype("float32") / 255, y_test, verbose=1)
    print("Test loss:", scores[0])
    print("Test accuracy:", scores[1])
    print("Saving best model")
    best_model.save(args.path)


if __name__ == "__main__":
    main()
