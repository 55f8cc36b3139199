"""Model definition of the multi-file example (reference:
core/tests/examples/multi_file_example/create_model.py): imported by train_model.py, so the whole
directory has to travel with the job."""
import cloud_b200 as tfc

keras = tfc.keras


def create_keras_model():
    model = keras.Sequential([
        keras.layers.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
        keras.layers.MaxPooling2D(),
        keras.layers.Flatten(),
        keras.layers.Dense(64, activation="relu"),
        keras.layers.Dense(10, activation="softmax"),
    ])
    model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(),
                  metrics=["accuracy"])
    return model
