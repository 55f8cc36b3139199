"""`tfc.run()` inside the training script itself (reference:
core/tests/examples/call_run_within_script_with_keras_fit.py): the first execution submits the
script as a job and exits; inside the job `tfc.run()` is a no-op and `tfc.remote()` is true.
ResNet-50 transfer-learning shape of the reference example, on synthetic images."""
import numpy as np

import cloud_b200 as tfc

keras = tfc.keras

tfc.run(chief_config=tfc.COMMON_MACHINE_CONFIGS["B200_2X"], stream_logs=True)

NUM_CLASSES, IMG = 120, 224
rng = np.random.RandomState(0)
images = rng.randint(0, 256, (256, IMG, IMG, 3)).astype("uint8")
labels = rng.randint(0, NUM_CLASSES, 256)

inputs = keras.Input(shape=(IMG, IMG, 3))
x = keras.layers.ImageNormalization()(inputs)
base_model = keras.applications.ResNet50(weights=None, include_top=False, input_tensor=x)
x = keras.layers.GlobalAveragePooling2D()(base_model.output)
x = keras.layers.Dropout(0.5)(x)
predictions = keras.layers.Dense(NUM_CLASSES, activation="softmax")(x)
model = keras.Model(inputs=inputs, outputs=predictions)

epochs = 2 if tfc.remote() else 1
callbacks = [keras.callbacks.TensorBoard(log_dir="./logs", histogram_freq=0),
             keras.callbacks.EarlyStopping(monitor="loss", patience=3)]
model.compile(optimizer=keras.optimizers.Adam(learning_rate=1e-2),
              loss="sparse_categorical_crossentropy", metrics=["accuracy"])
model.fit(images, labels, batch_size=64, epochs=epochs, callbacks=callbacks, verbose=2)
if tfc.remote():
    model.save("./resnet50_dogs_model")
