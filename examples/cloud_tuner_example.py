"""CloudTuner search, single tuner and distributed (reference:
tuner/tests/integration/tuner_integration_test.py and
tuner/tests/examples/ai_platform_optimizer_tuner.ipynb)."""
import os

import cloud_b200 as tfc
from cloud_b200.tuner import HyperParameters, TrialScheduler

keras = tfc.keras
STUDY = "mnist_mlp_demo"


def build_model(hp):
    model = keras.Sequential()
    model.add(keras.layers.Flatten(input_shape=(28, 28)))
    # CloudTuner needs the search space up-front: read values with hp.get(...)
    for _ in range(hp.get("num_layers")):
        model.add(keras.layers.Dense(64, activation="relu"))
    model.add(keras.layers.Dense(10, activation="softmax"))
    model.compile(optimizer=keras.optimizers.Adam(lr=hp.get("learning_rate")),
                  loss="sparse_categorical_crossentropy", metrics=["acc"])
    return model


def search_space():
    hps = HyperParameters()
    hps.Float("learning_rate", min_value=1e-4, max_value=1e-2, sampling="log")
    hps.Int("num_layers", 2, 10)
    return hps


def tuner_worker(tuner_id, directory="./tuner_demo"):
    (x, y), (vx, vy) = keras.datasets.mnist.load_data()
    x, vx = x[:10000].astype("float32") / 255, vx[:2000].astype("float32") / 255
    tuner = tfc.CloudTuner(build_model, project_id=os.environ.get("PROJECT_ID", "local"),
                           region=os.environ.get("REGION", "local"), objective="acc",
                           hyperparameters=search_space(), max_trials=8, study_id=STUDY,
                           directory=os.path.join(directory, tuner_id))
    tuner.tuner_id = tuner_id
    tuner.search(x=x, y=y[:10000], epochs=2, steps_per_epoch=20, validation_data=(vx, vy[:2000]), verbose=0)
    return tuner


if __name__ == "__main__":
    # one worker per GPU, all sharing the study
    records = TrialScheduler().run(lambda tid: tuner_worker(tid) and None, num_workers=4,
                                   study=(os.environ.get("PROJECT_ID", "local"),
                                          os.environ.get("REGION", "local"), "CloudTuner_study_" + STUDY))
    print(records)
    t = tuner_worker("summary")
    t.results_summary(3)
