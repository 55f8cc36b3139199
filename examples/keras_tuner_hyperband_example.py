"""Hyperband and Bayesian optimisation with the KerasTuner-shaped tuners of `cloud_b200.tuner`
(the engine the reference's CloudTuner builds on; these two searches need no optimizer service).

    python keras_tuner_hyperband_example.py [--samples 2000] [--max-epochs 9] [--directory ./tuner_demo]
"""
import argparse

import cloud_b200 as tfc
from cloud_b200 import tuner as tuners

keras = tfc.keras


def build_model(hp):
    model = keras.Sequential()
    model.add(keras.layers.Flatten(input_shape=(28, 28)))
    for i in range(hp.Int("num_layers", 1, 3)):
        model.add(keras.layers.Dense(hp.Int("units_%d" % i, 32, 256, step=32), activation="relu"))
    model.add(keras.layers.Dropout(hp.Float("dropout", 0.0, 0.5, step=0.1)))
    model.add(keras.layers.Dense(10, activation="softmax"))
    model.compile(optimizer=keras.optimizers.Adam(hp.Float("learning_rate", 1e-4, 1e-2, sampling="log")),
                  loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    return model


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=4000)
    ap.add_argument("--max-epochs", type=int, default=9)
    ap.add_argument("--bayesian-trials", type=int, default=8)
    ap.add_argument("--directory", default="./tuner_demo")
    args = ap.parse_args()
    (x, y), (vx, vy) = keras.datasets.mnist.load_data()
    x, y = x[:args.samples].astype("float32") / 255, y[:args.samples]
    vx, vy = vx[:args.samples // 4].astype("float32") / 255, vy[:args.samples // 4]

    hyperband = tuners.Hyperband(build_model, objective="val_accuracy", max_epochs=args.max_epochs, factor=3,
                                 directory=args.directory, project_name="hyperband", overwrite=True)
    hyperband.search_space_summary()
    hyperband.search(x=x, y=y, validation_data=(vx, vy), batch_size=64, verbose=0)
    hyperband.results_summary(3)
    print("Hyperband ran {} trials; best hyperparameters: {}".format(
        len(hyperband.oracle.trials),
        {k: v for k, v in hyperband.get_best_hyperparameters(1)[0].values.items() if not k.startswith("tuner/")}))

    bayes = tuners.BayesianOptimization(build_model, objective="val_accuracy", max_trials=args.bayesian_trials,
                                        directory=args.directory, project_name="bayesian", overwrite=True)
    bayes.search(x=x, y=y, epochs=3, validation_data=(vx, vy), batch_size=64, verbose=0)
    best = bayes.get_best_models(1)[0]
    print("Bayesian optimisation: best validation accuracy {:.4f}".format(best.evaluate(vx, vy, verbose=0)[1]))


if __name__ == "__main__":
    main()
