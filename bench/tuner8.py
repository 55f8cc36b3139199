"""BASELINE.json config 4: CloudTuner-style random search over MNIST-MLP hyper-parameters, 8
parallel trials packed across the GPUs of one box by the trial scheduler (one tuner worker per
GPU, shared study in the local optimizer), against the same 8 trials run one after another on one
GPU.  Wall-clock seconds for each arm.

    python bench/tuner8.py [--trials 8] [--model mlp|resnet18]
"""
import argparse
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker(tuner_id, directory, study_id, max_trials, model_kind, epochs):
    import numpy as np
    import cloud_b200 as tfc
    from cloud_b200.tuner import HyperParameters
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/gpu:0"))
    rng = np.random.RandomState(0)
    if model_kind == "mlp":
        x = rng.rand(60000, 28, 28).astype("float32")      # MNIST-sized
        y = rng.randint(0, 10, 60000)

        def build(hp):
            m = keras.Sequential()
            m.add(keras.layers.Flatten(input_shape=(28, 28)))
            for _ in range(hp.get("num_layers")):
                m.add(keras.layers.Dense(64, activation="relu"))
            m.add(keras.layers.Dense(10, activation="softmax"))
            m.compile(optimizer=keras.optimizers.Adam(lr=hp.get("learning_rate")),
                      loss="sparse_categorical_crossentropy", metrics=["acc"])
            return m
        fit = dict(x=x, y=y, epochs=epochs, batch_size=128, verbose=0)
    else:
        x = rng.randint(0, 256, (2048, 32, 32, 3)).astype("float32") / 255.0
        y = rng.randint(0, 10, 2048)

        def build(hp):
            m = keras.applications.ResNet18(classes=10, input_shape=(32, 32, 3))
            m.compile(optimizer=keras.optimizers.Adam(lr=hp.get("learning_rate")),
                      loss="sparse_categorical_crossentropy", metrics=["acc"])
            return m
        fit = dict(x=x, y=y, epochs=epochs, batch_size=128, verbose=0)
    hps = HyperParameters()
    hps.Float("learning_rate", 1e-4, 1e-2, sampling="log")
    hps.Int("num_layers", 2, 10)
    t = tfc.CloudTuner(build, project_id="bench", region="local", objective="acc", hyperparameters=hps,
                       max_trials=max_trials, study_id=study_id, directory=os.path.join(directory, tuner_id))
    t.tuner_id = tuner_id
    t.search(**fit)
    return {"tuner": tuner_id, "trials": len(t.oracle.trials)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=8)
    ap.add_argument("--model", default="mlp")
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--skip-sequential", action="store_true")
    args = ap.parse_args()
    from cloud_b200.tuner import TrialScheduler
    root = tempfile.mkdtemp(prefix="tuner8_")
    os.environ["TFC_JOB_ROOT"] = os.path.join(root, "jobs")
    out = {"bench": "tuner8", "model": args.model, "trials": args.trials}
    sched = TrialScheduler()
    n = max(sched.num_gpus, 1)
    t0 = time.perf_counter()
    res = sched.run(worker, args=(os.path.join(root, "packed"), "packed", args.trials, args.model, args.epochs),
                    num_workers=min(n, args.trials), study=("bench", "local", "CloudTuner_study_packed"))
    out["packed_seconds"] = time.perf_counter() - t0
    out["packed_workers"] = len(res)
    out["packed_exit_codes"] = [r["exit_code"] for r in res]
    if not args.skip_sequential:
        seq = TrialScheduler(num_gpus=min(n, 1))
        t0 = time.perf_counter()
        res = seq.run(worker, args=(os.path.join(root, "seq"), "seq", args.trials, args.model, args.epochs),
                      num_workers=1, study=("bench", "local", "CloudTuner_study_seq"))
        out["sequential_seconds"] = time.perf_counter() - t0
        out["sequential_exit_codes"] = [r["exit_code"] for r in res]
        out["speedup"] = out["sequential_seconds"] / out["packed_seconds"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
