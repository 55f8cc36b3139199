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
    t_start = time.time()
    marks = {}
    import numpy as np
    import cloud_b200 as tfc
    from cloud_b200.tuner import HyperParameters
    keras = tfc.keras
    marks["import_s"] = time.time() - t_start
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
    marks["data_s"] = time.time() - t_start
    hps = HyperParameters()
    hps.Float("learning_rate", 1e-4, 1e-2, sampling="log")
    hps.Int("num_layers", 2, 10)
    t = tfc.CloudTuner(build, project_id="bench", region="local", objective="acc", hyperparameters=hps,
                       max_trials=max_trials, study_id=study_id, directory=os.path.join(directory, tuner_id))
    t.tuner_id = tuner_id
    marks["tuner_s"] = time.time() - t_start
    t.search(**fit)
    marks["done_s"] = time.time() - t_start
    return {"tuner": tuner_id, "trials": len(t.oracle.trials), "marks": {k: round(v, 3) for k, v in marks.items()}}


class _UtilSampler:
    """Average SM utilisation of every GPU of the box while a run is in flight (NVML, 50 ms)."""

    def __init__(self):
        import threading
        self.samples, self._stop, self.ok = [], threading.Event(), False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.handles = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(pynvml.nvmlDeviceGetCount())]
            self.ok = bool(self.handles)
        except Exception:
            pass
        self.thread = threading.Thread(target=self._loop, daemon=True)

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append([self.nv.nvmlDeviceGetUtilizationRates(h).gpu for h in self.handles])
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.ok:
            self.thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self.ok:
            self.thread.join(timeout=2)

    def per_gpu(self):
        if not self.samples:
            return None
        n = len(self.samples)
        return [round(sum(s[i] for s in self.samples) / n, 1) for i in range(len(self.samples[0]))]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=8)
    ap.add_argument("--model", default="mlp")
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--skip-sequential", action="store_true")
    ap.add_argument("--cold", action="store_true", help="one cold process per worker (no pre-warmed pool)")
    args = ap.parse_args()
    from cloud_b200.tuner import TrialScheduler, WorkerPool
    root = tempfile.mkdtemp(prefix="tuner8_")
    os.environ["TFC_JOB_ROOT"] = os.path.join(root, "jobs")
    out = {"bench": "tuner8", "model": args.model, "trials": args.trials, "epochs": args.epochs,
           "workers": "cold processes" if args.cold else "pre-warmed pool"}
    sched = TrialScheduler()
    n = max(sched.num_gpus, 1)
    workers = min(n, args.trials)
    pool = None
    if not args.cold:
        pool = WorkerPool(num_workers=workers)
        out["pool_start_seconds"] = pool.start_seconds           # paid once, all workers in parallel
    try:
        with _UtilSampler() as util:
            t0 = time.perf_counter()
            res = sched.run(worker, args=(os.path.join(root, "packed"), "packed", args.trials, args.model,
                                          args.epochs),
                            num_workers=workers, study=("bench", "local", "CloudTuner_study_packed"), pool=pool)
            out["packed_seconds"] = time.perf_counter() - t0
        out["packed_workers"] = len(res)
        out["packed_worker_marks"] = [(r.get("result") or {}).get("marks") for r in res]
        out["packed_worker_seconds"] = [round(r["seconds"], 2) for r in res]
        out["packed_exit_codes"] = [r["exit_code"] for r in res]
        out["packed_gpu_util_percent"] = util.per_gpu()
        if not args.skip_sequential:
            # the same trials one after another on ONE worker (warm as well when the pool is used)
            seq = TrialScheduler(num_gpus=min(n, 1))
            seq_pool = None
            if pool is not None:
                class _One:                    # the first slot of the pool only
                    slots, procs = pool.slots[:1], pool.procs
                    submit = staticmethod(pool.submit)
                seq_pool = _One()
            t0 = time.perf_counter()
            res = seq.run(worker, args=(os.path.join(root, "seq"), "seq", args.trials, args.model, args.epochs),
                          num_workers=1, study=("bench", "local", "CloudTuner_study_seq"), pool=seq_pool)
            out["sequential_seconds"] = time.perf_counter() - t0
            out["sequential_exit_codes"] = [r["exit_code"] for r in res]
            out["speedup"] = out["sequential_seconds"] / out["packed_seconds"]
            if pool is not None:
                out["speedup_incl_pool_start"] = (out["sequential_seconds"] + out["pool_start_seconds"]) / \
                    (out["packed_seconds"] + out["pool_start_seconds"])
    finally:
        if pool is not None:
            pool.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
