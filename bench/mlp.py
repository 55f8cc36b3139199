"""README MNIST MLP (784-512-10, Adam, global batch 128; README.md:66-81) through the public
Keras-shaped API: steps/s and s/epoch on synthetic 60000x784 data.  BASELINE.json configs 1 and 2.

    python bench/mlp.py [--cpu] [--epochs 2]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/mlp.py
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true", help="OneDeviceStrategy on the CPU (config 1)")
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch", type=int, default=128, help="global batch")
    ap.add_argument("--samples", type=int, default=60000)
    ap.add_argument("--steps-per-execution", type=int, default=1,
                    help="Keras compile(steps_per_execution=K): K steps per CUDA graph")
    args = ap.parse_args()
    if args.cpu:
        os.environ["CUDA_VISIBLE_DEVICES"] = ""
    import torch
    import cloud_b200 as tfc
    keras = tfc.keras
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        strategy = tfc.distribute.MirroredStrategy()
    else:
        strategy = tfc.distribute.OneDeviceStrategy("/gpu:0")
    tfc.distribute.experimental_set_strategy(strategy)
    rng = np.random.RandomState(0)
    x = rng.rand(args.samples, 784).astype("float32")
    y = rng.randint(0, 10, args.samples).astype("int64")
    model = keras.Sequential([keras.layers.Dense(512, activation="relu", input_shape=(784,)),
                              keras.layers.Dropout(0.2),
                              keras.layers.Dense(10, activation="softmax")])
    model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(),
                  metrics=["accuracy"], steps_per_execution=args.steps_per_execution)
    if strategy.device.type == "cuda":
        x_t, y_t = torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory()
    else:
        x_t, y_t = x, y
    model.fit(x_t, y_t, epochs=1, batch_size=args.batch, verbose=0)           # warm-up epoch
    if strategy.device.type == "cuda":
        torch.cuda.synchronize()
    strategy.barrier()
    t0 = time.perf_counter()
    h = model.fit(x_t, y_t, epochs=args.epochs, batch_size=args.batch, verbose=0)
    if strategy.device.type == "cuda":
        torch.cuda.synchronize()
    strategy.barrier()
    dt = time.perf_counter() - t0
    steps = args.epochs * -(-args.samples // args.batch)
    if strategy.replica_id == 0:
        print(json.dumps({"bench": "mnist_mlp", "device": strategy.device.type, "n_replicas": world,
                          "global_batch": args.batch, "steps_per_execution": args.steps_per_execution,
                          "steps_per_s": steps / dt,
                          "s_per_epoch": dt / args.epochs, "samples_per_s": steps * args.batch / dt,
                          "final_loss": h.history["loss"][-1],
                          "backend": model._engine.comm.backend,
                          "cuda_graph": bool((model._graph and model._graph.get("graph") is not None) or
                                             (model._graph_k and model._graph_k.get("graph") is not None))}))


if __name__ == "__main__":
    main()
