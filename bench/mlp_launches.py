"""Kernel list of ONE training step of the README MLP (eager, no CUDA graph) for `ncu --metrics
gpu__time_duration.sum`: prints nothing itself; bench/summarize_launches.py reads the capture."""
import os
import sys

os.environ.setdefault("TFC_CUDA_GRAPH", "0")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import cloud_b200 as tfc  # noqa: E402

keras = tfc.keras
tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
model = keras.Sequential([keras.layers.Dense(512, activation="relu", input_shape=(784,)),
                          keras.layers.Dropout(0.2), keras.layers.Dense(10, activation="softmax")])
model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(), metrics=["accuracy"])
x = torch.rand(128, 784, device="cuda")
y = torch.randint(0, 10, (128,), device="cuda")
model._ensure_engine(x)
model.train(True)
xi = model._to_internal(x)
for _ in range(4):
    model._run_step(xi, y)
torch.cuda.synchronize()
