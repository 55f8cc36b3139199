import os, sys, traceback
sys.path.insert(0, "/root/repo")
import torch, numpy as np
import cloud_b200 as tfc
keras = tfc.keras
tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
model = keras.Sequential([keras.layers.Dense(512, activation="relu", input_shape=(784,)),
                          keras.layers.Dropout(0.2), keras.layers.Dense(10, activation="softmax")])
model.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.Adam(), metrics=["accuracy"])
x = torch.rand(128, 784, device="cuda"); y = torch.randint(0, 10, (128,), device="cuda")
model._ensure_engine(x); model.train(True)
xi = model._to_internal(x)
for _ in range(3):
    model._train_step(xi, y)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        out = model._train_step(xi, y)
    print("capture ok")
except Exception:
    traceback.print_exc()
