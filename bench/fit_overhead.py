"""Where does model.fit's host time go?  Runs the e2e arm of bench.py for several step counts and
prints fixed vs per-step cost plus a cProfile of the training thread."""
import cProfile
import io
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench as B  # noqa: E402
import cloud_b200 as tfc  # noqa: E402

keras = tfc.keras
tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
model = B.build_model(keras, 1000)
Bsz = 64
N = Bsz * 120
host_x = torch.randint(0, 256, (N, 224, 224, 3), dtype=torch.uint8).pin_memory()
host_y = torch.randint(0, 1000, (N,), dtype=torch.int64).pin_memory()
seen = []


class ReadLoss(keras.callbacks.Callback):
    def on_train_batch_end(self, batch, logs=None):
        seen.append(logs["loss"])


model.fit(host_x[:Bsz * 8], host_y[:Bsz * 8], batch_size=Bsz, epochs=1, shuffle=False, verbose=0,
          callbacks=[ReadLoss()])
torch.cuda.synchronize()
for steps in (10, 30, 60, 120):
    n = Bsz * steps
    t0 = time.perf_counter()
    model.fit(host_x[:n], host_y[:n], batch_size=Bsz, epochs=1, shuffle=False, verbose=0,
              callbacks=[ReadLoss()])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"steps {steps:4d}: total {dt * 1e3:8.2f} ms  per-step {dt / steps * 1e3:6.3f} ms", flush=True)
# no per-batch callback: how much does the lagged loss read-back cost?
t0 = time.perf_counter()
model.fit(host_x[:Bsz * 60], host_y[:Bsz * 60], batch_size=Bsz, epochs=1, shuffle=False, verbose=0)
torch.cuda.synchronize()
print(f"steps   60 without batch callback: per-step {(time.perf_counter() - t0) / 60 * 1e3:6.3f} ms", flush=True)
pr = cProfile.Profile()
pr.enable()
model.fit(host_x[:Bsz * 60], host_y[:Bsz * 60], batch_size=Bsz, epochs=1, shuffle=False, verbose=0,
          callbacks=[ReadLoss()])
torch.cuda.synchronize()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28)
print(s.getvalue()[:6000])
