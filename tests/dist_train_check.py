"""Worker script for distributed-equivalence tests (run under torch.distributed.run).

Trains the README MLP for a few steps under MirroredStrategy on WORLD_SIZE replicas and checks
that the weights match a single-replica run on the same global batches (sum of per-replica
gradients of a loss scaled by 1/global_batch == the large-batch gradient).
Writes "OK" / raises.  Works on CPU (gloo) and GPU (peer-memory kernels).
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cloud_b200 as tfc  # noqa: E402
from cloud_b200.keras import layers  # noqa: E402

keras = tfc.keras


def build(seed):
    torch.manual_seed(seed)
    keras.backend.clear_session()
    m = keras.Sequential([layers.Dense(64, activation="relu", input_shape=(32,)),
                          layers.Dense(16, activation="relu"),
                          layers.Dense(4, activation="softmax")])
    # CHECK_RATIO_METRIC: a metric whose denominator is a sum of its own (recall) next to accuracy
    metrics = ["accuracy", "recall"] if os.environ.get("CHECK_RATIO_METRIC") == "1" else ["accuracy"]
    m.compile(optimizer=keras.optimizers.Adam(1e-2), loss="sparse_categorical_crossentropy",
              metrics=metrics, run_eagerly=bool(int(os.environ.get("CHECK_EAGER", "1"))))
    return m


def main():
    use_gpu = torch.cuda.is_available() and os.environ.get("CHECK_DEVICE", "auto") != "cpu"
    rng = np.random.RandomState(0)
    x = rng.rand(256, 32).astype("float32")
    y = (x[:, :4].argmax(1)).astype("int64")

    strategy = tfc.distribute.MirroredStrategy()
    world, rank = strategy.num_replicas_in_sync, strategy.replica_id
    if not use_gpu:
        assert strategy.device.type == "cpu"
    with strategy.scope():
        m = build(1 + rank)            # different init per replica: the broadcast must fix it
        h = m.fit(x, y, batch_size=64, epochs=2, shuffle=False, verbose=0)
    w_dist = [w.copy() for w in m.get_weights()]
    backend = m._engine.comm.backend

    # single-replica oracle on the same global batches, same initial weights as replica 0
    single = tfc.distribute.OneDeviceStrategy(strategy.device)
    with single.scope():
        ref = build(1)
        ref._strategy = single
        h_ref = ref.fit(x, y, batch_size=64, epochs=2, shuffle=False, verbose=0)
    w_ref = ref.get_weights()
    tol = 5e-2 if use_gpu else 1e-5       # bf16 compute on GPU
    worst = max(float(np.abs(a - b).max()) for a, b in zip(w_dist, w_ref))
    assert worst < tol, f"rank {rank}: weights differ from the single-replica run by {worst}"
    assert abs(h.history["loss"][-1] - h_ref.history["loss"][-1]) < (5e-2 if use_gpu else 1e-5), \
        (h.history, h_ref.history)
    if "recall" in h.history:       # numerators and denominators were summed over the replicas
        assert abs(h.history["recall"][-1] - h_ref.history["recall"][-1]) < (5e-2 if use_gpu else 1e-6), \
            (h.history["recall"], h_ref.history["recall"])
    # replicas agree exactly
    t = torch.tensor(np.concatenate([w.reshape(-1) for w in w_dist]))
    lst = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        import torch.distributed as dist
        dev_t = t.to(strategy.device)
        gathered = [torch.zeros_like(dev_t) for _ in range(world)]
        dist.all_gather(gathered, dev_t)
        for g in gathered[1:]:
            assert torch.equal(g, gathered[0]), "replicas diverged"
    # chief-only save + reload
    out_dir = os.environ.get("CHECK_OUT")
    if out_dir:
        m.save(os.path.join(out_dir, "model"))
        strategy.barrier()
        if rank == 0:
            again = keras.models.load_model(os.path.join(out_dir, "model"))
            for a, b in zip(again.get_weights(), w_dist):
                assert np.allclose(a, b, atol=1e-6)
    if rank == 0:
        print(f"OK world={world} backend={backend} worst_diff={worst:.3g}")


if __name__ == "__main__":
    main()
