"""tf.data-shaped input pipeline and the self-contained HDF5 reader/writer (C++ runtime library)."""
import struct

import numpy as np
import pytest

import cloud_b200 as tfc
from cloud_b200.io import hdf5

Dataset = tfc.data.Dataset


def test_dataset_batch_map_take_repeat_and_cardinality():
    ds = Dataset.range(10)
    assert ds.cardinality() == 10
    assert [int(v) for v in ds.as_numpy_iterator()] == list(range(10))
    b = Dataset.from_tensor_slices((np.arange(10, dtype="float32"), np.arange(10) * 2)).batch(4)
    batches = list(b.as_numpy_iterator())
    assert [len(x) for x, _ in batches] == [4, 4, 2] and b.cardinality() == 3
    assert np.array_equal(batches[2][1], [16, 18])
    b2 = Dataset.from_tensor_slices(np.arange(10)).batch(4, drop_remainder=True)
    assert [list(map(int, x)) for x in b2.as_numpy_iterator()] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    m = Dataset.range(5).map(lambda v: v * v).skip(1).take(3)
    assert [int(v) for v in m.as_numpy_iterator()] == [1, 4, 9]
    r = Dataset.range(3).repeat(2)
    assert [int(v) for v in r.as_numpy_iterator()] == [0, 1, 2, 0, 1, 2]
    assert Dataset.range(3).repeat().cardinality() == -1
    u = Dataset.from_tensor_slices(np.arange(6)).batch(3).unbatch()
    assert [int(v) for v in u.as_numpy_iterator()] == list(range(6))


def test_dataset_shuffle_is_a_seeded_permutation_and_prefetch_cache_are_transparent():
    base = Dataset.from_tensor_slices(np.arange(100))
    a = [int(v) for v in base.shuffle(100, seed=3).as_numpy_iterator()]
    b = [int(v) for v in base.shuffle(100, seed=3).as_numpy_iterator()]
    assert a == b and sorted(a) == list(range(100)) and a != list(range(100))
    c = base.cache().prefetch(Dataset.AUTOTUNE).batch(25)
    assert sum(len(x) for x in c.as_numpy_iterator()) == 100


def test_dataset_from_generator_and_from_tensors():
    def gen():
        for i in range(4):
            yield np.full((2,), i, dtype="float32"), i
    g = Dataset.from_generator(gen)
    out = list(g.as_numpy_iterator())
    assert len(out) == 4 and int(out[3][1]) == 3 and np.array_equal(out[2][0], [2, 2])
    t = Dataset.from_tensors((np.ones((3, 2)), np.zeros(3)))
    only = list(t.as_numpy_iterator())
    assert len(only) == 1 and only[0][0].shape == (3, 2)


def test_hdf5_groups_datasets_and_attributes_round_trip(tmp_path):
    path = str(tmp_path / "t.h5")
    w = np.random.RandomState(0).randn(7, 5).astype("float32")
    idx = np.arange(12, dtype="int64").reshape(3, 4)
    with hdf5.File(path, "w") as f:
        f.attrs["backend"] = b"cloud_b200"
        f.attrs["layer_names"] = [b"dense", b"bn"]
        f.attrs["version"] = 3
        g = f.create_group("dense")
        g.attrs["weight_names"] = [b"dense/kernel:0"]
        g.create_group("dense").create_dataset("kernel:0", data=w)
        f.require_group("bn").create_dataset("steps", data=idx)
        f.create_dataset("scalar", data=np.float64(2.5))
    with open(path, "rb") as fh:
        assert fh.read(8) == b"\x89HDF\r\n\x1a\n"              # the HDF5 signature
    with hdf5.File(path, "r") as f:
        assert sorted(f.keys()) == ["bn", "dense", "scalar"]
        assert f.attrs["backend"] in (b"cloud_b200", "cloud_b200")
        assert [bytes(n) for n in f.attrs["layer_names"]] == [b"dense", b"bn"]
        assert int(f.attrs["version"]) == 3
        assert "weight_names" in f["dense"].attrs
        k = f["dense"]["dense"]["kernel:0"]
        assert k.shape == (7, 5) and np.array_equal(np.asarray(k), w)
        assert np.array_equal(np.asarray(f["bn/steps"]), idx)
        assert float(np.asarray(f["scalar"])) == 2.5
        assert "nope" not in f and "dense" in f


def test_hdf5_rejects_garbage_and_oversized_attributes(tmp_path):
    bad = tmp_path / "bad.h5"
    bad.write_bytes(b"not an hdf5 file" + struct.pack("<Q", 0) * 64)
    with pytest.raises(Exception):
        hdf5.File(str(bad), "r")
    with hdf5.File(str(tmp_path / "big.h5"), "w") as f:
        with pytest.raises(hdf5.AttributeTooLarge):
            f.attrs["model_config"] = b"x" * 70000               # > one 64 KB object-header message


# -- parallel map and prefetch (tf.data runs these in C++ threads; here: Python threads around numpy) --
def test_parallel_map_keeps_order_and_uses_several_threads():
    import threading
    import time
    from cloud_b200 import data
    seen = set()

    def slow_square(v):
        seen.add(threading.current_thread().name)
        time.sleep(0.01)
        return v * v
    ds = data.Dataset.range(40)
    t0 = time.perf_counter()
    serial = [int(v) for v in ds.map(slow_square)]
    t_serial = time.perf_counter() - t0
    seen.clear()
    t0 = time.perf_counter()
    parallel = [int(v) for v in ds.map(slow_square, num_parallel_calls=4)]
    t_parallel = time.perf_counter() - t0
    assert parallel == serial == [i * i for i in range(40)]
    assert len(seen) >= 2 and all(n.startswith("tfc-data-map") for n in seen)
    assert t_parallel < 0.6 * t_serial
    auto = ds.map(slow_square, num_parallel_calls=data.experimental.AUTOTUNE).batch(8)
    assert [b.tolist() for b in auto][0] == [0, 1, 4, 9, 16, 25, 36, 49] and auto.cardinality() == 5
    # a consumer that stops early does not leave the pool hanging
    assert [int(v) for v in ds.map(slow_square, num_parallel_calls=4).take(3)] == [0, 1, 4]


def test_prefetch_overlaps_producer_and_consumer_and_forwards_errors():
    import time
    from cloud_b200 import data

    def slow(v):
        time.sleep(0.01)
        return v
    ds = data.Dataset.range(30).map(slow)
    t0 = time.perf_counter()
    for _ in ds:
        time.sleep(0.01)                       # the "training step"
    t_plain = time.perf_counter() - t0
    t0 = time.perf_counter()
    got = []
    for v in ds.prefetch(4):
        time.sleep(0.01)
        got.append(int(v))
    t_prefetch = time.perf_counter() - t0
    assert got == list(range(30)) and t_prefetch < 0.75 * t_plain
    assert ds.prefetch(data.AUTOTUNE).cardinality() == 30
    assert [int(v) for v in ds.prefetch(2).take(2)] == [0, 1]          # early stop: the thread lets go

    def boom(v):
        if v == 5:
            raise ValueError("bad element")
        return v
    with pytest.raises(ValueError, match="bad element"):
        list(data.Dataset.range(10).map(boom).prefetch(2))
    # two epochs over the same dataset object
    twice = data.Dataset.range(6).shuffle(6, seed=1).batch(2).prefetch(2)
    assert sorted(sum((b.tolist() for b in twice), [])) == sorted(sum((b.tolist() for b in twice), [])) == list(range(6))


def test_zip_concatenate_enumerate_filter_reduce():
    from cloud_b200 import data
    a, b = data.Dataset.range(5), data.Dataset.from_tensor_slices(np.arange(10, 17))
    assert [(int(x), int(y)) for x, y in data.Dataset.zip((a, b))] == [(0, 10), (1, 11), (2, 12), (3, 13), (4, 14)]
    assert [int(v) for v in a.concatenate(b.take(2))] == [0, 1, 2, 3, 4, 10, 11]
    assert [(int(i), int(v)) for i, v in b.enumerate(start=1).take(2)] == [(1, 10), (2, 11)]
    evens = a.concatenate(b).filter(lambda v: v % 2 == 0)
    assert [int(v) for v in evens] == [0, 2, 4, 10, 12, 14, 16] and evens.cardinality() == data.experimental.UNKNOWN_CARDINALITY
    assert [t.tolist() for t in evens.batch(3)] == [[0, 2, 4], [10, 12, 14], [16]]
    pairs = data.Dataset.from_tensor_slices((np.arange(6), np.arange(6) * 10)).filter(lambda x, y: y >= 30)
    assert [(int(x), int(y)) for x, y in pairs] == [(3, 30), (4, 40), (5, 50)]
    assert int(a.reduce(np.int64(0), lambda s, v: s + v)) == 10
    assert [int(v) for v in a.apply(lambda ds: ds.skip(3))] == [3, 4]
    # zipped features and labels feed fit like a dataset of pairs
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    x = data.Dataset.from_tensor_slices(np.random.rand(32, 4).astype("float32"))
    y = data.Dataset.from_tensor_slices(np.random.randint(0, 2, 32))
    model = tfc.keras.Sequential([tfc.keras.layers.Dense(2, activation="softmax", input_shape=(4,))])
    model.compile("adam", "sparse_categorical_crossentropy")
    hist = model.fit(data.Dataset.zip((x, y)).batch(8), epochs=2, verbose=0)
    assert len(hist.history["loss"]) == 2
