"""A small ``tf.data``-shaped input pipeline on numpy (host side of K12).

Covers what the reference's scripts build: ``Dataset.from_tensor_slices((x, y)).map(f).cache()
.shuffle(n).batch(b).repeat().prefetch(k).take(n)`` (core/tests/testdata/mnist_example_using_fit.py:44-51,
tuner/tests/integration/tuner_integration_test.py:200-207) plus ``with_options`` /
``experimental_distribute_dataset`` hand-off to a strategy.  Batches are numpy arrays; the Keras
front end stages them through pinned memory and overlaps the H2D copy with compute.
"""
from __future__ import annotations

import numpy as np


def _map_structure(fn, s):
    if isinstance(s, (tuple, list)):
        return tuple(_map_structure(fn, e) for e in s)
    if isinstance(s, dict):
        return {k: _map_structure(fn, v) for k, v in s.items()}
    return fn(s)


def _first_leaf(s):
    while isinstance(s, (tuple, list, dict)):
        s = next(iter(s.values())) if isinstance(s, dict) else s[0]
    return s


def _to_numpy(a):
    try:
        import torch
        if isinstance(a, torch.Tensor):
            return a.detach().cpu().numpy()
    except ImportError:  # pragma: no cover
        pass
    return np.asarray(a)


class Dataset:
    AUTOTUNE = -1

    def __init__(self, source, ops=None):
        self._source = source          # ("slices", structure) | ("generator", callable)
        self._ops = list(ops or [])

    # -- constructors ---------------------------------------------------------------------------
    @staticmethod
    def from_tensor_slices(tensors):
        return Dataset(("slices", _map_structure(_to_numpy, tensors)))

    @staticmethod
    def from_tensors(tensors):
        return Dataset(("slices", _map_structure(lambda a: _to_numpy(a)[None], tensors)))

    @staticmethod
    def from_generator(generator, **_):
        return Dataset(("generator", generator))

    @staticmethod
    def range(n):
        return Dataset(("slices", np.arange(n)))

    @staticmethod
    def zip(datasets):
        """Elements are tuples of the corresponding elements of ``datasets``; ends with the shortest."""
        parts = tuple(datasets)

        def gen():
            yield from zip(*parts)
        return Dataset(("generator", gen))

    def _with(self, op):
        return Dataset(self._source, self._ops + [op])

    def concatenate(self, other):
        first = self

        def gen():
            yield from first
            yield from other
        return Dataset(("generator", gen))

    def enumerate(self, start=0):
        base = self

        def gen():
            for i, e in enumerate(base, start):
                yield (np.int64(i), e)
        return Dataset(("generator", gen))

    def filter(self, predicate):
        return self._with(("filter", predicate))

    def reduce(self, initial_state, reduce_func):
        state = initial_state
        for e in self:
            state = reduce_func(state, e)
        return state

    def apply(self, transformation_func):
        return transformation_func(self)

    # -- transformations ------------------------------------------------------------------------
    def map(self, fn, num_parallel_calls=None, deterministic=True, **_):
        """``num_parallel_calls`` (or AUTOTUNE): the function runs on a pool of threads, results come
        out in input order; numpy releases the GIL in the heavy calls (resize, casts), so decoding /
        augmenting scales with cores like tf.data's parallel map."""
        workers = 0
        if num_parallel_calls is not None:
            workers = _autotune_workers() if int(num_parallel_calls) < 0 else int(num_parallel_calls)
        return self._with(("map", fn, workers if workers > 1 else 0))

    def cache(self, filename=""):
        return self._with(("cache",))

    def shuffle(self, buffer_size, seed=None, reshuffle_each_iteration=True):
        return self._with(("shuffle", int(buffer_size), seed))

    def batch(self, batch_size, drop_remainder=False, **_):
        return self._with(("batch", int(batch_size), bool(drop_remainder)))

    def repeat(self, count=None):
        return self._with(("repeat", count))

    def take(self, count):
        return self._with(("take", int(count)))

    def skip(self, count):
        return self._with(("skip", int(count)))

    def prefetch(self, buffer_size=None):
        """Elements are produced by a background thread, up to ``buffer_size`` ahead of the consumer
        (AUTOTUNE: 2), overlapping the input pipeline with the training step."""
        n = 2 if buffer_size is None or int(buffer_size) < 0 else int(buffer_size)
        return self._with(("prefetch", n)) if n > 0 else self

    def with_options(self, options):
        return self

    def unbatch(self):
        return self._with(("unbatch",))

    # -- introspection --------------------------------------------------------------------------
    @property
    def batched(self) -> bool:
        return any(op[0] == "batch" for op in self._ops)

    def cardinality(self):
        n = self._source_len()
        for op in self._ops:
            if n is None:
                return -2
            if op[0] == "batch":
                n = n // op[1] if op[2] else -(-n // op[1])
            elif op[0] == "repeat":
                if op[1] is None:
                    return -1
                n = n * op[1]
            elif op[0] == "take":
                n = min(n, op[1])
            elif op[0] == "skip":
                n = max(n - op[1], 0)
            elif op[0] in ("unbatch", "filter"):
                return -2
        return n                        # map / cache / shuffle / prefetch keep the count

    def __len__(self):
        n = self.cardinality()
        if n is None or n < 0:
            raise TypeError("dataset length is infinite or unknown")
        return n

    def _source_len(self):
        if self._source[0] == "slices":
            return len(_first_leaf(self._source[1]))
        return None

    # -- iteration ------------------------------------------------------------------------------
    def _elements(self):
        kind, payload = self._source
        if kind == "slices":
            n = len(_first_leaf(payload))
            for i in range(n):
                yield _map_structure(lambda a: a[i], payload)
        else:
            yield from payload()

    def _vectorised(self):
        """Fast path: slices -> [element-wise map]* -> cache -> shuffle -> batch ... handled on
        whole arrays when the map functions accept batched arrays."""
        return None

    def __iter__(self):
        it = self._elements()
        epoch_rng = np.random.default_rng()
        for idx, op in enumerate(self._ops):
            it = _apply(it, op, epoch_rng, self, idx)
        return it

    def as_numpy_iterator(self):
        return iter(self)


def _autotune_workers() -> int:
    import os
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        cores = os.cpu_count() or 1
    return max(1, min(8, cores))


def _apply(it, op, rng, ds, idx):
    kind = op[0]
    if kind == "map":
        fn = op[1]
        workers = op[2] if len(op) > 2 else 0

        def call(e):
            out = fn(*e) if isinstance(e, tuple) else fn(e)
            return _map_structure(_to_numpy, out) if not isinstance(out, tuple) else \
                tuple(_map_structure(_to_numpy, o) for o in out)

        def gen():
            for e in it:
                yield call(e)

        def parallel_gen():
            import collections
            from concurrent.futures import ThreadPoolExecutor
            window = collections.deque()
            pool = ThreadPoolExecutor(max_workers=workers, thread_name_prefix="tfc-data-map")
            try:
                for e in it:
                    window.append(pool.submit(call, e))
                    if len(window) >= 2 * workers:
                        yield window.popleft().result()      # input order is kept
                while window:
                    yield window.popleft().result()
            finally:
                for f in window:
                    f.cancel()
                pool.shutdown(wait=False)
        return parallel_gen() if workers else gen()
    if kind == "prefetch":
        def gen():
            import queue
            import threading
            q = queue.Queue(maxsize=op[1])
            stop = threading.Event()
            end = object()

            def produce():
                try:
                    for e in it:
                        while not stop.is_set():
                            try:
                                q.put(e, timeout=0.1)
                                break
                            except queue.Full:
                                continue
                        if stop.is_set():
                            return
                    item = end
                except BaseException as err:             # surfaces in the consumer
                    item = err
                while not stop.is_set():
                    try:
                        q.put(item, timeout=0.1)
                        return
                    except queue.Full:
                        continue
            t = threading.Thread(target=produce, name="tfc-data-prefetch", daemon=True)
            t.start()
            try:
                while True:
                    e = q.get()
                    if e is end:
                        return
                    if isinstance(e, BaseException):
                        raise e
                    yield e
            finally:
                stop.set()
        return gen()
    if kind == "filter":
        def gen():
            for e in it:
                if bool(op[1](*e) if isinstance(e, tuple) else op[1](e)):
                    yield e
        return gen()
    if kind == "cache":
        key = ("_cache", idx)
        if not hasattr(ds, "_cache_store"):
            ds._cache_store = {}
        if key in ds._cache_store:
            return iter(ds._cache_store[key])

        def gen():
            buf = []
            for e in it:
                buf.append(e)
                yield e
            ds._cache_store[key] = buf
        return gen()
    if kind == "shuffle":
        size, seed = op[1], op[2]
        r = np.random.default_rng(seed) if seed is not None else rng

        def gen():
            buf = []
            for e in it:
                buf.append(e)
                if len(buf) >= size:
                    j = int(r.integers(len(buf)))
                    buf[j], buf[-1] = buf[-1], buf[j]
                    yield buf.pop()
            r.shuffle(buf)
            yield from buf
        return gen()
    if kind == "batch":
        size, drop = op[1], op[2]

        def gen():
            buf = []
            for e in it:
                buf.append(e)
                if len(buf) == size:
                    yield _stack(buf)
                    buf = []
            if buf and not drop:
                yield _stack(buf)
        return gen()
    if kind == "repeat":
        count = op[1]

        def gen():
            first = list(it)
            yield from first
            n = 1
            while count is None or n < count:
                # re-run the pipeline prefix so shuffles reshuffle each repetition
                sub = Dataset(ds._source, ds._ops[:idx])
                sub._cache_store = getattr(ds, "_cache_store", {})
                got = False
                for e in sub:
                    got = True
                    yield e
                if not got:
                    return
                n += 1
        return gen()
    if kind == "take":
        def gen():
            for i, e in enumerate(it):
                if i >= op[1]:
                    return
                yield e
        return gen()
    if kind == "skip":
        def gen():
            for i, e in enumerate(it):
                if i >= op[1]:
                    yield e
        return gen()
    if kind == "unbatch":
        def gen():
            for e in it:
                n = len(_first_leaf(e))
                for i in range(n):
                    yield _map_structure(lambda a: a[i], e)
        return gen()
    raise ValueError(f"unknown dataset op {kind}")


def _stack(elems):
    first = elems[0]
    if isinstance(first, tuple):
        return tuple(_stack([e[i] for e in elems]) for i in range(len(first)))
    if isinstance(first, dict):
        return {k: _stack([e[k] for e in elems]) for k in first}
    return np.stack([np.asarray(e) for e in elems])


class Options:
    """``tf.data.Options`` placeholder (auto-shard policy is always DATA here)."""

    class _Distribute:
        auto_shard_policy = None

    def __init__(self):
        self.experimental_distribute = Options._Distribute()


class experimental:
    AUTOTUNE = -1

    class AutoShardPolicy:
        OFF, AUTO, FILE, DATA = 0, 1, 2, 3

    INFINITE_CARDINALITY = -1
    UNKNOWN_CARDINALITY = -2

    @staticmethod
    def cardinality(dataset):
        """``tf.data.experimental.cardinality(ds)`` (dogs_classification.ipynb prints it)."""
        return dataset.cardinality()


AUTOTUNE = -1
