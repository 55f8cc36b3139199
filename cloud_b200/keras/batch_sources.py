"""Where ``fit`` / ``evaluate`` get their batches from (host side of K12):
array sources that slice a replica's share of every global batch (Keras: ``batch_size`` is the
global batch), dataset / iterable sources that re-split global batches over replicas, the grouping
of K batches for ``steps_per_execution``, the pinned-memory prefetcher that overlaps the host-to-
device copy of the next batch with the current step, and the launcher heartbeat.
(Moved out of ``models.py``; the reference gets all of this from ``tf.keras`` / ``tf.data``.)
"""
from __future__ import annotations

import os
import queue
import threading
import time

import numpy as np
import torch

from ..data import Dataset
from ..distribute.strategy import DistributedDataset

def _np(a):
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


# ----------------------------------------------------------------------------------------------
# Batch sources
# ----------------------------------------------------------------------------------------------
_HEARTBEAT = {"path": None, "last": 0.0, "resolved": False}


def _heartbeat():
    """Under the launcher: touch <job_dir>/heartbeat/rank<r> at most once a second.  The host blocks
    on the device within a step or two (loss read-back, launch queue), so a hung kernel or
    collective silences it and the supervisor's TFC_HEARTBEAT_TIMEOUT check ends the job."""
    hb = _HEARTBEAT
    if not hb["resolved"]:
        hb["resolved"] = True
        jdir = os.environ.get("TFC_JOB_DIR")
        if jdir:
            d = os.path.join(jdir, "heartbeat")
            try:
                os.makedirs(d, exist_ok=True)
                hb["path"] = os.path.join(d, "rank{}".format(os.environ.get("RANK", "0")))
            except OSError:
                hb["path"] = None
    if hb["path"] is None:
        return
    now = time.monotonic()
    if now - hb["last"] >= 1.0:
        hb["last"] = now
        try:
            with open(hb["path"], "a"):
                os.utime(hb["path"], None)
        except OSError:
            pass


def _grouped(batches, k):
    """Lists of ``k`` consecutive batches with identical shapes (``steps_per_execution``); whatever
    does not fill a group - the tail of an epoch, a short last batch - comes out one by one."""
    if k <= 1:
        for b in batches:
            yield [b]
        return
    buf = []
    for b in batches:
        if b[1] is None or getattr(b[3], "k", 1) > 1:     # unlabelled / already stacked by the source
            for old in buf:
                yield [old]
            buf = []
            yield [b]
            continue
        if buf and (b[0].shape != buf[0][0].shape or b[3] != buf[0][3]):
            for old in buf:
                yield [old]
            buf = []
        buf.append(b)
        if len(buf) == k:
            yield buf
            buf = []
    for old in buf:
        yield [old]


class _NG(int):
    """Global batch size of a source item that carries ``k`` stacked batches ([k, per, ...])."""
    k = 1


class _ArraySource:
    """numpy/torch arrays -> per-replica batches of a global batch (Keras: batch_size is global).
    With ``group=K`` (``steps_per_execution``) K consecutive full batches travel as one
    ``[K, per_replica, ...]`` item - one host slice/gather and one H2D copy per execution."""

    def __init__(self, x, y, batch_size, shuffle, world, rank, seed=0, steps=None, drop_last=False,
                 group=1):
        self.group = max(1, int(group))
        self.x = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(_np(x)))
        self.y = None
        if y is not None:
            self.y = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(_np(y)))
        self.n = len(self.x)
        self.batch = int(batch_size)
        self.shuffle = shuffle
        self.world, self.rank = world, rank
        self.seed = seed
        self.epoch = 0
        self.steps = steps
        self.drop_last = drop_last

    def __len__(self):
        full = self.n // self.batch if self.drop_last else -(-self.n // self.batch)
        rem = self.n - (full - 1) * self.batch if full else 0
        if full and self.world > 1 and 0 < rem < self.world:
            # a tail with fewer samples than replicas would leave some replica with an empty batch
            # while the others wait for it in the gradient all-reduce: every replica drops that step
            full -= 1
        return min(full, self.steps) if self.steps else full

    def __iter__(self):
        n_steps = len(self)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)      # same permutation on every replica
            perm = torch.randperm(self.n, generator=g)
        else:
            perm = None
        self.epoch += 1
        s = 0
        K = self.group
        per_full = -(-self.batch // self.world)
        while K > 1 and self.y is not None and s + K <= n_steps and (s + K) * self.batch <= self.n \
                and per_full * self.world == self.batch:
            # K full global batches: this replica's rows of each, stacked
            if perm is None and self.world == 1:
                xb = self.x[s * self.batch:(s + K) * self.batch]
                yb = self.y[s * self.batch:(s + K) * self.batch]
            else:
                base = torch.arange(K).repeat_interleave(per_full) * self.batch + \
                    torch.arange(per_full).repeat(K) + self.rank * per_full + s * self.batch
                idx = perm[base] if perm is not None else base
                xb, yb = self.x.index_select(0, idx), self.y.index_select(0, idx)
            ng = _NG(self.batch)
            ng.k = K
            yield xb.reshape(K, per_full, *self.x.shape[1:]), yb.reshape(K, per_full, *self.y.shape[1:]), ng
            s += K
        for s in range(s, n_steps):
            lo = s * self.batch
            hi = min(self.n, lo + self.batch)
            # tail batches are split as evenly as they go: replica r gets floor(rem / world) rows
            # plus one of the rem % world left-over rows if r is among the first ones
            q, extra = divmod(hi - lo, self.world)
            a = lo + self.rank * q + min(self.rank, extra)
            b = a + q + (1 if self.rank < extra else 0)
            if perm is not None:
                idx = perm[a:b]
                xb = self.x.index_select(0, idx)
                yb = self.y.index_select(0, idx) if self.y is not None else None
            else:
                xb = self.x[a:b]
                yb = self.y[a:b] if self.y is not None else None
            yield xb, yb, hi - lo


class _IterSource:
    """Dataset / iterable of (x, y) global batches, re-split over replicas."""

    def __init__(self, ds, world, rank, steps=None):
        self.ds = ds
        self.world, self.rank = world, rank
        self.steps = steps
        self._it = None

    def __len__(self):
        if self.steps:
            return self.steps
        return len(self.ds)            # TypeError when the dataset's size is unknown (generators, filters)

    def __iter__(self):
        infinite = False
        if isinstance(self.ds, Dataset):
            infinite = self.ds.cardinality() == -1
        if self._it is None or not infinite:
            self._it = iter(self.ds)
        count = 0
        while self.steps is None or count < self.steps:
            try:
                batch = next(self._it)
            except StopIteration:
                if self.steps is not None and count < self.steps and not infinite:
                    self._it = iter(self.ds)     # Keras would warn "ran out of data"; we wrap
                    try:
                        batch = next(self._it)
                    except StopIteration:
                        return
                else:
                    return
            count += 1
            if isinstance(batch, (tuple, list)):
                xb, yb = batch[0], (batch[1] if len(batch) > 1 else None)
            else:
                xb, yb = batch, None
            xb = torch.as_tensor(_np(xb)) if not isinstance(xb, torch.Tensor) else xb
            if yb is not None and not isinstance(yb, torch.Tensor):
                yb = torch.as_tensor(_np(yb))
            n_global = len(xb)
            if self.world > 1 and not isinstance(self.ds, DistributedDataset):
                per = -(-n_global // self.world)
                a, b = self.rank * per, min(n_global, (self.rank + 1) * per)
                xb = xb[a:b]
                yb = yb[a:b] if yb is not None else None
            elif self.world > 1:
                n_global = None          # already split by the strategy: assume an even split
            yield xb, yb, n_global


class _Prefetcher:
    """Background thread: source batch -> pinned staging -> async H2D on a copy stream."""

    _PIN_CACHE: dict = {}      # staging buffers survive across epochs and fit() calls

    def __init__(self, source, device, depth=3):
        self.source = source
        self.device = device
        self.cuda = device.type == "cuda"
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(device) if self.cuda else None
        self.h2d_bytes = 0

    def __iter__(self):
        if not self.cuda:
            for xb, yb, ng in self.source:
                yield xb, yb, None, ng
            return
        q: queue.Queue = queue.Queue(maxsize=self.depth)
        stop = threading.Event()
        pinned = _Prefetcher._PIN_CACHE        # pinning (cudaHostAlloc) is slow: do it once

        def pin(slot, name, t):
            key = (slot, name, tuple(t.shape), t.dtype)
            buf = pinned.get(key)
            if buf is None:
                buf = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                pinned[key] = buf
            buf.copy_(t)
            return buf

        def worker():
            torch.cuda.set_device(self.device)
            try:
                for i, (xb, yb, ng) in enumerate(self.source):
                    if stop.is_set():
                        return
                    slot = i % (self.depth + 2)
                    xp = xb if xb.is_pinned() else pin(slot, "x", xb)
                    yp = None
                    if yb is not None:
                        yp = yb if yb.is_pinned() else pin(slot, "y", yb)
                    with torch.cuda.stream(self.copy_stream):
                        xd = xp.to(self.device, non_blocking=True)
                        yd = yp.to(self.device, non_blocking=True) if yp is not None else None
                        ev = torch.cuda.Event()
                        ev.record(self.copy_stream)
                    self.h2d_bytes += xp.numel() * xp.element_size() + \
                        (yp.numel() * yp.element_size() if yp is not None else 0)
                    q.put((xd, yd, ev, ng))
                q.put(None)
            except BaseException as e:  # surface loader errors in the training thread
                q.put(e)

        th = threading.Thread(target=worker, daemon=True)
        th.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
            while th.is_alive():
                try:
                    q.get_nowait()
                except queue.Empty:
                    time.sleep(0.001)


# ----------------------------------------------------------------------------------------------


class Sequence:
    """``tf.keras.utils.Sequence``: ``__len__`` batches per epoch, ``__getitem__(i)`` returns batch i
    as ``(x, y)``, ``on_epoch_end`` runs after every pass (reshuffle there)."""

    def __len__(self):
        raise NotImplementedError

    def __getitem__(self, index):
        raise NotImplementedError

    def on_epoch_end(self):
        pass

    def __iter__(self):
        return iter(_SequenceAdapter(self))


class _SequenceAdapter:
    """Makes anything with ``__len__`` and ``__getitem__`` (a Sequence, also one that does not derive
    from the class above) an iterable of global batches for ``_IterSource``."""

    def __init__(self, seq):
        self.seq = seq

    def __len__(self):
        return len(self.seq)

    def __iter__(self):
        for i in range(len(self.seq)):
            yield self.seq[i]
        hook = getattr(self.seq, "on_epoch_end", None)
        if callable(hook):
            hook()
