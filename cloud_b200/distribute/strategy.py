"""Strategy objects (see package docstring).  One process per device; `torch.distributed` is the
control plane, the DPEngine + peer-memory kernels are the data plane."""
from __future__ import annotations

import contextlib
import enum
import json
import os
import threading
import warnings

import torch
import torch.distributed as dist


_exit_hook_installed = False


def _destroy_group_at_exit():
    """The strategy created the process group, so it also tears it down (NCCL warns otherwise)."""
    global _exit_hook_installed
    if _exit_hook_installed:
        return
    _exit_hook_installed = True
    import atexit

    def _cleanup():
        try:
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:      # interpreter shutdown: nothing useful can be done with an error here
            pass
    atexit.register(_cleanup)


class ReduceOp(enum.Enum):
    SUM = "SUM"
    MEAN = "MEAN"


class PerReplica:
    """Result of ``strategy.run`` - in a process-per-replica design it holds the local value."""

    def __init__(self, value):
        self.values = (value,)
        self.local = value


_global = {"strategy": None}
_tls = threading.local()


def _scope_stack():
    if not hasattr(_tls, "stack"):
        _tls.stack = []
    return _tls.stack


def _env_world():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), \
        int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))


def _parse_device(device: str | torch.device | None) -> torch.device:
    """Accepts TF spellings ('/gpu:0', '/cpu:0', 'GPU:1') and torch spellings."""
    if isinstance(device, torch.device):
        return device
    if device is None:
        return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
            else torch.device("cpu")
    d = device.lower().strip("/")
    if d.startswith("device:"):
        d = d[len("device:"):]
    kind, _, idx = d.partition(":")
    if kind in ("gpu", "cuda"):
        if not torch.cuda.is_available():
            # the reference emits OneDeviceStrategy('/gpu:0') even for the CPU preset
            # (core/preprocess.py:144-146); on a GPU-less replica that means "the one device we have"
            return torch.device("cpu")
        return torch.device("cuda", int(idx or 0))
    if kind == "cpu":
        return torch.device("cpu")
    raise ValueError(f"cannot parse device {device!r}")


class Strategy:
    """Base: a replica group + device.  Subclasses only differ in how the group is formed."""

    def __init__(self, device: torch.device, use_process_group: bool):
        self.device = device
        if device.type == "cuda":
            torch.cuda.set_device(device)
            # cuDNN times its algorithms during the eager warm-up steps, before the step is frozen into
            # a CUDA graph: 19.31 -> 19.07 ms/step on ResNet-50 at batch 256 (TFC_CUDNN_BENCHMARK=0: off)
            if os.environ.get("TFC_CUDNN_BENCHMARK", "1") != "0":
                torch.backends.cudnn.benchmark = True
        self._group = None
        world, rank, _local = _env_world()
        if use_process_group and world > 1:
            if not dist.is_initialized():
                backend = "nccl" if device.type == "cuda" else "gloo"
                kwargs = {}
                if device.type == "cuda":
                    kwargs["device_id"] = device
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29512")
                dist.init_process_group(backend, rank=rank, world_size=world, **kwargs)
                _destroy_group_at_exit()
            self._world, self._rank = dist.get_world_size(), dist.get_rank()
        else:
            self._world, self._rank = 1, 0
        self._communicator = None

    # --- properties user code reads ------------------------------------------------------------
    @property
    def num_replicas_in_sync(self) -> int:
        return self._world

    @property
    def replica_id(self) -> int:
        return self._rank

    @property
    def is_chief(self) -> bool:
        """Chief detection as in experimental/cloud_fit/remote.py:148-156: TF_CONFIG task of type
        'chief' or index 0; without TF_CONFIG, replica 0."""
        cfg = os.environ.get("TF_CONFIG")
        if cfg:
            try:
                parsed = json.loads(cfg)
                task = parsed.get("task", {})
                if task:
                    # a cluster with a dedicated chief: only that task is chief; without one,
                    # worker 0 plays the role (remote.py:148-156)
                    if "chief" in parsed.get("cluster", {}):
                        is_chief_task = task.get("type") == "chief"
                    else:
                        is_chief_task = task.get("type") == "chief" or int(task.get("index", 0)) == 0
                    return is_chief_task and self._rank_in_task() == 0
            except (ValueError, TypeError):
                pass
        return self._rank == 0

    def _rank_in_task(self) -> int:
        return int(os.environ.get("TFC_RANK_IN_TASK", "0"))

    @property
    def extended(self):
        return self

    # --- communicator --------------------------------------------------------------------------
    @property
    def communicator(self):
        if self._communicator is None:
            from ..parallel.comm import Communicator
            self._communicator = Communicator(self.device, self._group, single=self._world == 1)
        return self._communicator

    # --- scope ---------------------------------------------------------------------------------
    @contextlib.contextmanager
    def scope(self):
        _scope_stack().append(self)
        try:
            yield self
        finally:
            _scope_stack().pop()

    # --- data ----------------------------------------------------------------------------------
    def experimental_distribute_dataset(self, dataset):
        return DistributedDataset(dataset, self)

    def distribute_datasets_from_function(self, fn):
        return DistributedDataset(fn(self), self, already_sharded=True)

    # --- compute -------------------------------------------------------------------------------
    def run(self, fn, args=(), kwargs=None):
        with self.scope():
            return PerReplica(fn(*args, **(kwargs or {})))

    experimental_run_v2 = run   # name used by testdata/mnist_example_using_ctl.py:152-154

    def reduce(self, reduce_op, value, axis=None):
        """Cross-replica reduction of a per-replica value (mnist_example_using_ctl.py:155-157)."""
        if isinstance(value, PerReplica):
            value = value.local
        if isinstance(reduce_op, str):
            reduce_op = ReduceOp[reduce_op.upper()]
        t = value if isinstance(value, torch.Tensor) else torch.as_tensor(value, dtype=torch.float32)
        t = t.detach().to(self.device, dtype=torch.float32).clone()
        if axis is not None:
            t = t.sum(dim=axis) if reduce_op == ReduceOp.SUM else t.mean(dim=axis)
        if self._world > 1:
            self.communicator.all_reduce_sum(t)
            if reduce_op == ReduceOp.MEAN:
                t /= self._world
        return t

    def barrier(self):
        if self._world > 1:
            dist.barrier()

    def __repr__(self):
        return f"{type(self).__name__}(device={self.device}, replicas={self._world}, rank={self._rank})"


class OneDeviceStrategy(Strategy):
    def __init__(self, device="/gpu:0"):
        super().__init__(_parse_device(device), use_process_group=False)


class MirroredStrategy(Strategy):
    """One replica per GPU of the job.  Under the launcher (or torchrun) every process is one
    replica bound to ``LOCAL_RANK``; started by hand in a single process it degrades to one
    replica on the current device with a warning (one process cannot drive several GPUs in this
    design - use ``tfc.run`` or the launcher)."""

    def __init__(self, devices=None, cross_device_ops=None):
        world, _rank, local = _env_world()
        if torch.cuda.is_available():
            device = torch.device("cuda", local % max(torch.cuda.device_count(), 1))
        else:
            device = torch.device("cpu")
        if world == 1 and torch.cuda.is_available() and torch.cuda.device_count() > 1 \
                and not os.environ.get("TF_KERAS_RUNNING_REMOTELY"):
            warnings.warn("MirroredStrategy created outside the cloud_b200 launcher: using one "
                          "replica on the current GPU.  Launch with tfc.run(...) or "
                          "`python -m cloud_b200.launch` to use every GPU.")
        super().__init__(device, use_process_group=True)


class MultiWorkerMirroredStrategy(MirroredStrategy):
    """Chief + workers.  The launcher synthesises ``TF_CONFIG`` per rank group so that
    chief-only saving code keeps working (experimental/cloud_fit/remote.py:130-156)."""

    def __init__(self, communication=None, cluster_resolver=None):
        super().__init__()


class TPUStrategy(Strategy):
    def __init__(self, *a, **k):
        raise NotImplementedError("There is no TPU on this box: cloud_b200 targets NVIDIA B200 GPUs. "
                                  "Use a B200 machine config instead of TPU_V2/TPU_V3.")


def experimental_set_strategy(strategy: Strategy | None) -> None:
    """Install a process-global default strategy (core/preprocess.py:148-149)."""
    _global["strategy"] = strategy


def has_strategy() -> bool:
    return bool(_scope_stack()) or _global["strategy"] is not None


_default = {"strategy": None}


def get_strategy() -> Strategy:
    """Innermost ``scope()``, else the global default, else a default one-device strategy."""
    st = _scope_stack()
    if st:
        return st[-1]
    if _global["strategy"] is not None:
        return _global["strategy"]
    if _default["strategy"] is None:
        _default["strategy"] = OneDeviceStrategy(None)
    return _default["strategy"]


def in_cross_replica_context() -> bool:
    return True


def get_replica_context():
    return None


class DistributedDataset:
    """Splits each *global* batch of ``dataset`` across replicas (B1 in SURVEY.md appendix B: a
    batched dataset handed to a strategy is treated as global batches and re-split)."""

    def __init__(self, dataset, strategy: Strategy, already_sharded: bool = False):
        self.dataset = dataset
        self.strategy = strategy
        self.already_sharded = already_sharded

    def __iter__(self):
        w, r = self.strategy.num_replicas_in_sync, self.strategy.replica_id
        for batch in self.dataset:
            if self.already_sharded or w == 1:
                yield batch
                continue
            yield _shard(batch, w, r)

    def __len__(self):
        return len(self.dataset)


def _shard(batch, world, rank):
    if isinstance(batch, (tuple, list)):
        return type(batch)(_shard(b, world, rank) for b in batch)
    if isinstance(batch, dict):
        return {k: _shard(v, world, rank) for k, v in batch.items()}
    n = len(batch)
    per = (n + world - 1) // world
    return batch[rank * per: min(n, (rank + 1) * per)]
