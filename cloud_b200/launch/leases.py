"""GPU leases: which job holds which GPU of the box.

AI Platform Training gave every job of the reference its own machines (deploy.py:82-88), so two
``tfc.run`` calls, or the parallel trial jobs of a ``DistributingCloudTuner``, never met.  On one box
they do: without bookkeeping every job would start at GPU 0.  The lease table is a JSON file under
the job root, mutated under an exclusive ``flock``:

    {"gpus": {"0": {"job": "tf_cloud_train_...", "pid": 1234, "time": ...}, ...}, "queue": [...]}

* a supervisor asks for ``n`` GPUs before it starts its ranks; it gets the lowest-numbered free ones
  (NVSwitch: every GPU reaches every other at full bandwidth, so there is nothing to optimise in the
  placement) or waits in FIFO order while the job state stays ``QUEUED``;
* a lease whose holder process has died is free again (no clean-up daemon needed);
* the trial scheduler takes its worker slots from the same table, so tuner workers and launched
  jobs share the box without stepping on each other.
"""
from __future__ import annotations

import contextlib
import fcntl
import json
import os
import time

from ..core.containerize import job_root


class LeaseTimeout(RuntimeError):
    pass


def _path() -> str:
    root = job_root()
    os.makedirs(root, exist_ok=True)
    return os.path.join(root, "gpu_leases.json")


@contextlib.contextmanager
def _locked():
    path = _path()
    fd = os.open(path + ".lock", os.O_CREAT | os.O_RDWR, 0o644)
    try:
        fcntl.flock(fd, fcntl.LOCK_EX)
        try:
            with open(path) as f:
                table = json.load(f)
        except (OSError, ValueError):
            table = {}
        table.setdefault("gpus", {})
        table.setdefault("queue", [])
        yield table
        tmp = "{}.{}.tmp".format(path, os.getpid())
        with open(tmp, "w") as f:
            json.dump(table, f, indent=1)
        os.replace(tmp, path)
    finally:
        fcntl.flock(fd, fcntl.LOCK_UN)
        os.close(fd)


def _alive(pid) -> bool:
    try:
        os.kill(int(pid), 0)
    except ProcessLookupError:
        return False
    except (PermissionError, ValueError, TypeError):
        return True
    # a zombie still answers signal 0: look at its state
    try:
        with open("/proc/{}/stat".format(int(pid))) as f:
            return f.read().rsplit(")", 1)[1].split()[0] != "Z"
    except (OSError, IndexError):
        return True


def _sweep(table) -> None:
    for gpu, lease in list(table["gpus"].items()):
        if not _alive(lease.get("pid")):
            del table["gpus"][gpu]
    table["queue"] = [q for q in table["queue"] if _alive(q.get("pid"))]


def snapshot(total_gpus: int | None = None) -> dict:
    """The table after dropping dead holders: ``{"gpus": {id: lease}, "queue": [...], "free": [...]}``."""
    with _locked() as table:
        _sweep(table)
        out = {"gpus": dict(table["gpus"]), "queue": list(table["queue"])}
    if total_gpus is not None:
        out["free"] = [g for g in range(total_gpus) if str(g) not in out["gpus"]]
    return out


def try_acquire(owner: str, n: int, total_gpus: int, pid: int | None = None, candidates=None):
    """``n`` free GPU ids for ``owner`` (held until ``release`` or until ``pid`` dies), or None when
    fewer than ``n`` are free or an earlier request is still waiting."""
    pid = os.getpid() if pid is None else pid
    if n <= 0:
        return []
    if n > total_gpus:
        raise ValueError("{} asks for {} GPUs but the box has {}".format(owner, n, total_gpus))
    with _locked() as table:
        _sweep(table)
        queue = table["queue"]
        if not any(q["owner"] == owner for q in queue):
            queue.append({"owner": owner, "pid": pid, "n": n, "time": time.time()})
        pool = list(range(total_gpus)) if candidates is None else list(candidates)
        free = [g for g in pool if str(g) not in table["gpus"]]
        first = queue[0]["owner"] == owner
        if not first or len(free) < n:
            return None
        got = free[:n]
        for g in got:
            table["gpus"][str(g)] = {"job": owner, "pid": pid, "time": time.time()}
        table["queue"] = [q for q in queue if q["owner"] != owner]
        return got


def acquire(owner: str, n: int, total_gpus: int, pid: int | None = None, timeout: float | None = None,
            poll: float = 0.2, on_wait=None, should_stop=None):
    """Blocks (FIFO) until ``n`` GPUs are free; ``LeaseTimeout`` after ``timeout`` seconds; None when
    ``should_stop()`` turns true while waiting (the request leaves the queue)."""
    t0 = time.time()
    told = False
    while True:
        got = try_acquire(owner, n, total_gpus, pid)
        if got is not None:
            return got
        if should_stop is not None and should_stop():
            cancel(owner)
            return None
        if on_wait is not None and not told:
            on_wait()
            told = True
        if timeout is not None and time.time() - t0 > timeout:
            cancel(owner)
            raise LeaseTimeout("{} waited {:.0f} s for {} free GPU(s)".format(owner, time.time() - t0, n))
        time.sleep(poll)


def cancel(owner: str) -> None:
    with _locked() as table:
        table["queue"] = [q for q in table["queue"] if q["owner"] != owner]


def release(owner: str) -> list:
    """Frees everything ``owner`` holds; returns the GPU ids."""
    with _locked() as table:
        mine = [g for g, lease in table["gpus"].items() if lease.get("job") == owner]
        for g in mine:
            del table["gpus"][g]
        table["queue"] = [q for q in table["queue"] if q["owner"] != owner]
    return sorted(int(g) for g in mine)
