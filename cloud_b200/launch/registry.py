"""File-based job registry: ``<job_root>/jobs/<job_id>/{job.json,status.json,logs/}``.
Plays the role of the AI Platform job list the reference submits into (deploy.py:82-88) and that
DistributingCloudTuner polls.  Writes are atomic (tmp + rename) so several processes can read."""
from __future__ import annotations

import json
import os
import time

from ..core.containerize import job_root

TERMINAL = ("SUCCEEDED", "FAILED", "CANCELLED")


def job_dir(job_id: str) -> str:
    return os.path.join(job_root(), "jobs", job_id)


def _atomic_write(path: str, obj) -> None:
    tmp = f"{path}.{os.getpid()}.tmp"
    with open(tmp, "w") as f:
        json.dump(obj, f, indent=1)
    os.replace(tmp, path)


def create(job_id: str, spec: dict) -> str:
    d = job_dir(job_id)
    os.makedirs(os.path.join(d, "logs"), exist_ok=True)
    _atomic_write(os.path.join(d, "job.json"), spec)
    set_status(job_id, "QUEUED")
    return d


def set_status(job_id: str, state: str, **extra) -> None:
    rec = {"jobId": job_id, "state": state, "time": time.time()}
    rec.update(extra)
    _atomic_write(os.path.join(job_dir(job_id), "status.json"), rec)


def get(job_id: str) -> dict:
    with open(os.path.join(job_dir(job_id), "job.json")) as f:
        return json.load(f)


def status(job_id: str) -> dict:
    path = os.path.join(job_dir(job_id), "status.json")
    for _ in range(5):
        try:
            with open(path) as f:
                return json.load(f)
        except (OSError, ValueError):
            time.sleep(0.05)
    return {"jobId": job_id, "state": "UNKNOWN"}


def list_jobs() -> list:
    base = os.path.join(job_root(), "jobs")
    if not os.path.isdir(base):
        return []
    return sorted(os.listdir(base))


def wait(job_id: str, timeout: float | None = None, poll: float = 0.2) -> dict:
    t0 = time.time()
    while True:
        st = status(job_id)
        if st.get("state") in TERMINAL:
            return st
        if timeout is not None and time.time() - t0 > timeout:
            return st
        time.sleep(poll)


def log_paths(job_id: str) -> list:
    d = os.path.join(job_dir(job_id), "logs")
    return sorted(os.path.join(d, f) for f in os.listdir(d)) if os.path.isdir(d) else []
