"""Per-job supervisor process: ``python -m cloud_b200.launch.supervisor <job_id>``.

Spawns one rank per planned replica (one process per GPU), each with
``RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT / TF_CONFIG / TFC_RANK_IN_TASK`` and the
job's GPU subset in ``CUDA_VISIBLE_DEVICES``; tees every rank's stdout+stderr to
``logs/rank<r>.log``; kills the whole group as soon as one rank fails and records the failing rank
with the tail of its log; honours ``TFC_JOB_TIMEOUT``.  This is what AI Platform Training did for
the reference after ``jobs.create`` (deploy.py:82-88); failure detection is new.
"""
from __future__ import annotations

import os
import signal
import socket
import subprocess
import sys
import time

from . import leases, registry, topology


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tail(path: str, n: int = 30) -> str:
    try:
        with open(path, "rb") as f:
            f.seek(0, os.SEEK_END)
            size = f.tell()
            f.seek(max(0, size - 8192))
            return "\n".join(f.read().decode(errors="replace").splitlines()[-n:])
    except OSError:
        return ""


def _lease_gpus(job_id: str, n: int, cancelled: dict):
    """Physical GPU ids for the job, or None after recording why it did not get them."""
    if os.environ.get("TFC_GPU_LEASES", "1") == "0":
        return list(range(n))
    total = topology.gpu_count()
    if n > total:
        registry.set_status(job_id, "FAILED", error_tail="the job needs {} GPUs, the box has {}".format(n, total))
        return None
    wait = float(os.environ.get("TFC_QUEUE_TIMEOUT", "0") or 0) or None
    try:
        got = leases.acquire(job_id, n, total, timeout=wait, should_stop=lambda: cancelled["flag"],
                             on_wait=lambda: registry.set_status(job_id, "QUEUED", supervisor_pid=os.getpid(),
                                                                 waiting_for_gpus=n))
    except leases.LeaseTimeout as e:
        registry.set_status(job_id, "FAILED", error_tail=str(e))
        return None
    if got is None:
        registry.set_status(job_id, "CANCELLED", exit_codes=[], seconds=0.0)
    return got


def run_job(job_id: str) -> int:
    spec = registry.get(job_id)
    jdir = registry.job_dir(job_id)
    plan = spec["plan"]
    image = spec["image"]
    workdir = image["workdir"]
    cmd = [spec.get("python") or sys.executable] + image["entrypoint"][1:] + list(spec.get("args") or [])
    port = _free_port()
    base_env = dict(os.environ)
    base_env.update(spec.get("env") or {})
    base_env["MASTER_ADDR"] = "127.0.0.1"
    base_env["MASTER_PORT"] = str(port)
    base_env["WORLD_SIZE"] = str(plan["world_size"])
    base_env["TF_KERAS_RUNNING_REMOTELY"] = "1"
    base_env["TFC_JOB_ID"] = job_id
    base_env["TFC_JOB_DIR"] = jdir
    base_env["PYTHONUNBUFFERED"] = "1"
    # ranks share the host: each gets its share of the cores for OpenMP/MKL (torchrun pins this to
    # 1; a data loader thread plus the training thread want a little more when cores allow)
    share = str(max(1, (os.cpu_count() or 1) // max(1, int(plan["world_size"]))))
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        base_env.setdefault(k, share)
    repo_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    base_env["PYTHONPATH"] = os.pathsep.join([repo_root, workdir, base_env.get("PYTHONPATH", "")])
    cancelled = {"flag": False}

    def _on_term(signum, frame):
        cancelled["flag"] = True
    signal.signal(signal.SIGTERM, _on_term)
    signal.signal(signal.SIGINT, _on_term)
    if plan["uses_gpu"]:
        # the plan numbers the job's GPUs 0..n-1; which physical GPUs those are is decided here,
        # under the box-wide lease table, so concurrent jobs never share a GPU.  Until enough GPUs
        # are free the job stays QUEUED (FIFO); leases die with this process.
        gpu_ids = _lease_gpus(job_id, len(plan["gpu_ids"]), cancelled)
        if gpu_ids is None:
            return 1
        base_env["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in gpu_ids)
    else:
        gpu_ids = []
        base_env["CUDA_VISIBLE_DEVICES"] = ""
    timeout = float(base_env.get("TFC_JOB_TIMEOUT", "0") or 0)      # own environment or the job's `env`
    # Heartbeats: the training engine touches <job_dir>/heartbeat/rank<r>
    # after every step.  A rank that has started beating and then stays silent for longer than
    # TFC_HEARTBEAT_TIMEOUT seconds (0 = off) - a hung collective, a wedged data loader - fails
    # the job instead of holding the GPUs forever.
    hb_timeout = float(base_env.get("TFC_HEARTBEAT_TIMEOUT", "0") or 0)
    hb_dir = os.path.join(jdir, "heartbeat")
    os.makedirs(hb_dir, exist_ok=True)
    # Restarts: after a failed rank the whole group is started again, up to TFC_JOB_RESTARTS times
    # (default 0, like AI Platform).  Training resumes where a `BackupAndRestore` callback left it;
    # ranks see the attempt number in TFC_RESTART_COUNT.
    max_restarts = int(base_env.get("TFC_JOB_RESTARTS", "0") or spec.get("max_restarts", 0) or 0)
    t0 = time.time()
    attempt = 0
    failures = []
    while True:
        for name in os.listdir(hb_dir):                  # beats of a previous attempt say nothing
            os.remove(os.path.join(hb_dir, name))
        base_env["MASTER_PORT"] = str(port if attempt == 0 else _free_port())
        base_env["TFC_RESTART_COUNT"] = str(attempt)
        registry.set_status(job_id, "RUNNING", supervisor_pid=os.getpid(),
                            master_port=int(base_env["MASTER_PORT"]), gpus=gpu_ids, restarts=attempt,
                            failures=failures)
        procs, logs = [], []
        for r in plan["ranks"]:
            env = dict(base_env)
            env["RANK"] = str(r["rank"])
            env["LOCAL_RANK"] = str(r["local_rank"])
            env["TF_CONFIG"] = r["tf_config"]
            env["TFC_RANK_IN_TASK"] = str(r["rank_in_task"])
            env["TFC_TASK_TYPE"] = r["task_type"]
            log_path = os.path.join(jdir, "logs", "rank{}.log".format(r["rank"]))
            fh = open(log_path, "ab", buffering=0)
            if attempt:
                fh.write("---- restart {} of {} ----\n".format(attempt, max_restarts).encode())
            logs.append((fh, log_path))
            procs.append(subprocess.Popen(cmd, cwd=workdir, env=env, stdout=fh, stderr=subprocess.STDOUT,
                                          start_new_session=True))
        failed = _watch(procs, cancelled, timeout, t0, hb_timeout, hb_dir)
        if failed is not None or cancelled["flag"]:
            _kill_group(procs)
        for fh, _ in logs:
            fh.close()
        retry = (failed is not None and failed[0] >= 0 and not cancelled["flag"] and attempt < max_restarts)
        if not retry:
            break
        failures.append({"attempt": attempt, "rank": failed[0], "exit_code": failed[1],
                         "error_tail": _tail(logs[failed[0]][1], 8)})
        attempt += 1
    if gpu_ids:
        leases.release(job_id)
    exit_codes = [p.poll() for p in procs]
    common = dict(exit_codes=exit_codes, seconds=time.time() - t0, gpus=gpu_ids, restarts=attempt,
                  failures=failures)
    if cancelled["flag"] and failed is None:
        registry.set_status(job_id, "CANCELLED", **common)
        return 130
    if failed is not None:
        rank, rc = failed
        tail = _tail(logs[rank][1]) if rank >= 0 else ""
        registry.set_status(job_id, "FAILED", failed_rank=rank, exit_code=rc, error_tail=tail, **common)
        return 1
    registry.set_status(job_id, "SUCCEEDED", **common)
    return 0


def _watch(procs, cancelled, timeout, t0, hb_timeout, hb_dir):
    """Blocks until every rank has exited, one has failed, the job is cancelled or out of time.
    Returns None or (rank, exit code or reason); rank -1 = the job as a whole."""
    while True:
        alive, failed = 0, None
        for i, p in enumerate(procs):
            rc = p.poll()
            if rc is None:
                alive += 1
            elif rc != 0 and failed is None:
                failed = (i, rc)
        if failed is not None or alive == 0 or cancelled["flag"]:
            return failed
        if timeout and time.time() - t0 > timeout:
            return (-1, "timeout after {}s".format(timeout))
        if hb_timeout:
            now = time.time()
            for i, p in enumerate(procs):
                hb = os.path.join(hb_dir, "rank{}".format(i))
                if p.poll() is None and os.path.exists(hb) and now - os.path.getmtime(hb) > hb_timeout:
                    return (i, "no heartbeat for {:.0f}s (stalled)".format(now - os.path.getmtime(hb)))
        time.sleep(0.1)


def _kill_group(procs) -> None:
    for p in procs:
        if p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGTERM)
            except (ProcessLookupError, PermissionError):
                pass
    deadline = time.time() + 5
    for p in procs:
        try:
            p.wait(timeout=max(0.1, deadline - time.time()))
        except subprocess.TimeoutExpired:
            try:
                os.killpg(p.pid, signal.SIGKILL)
            except (ProcessLookupError, PermissionError):
                pass
            p.wait()


def main(argv=None) -> int:
    argv = argv if argv is not None else sys.argv[1:]
    if len(argv) != 1:
        print("usage: python -m cloud_b200.launch.supervisor <job_id>", file=sys.stderr)
        return 2
    try:
        return run_job(argv[0])
    except Exception as e:  # the supervisor itself must never leave a job RUNNING forever
        registry.set_status(argv[0], "FAILED", error_tail="supervisor error: {!r}".format(e))
        raise


if __name__ == "__main__":
    sys.exit(main())
