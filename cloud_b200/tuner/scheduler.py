"""Multi-GPU trial scheduler: packs parallel hyper-parameter trials onto the GPUs of one box.

The reference distributes a search by running K independent tuner processes that share a
``study_id`` and differ in ``tuner_id`` (tuner/tests/integration/tuner_integration_test.py:82-116,
281-294, with ``multiprocessing.Pool(4)``); the optimizer service is the only shared state.  This
scheduler is that pattern made B200-native: one worker process per GPU slice
(``CUDA_VISIBLE_DEVICES`` = ``gpus_per_trial`` devices), ``KERASTUNER_TUNER_ID=tuner<i>``, the
shared study living in the local file-backed optimizer store.  A worker that crashes has its
pending trial marked infeasible (``INVALID``) and its GPUs are handed to a replacement worker.

    sched = TrialScheduler(gpus_per_trial=1)
    results = sched.run(search_fn, num_workers=8)     # search_fn(tuner_id) builds a CloudTuner
                                                      # with a common study_id and calls .search()

Starting a worker costs far more than a short trial (interpreter + ``import torch`` + CUDA context +
loading the kernel library: seconds, against a 2 s MNIST trial), so workers can be kept warm:

    with WorkerPool(num_workers=8) as pool:           # one pre-warmed process per GPU slot
        sched.run(search_fn, num_workers=8, pool=pool)
        sched.run(other_search, num_workers=8, pool=pool)      # no start-up this time

A pool worker imports the framework, creates its CUDA context and loads the kernels once, then takes
jobs from a spool directory until the pool is closed (or its parent dies).  A worker that crashes
takes only its current job with it: the job reports a non-zero exit code (the scheduler marks the
trial infeasible and frees the slot) and the slot gets a fresh worker on its next job.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile
import time

import cloudpickle

from . import local_vizier
from ..launch import topology


class TrialScheduler:
    def __init__(self, num_gpus: int | None = None, gpus_per_trial: int = 1,
                 max_restarts: int = 1, poll: float = 0.1):
        self._concurrent = 1
        topo = topology.discover()
        self.num_gpus = topo.gpu_count if num_gpus is None else int(num_gpus)
        self.gpus_per_trial = max(int(gpus_per_trial), 1)
        self.max_restarts = max_restarts
        self.poll = poll
        self._pool = None

    def slots(self, num_workers: int | None = None) -> list:
        """GPU id lists, one per concurrent worker; a GPU-less box gets CPU slots."""
        if self.num_gpus == 0:
            return [[] for _ in range(num_workers or 1)]
        n = self.num_gpus // self.gpus_per_trial
        if num_workers:
            n = min(n, num_workers)
        return [list(range(i * self.gpus_per_trial, (i + 1) * self.gpus_per_trial)) for i in range(n)]

    def run(self, target, args=(), kwargs=None, num_workers: int | None = None,
            study: tuple | None = None, env: dict | None = None, timeout: float | None = None,
            log_dir: str | None = None, pool: "WorkerPool | None" = None) -> list:
        """Runs ``target(tuner_id, *args, **kwargs)`` in ``num_workers`` processes, at most one per
        GPU slot at a time.  ``study=(project_id, region, full_study_id)`` enables crash clean-up
        in the optimizer store.  Returns one record per worker."""
        if pool is not None:
            # the pool already holds its GPUs (and their leases) and its warm processes
            slots = pool.slots
            num_workers = num_workers or len(slots)
            self._concurrent = min(num_workers, len(slots))
            self._pool = pool
            try:
                return self._run(target, args, kwargs, num_workers, slots, study, env, timeout, log_dir)
            finally:
                self._pool = None
        slots = self.slots(num_workers)
        num_workers = num_workers or len(slots)
        self._concurrent = min(num_workers, len(slots))
        # the GPUs of the slots come from the box-wide lease table (launch/leases.py), so a search
        # and the jobs launched with tfc.run share the box without sharing a GPU
        lease_owner = None
        wanted = sum(len(s) for s in slots)
        if wanted and os.environ.get("TFC_GPU_LEASES", "1") != "0":
            from ..launch import leases
            lease_owner = "trial_scheduler_{}_{}".format(os.getpid(), int(time.time() * 1000) % 100000)
            # take the slots that are free right now (at least one): a search next to a running
            # job uses the rest of the box instead of waiting for all of it
            free = len(leases.snapshot(self.num_gpus)["free"])
            n_slots = min(len(slots), max(1, free // self.gpus_per_trial))
            got = leases.acquire(lease_owner, n_slots * self.gpus_per_trial, self.num_gpus,
                                 timeout=float(os.environ.get("TFC_QUEUE_TIMEOUT", "0") or 0) or None)
            slots = [got[i * self.gpus_per_trial:(i + 1) * self.gpus_per_trial] for i in range(n_slots)]
            self._concurrent = min(num_workers, len(slots))
        try:
            return self._run(target, args, kwargs, num_workers, slots, study, env, timeout, log_dir)
        finally:
            if lease_owner:
                from ..launch import leases
                leases.release(lease_owner)

    def _run(self, target, args, kwargs, num_workers, slots, study, env, timeout, log_dir) -> list:
        log_dir = log_dir or tempfile.mkdtemp(prefix="cb2_trials_")
        os.makedirs(log_dir, exist_ok=True)
        blob = os.path.join(log_dir, "target.pkl")
        # a function living in the user's own script/module travels by value: the worker process
        # cannot be assumed to be able to import that module
        mod = sys.modules.get(getattr(target, "__module__", None) or "")
        by_value = (mod is not None and mod.__name__ not in ("__main__",)
                    and not mod.__name__.startswith("cloud_b200")
                    and "site-packages" not in (getattr(mod, "__file__", "") or ""))
        if by_value:
            cloudpickle.register_pickle_by_value(mod)
        try:
            payload = cloudpickle.dumps((target, tuple(args), dict(kwargs or {})))
        finally:
            if by_value:
                cloudpickle.unregister_pickle_by_value(mod)
        with open(blob, "wb") as f:
            f.write(payload)
        pending = [{"index": i, "tuner_id": "tuner{}".format(i), "restarts": 0} for i in range(num_workers)]
        free = list(range(len(slots)))
        running, done = [], []
        t0 = time.time()
        while pending or running:
            while pending and free:
                job = pending.pop(0)
                slot = free.pop(0)
                job["slot"] = slot
                job.update(proc=self._spawn(job, slots[slot], blob, log_dir, env),
                           started=time.time())
                running.append(job)
            time.sleep(min(self.poll, 0.02) if self._pool is not None else self.poll)
            for job in list(running):
                rc = job["proc"].poll()
                if rc is None:
                    if timeout and time.time() - t0 > timeout:
                        job["proc"].kill()
                    continue
                running.remove(job)
                free.append(job["slot"])
                job["seconds"] = time.time() - job["started"]
                job["exit_code"] = rc
                if rc != 0:
                    self._mark_crashed(study, job["tuner_id"])
                    if job["restarts"] < self.max_restarts and not (timeout and time.time() - t0 > timeout):
                        pending.append({"index": job["index"], "tuner_id": job["tuner_id"],
                                        "restarts": job["restarts"] + 1})
                        continue
                done.append({k: job[k] for k in ("index", "tuner_id", "slot", "exit_code", "seconds", "restarts")}
                            | {"result": getattr(job["proc"], "result", None)}
                            | {"gpus": slots[job["slot"]],
                               "log": os.path.join(log_dir, job["tuner_id"] + ".log")})
        done.sort(key=lambda r: r["index"])
        return done

    def _spawn(self, job, gpu_ids, blob, log_dir, env):
        if self._pool is not None:
            return self._pool.submit(job["slot"], blob, job["tuner_id"], log_dir)
        e = dict(os.environ)
        e.update(env or {})
        e["KERASTUNER_TUNER_ID"] = job["tuner_id"]
        e["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in gpu_ids)
        e["TFC_TRIAL_SLOT"] = str(job["slot"])
        # workers share the host: give each its share of the cores instead of letting every
        # process start one OpenMP/MKL thread per core (8 workers x 8 spinning threads made a
        # packed search slower than a sequential one)
        share = str(max(1, (os.cpu_count() or 1) // max(1, self._concurrent)))
        for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
            if k not in os.environ and k not in (env or {}):
                e[k] = share
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):   # a worker is its own one-replica job
            e.pop(k, None)
        repo_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        e["PYTHONPATH"] = os.pathsep.join([repo_root, e.get("PYTHONPATH", "")])
        log = open(os.path.join(log_dir, job["tuner_id"] + ".log"), "ab", buffering=0)
        return subprocess.Popen([sys.executable, "-m", "cloud_b200.tuner.scheduler", blob,
                                 job["tuner_id"]], env=e, stdout=log, stderr=subprocess.STDOUT)

    @staticmethod
    def _mark_crashed(study, tuner_id):
        """A dead worker leaves an ACTIVE trial behind: complete it as infeasible so the study
        does not wait for it."""
        if not study:
            return
        project_id, region, study_id = study
        name = "projects/{}/locations/{}/studies/{}".format(project_id, region, study_id)
        svc = local_vizier.LocalVizier()
        try:
            for t in svc.list_trials(name).get("trials", []):
                if t.get("state") == "ACTIVE" and t.get("clientId") == tuner_id:
                    svc.complete_trial(t["name"], {"trial_infeasible": True,
                                                   "infeasible_reason": "worker crashed"})
        except local_vizier.HttpError:
            pass


def _worker_env(gpu_ids, slot, concurrent, env):
    e = dict(os.environ)
    e.update(env or {})
    e["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in gpu_ids)
    e["TFC_TRIAL_SLOT"] = str(slot)
    share = str(max(1, (os.cpu_count() or 1) // max(1, concurrent)))
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        if k not in os.environ and k not in (env or {}):
            e[k] = share
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    repo_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    e["PYTHONPATH"] = os.pathsep.join([repo_root, e.get("PYTHONPATH", "")])
    return e


class _PoolJob:
    """What ``TrialScheduler._run`` needs from a process handle: ``poll()`` and ``kill()``."""

    def __init__(self, pool, slot, path):
        self.pool, self.slot, self.path = pool, slot, path
        self.result = None                     # what the target returned (JSON-able), once done

    def poll(self):
        done = self.path + ".done"
        if os.path.exists(done):
            try:
                with open(done) as f:
                    rec = json.load(f)
                self.result = rec.get("result")
                return int(rec.get("exit_code", 1))
            except (ValueError, OSError):
                return None                    # still being written
        proc = self.pool.procs[self.slot]
        rc = proc.poll() if proc is not None else 1
        if rc is not None:                     # the worker died under this job
            self.pool.procs[self.slot] = None
            return rc if rc != 0 else 1
        return None

    def kill(self):
        proc = self.pool.procs[self.slot]
        if proc is not None:
            proc.kill()


class WorkerPool:
    """Pre-warmed trial workers, one per GPU slot, for any number of ``TrialScheduler.run`` calls.

    ``start_seconds`` is what the warm-up cost (all workers start in parallel); ``close()`` - or
    leaving the ``with`` block - stops them and returns the GPU leases."""

    def __init__(self, num_workers: int | None = None, num_gpus: int | None = None, gpus_per_trial: int = 1,
                 env: dict | None = None, log_dir: str | None = None, warm_timeout: float = 300.0):
        sched = TrialScheduler(num_gpus=num_gpus, gpus_per_trial=gpus_per_trial)
        self.slots = sched.slots(num_workers)
        self.env = dict(env or {})
        self.dir = log_dir or tempfile.mkdtemp(prefix="cb2_pool_")
        os.makedirs(self.dir, exist_ok=True)
        self._lease_owner = None
        wanted = sum(len(s) for s in self.slots)
        if wanted and os.environ.get("TFC_GPU_LEASES", "1") != "0":
            from ..launch import leases
            self._lease_owner = "trial_pool_{}_{}".format(os.getpid(), int(time.time() * 1000) % 100000)
            free = len(leases.snapshot(sched.num_gpus)["free"])
            n = min(len(self.slots), max(1, free // sched.gpus_per_trial))
            got = leases.acquire(self._lease_owner, n * sched.gpus_per_trial, sched.num_gpus,
                                 timeout=float(os.environ.get("TFC_QUEUE_TIMEOUT", "0") or 0) or None)
            self.slots = [got[i * sched.gpus_per_trial:(i + 1) * sched.gpus_per_trial] for i in range(n)]
        self.procs = [None] * len(self.slots)
        self._seq = 0
        t0 = time.time()
        try:
            for i in range(len(self.slots)):
                self._start(i)
            self._wait_ready(range(len(self.slots)), warm_timeout)
        except BaseException:
            self.close()
            raise
        self.start_seconds = time.time() - t0

    def _slot_dir(self, i):
        d = os.path.join(self.dir, "slot{}".format(i))
        os.makedirs(d, exist_ok=True)
        return d

    def _start(self, i):
        d = self._slot_dir(i)
        for name in ("ready", "stop"):
            try:
                os.unlink(os.path.join(d, name))
            except OSError:
                pass
        log = open(os.path.join(d, "worker.log"), "ab", buffering=0)
        self.procs[i] = subprocess.Popen(
            [sys.executable, "-m", "cloud_b200.tuner.scheduler", "--serve", d, str(os.getpid())],
            env=_worker_env(self.slots[i], i, len(self.slots), self.env), stdout=log, stderr=subprocess.STDOUT)

    def _wait_ready(self, which, timeout):
        t0 = time.time()
        for i in which:
            flag = os.path.join(self._slot_dir(i), "ready")
            while not os.path.exists(flag):
                if self.procs[i].poll() is not None:
                    raise RuntimeError("trial worker {} died while warming up (see {})".format(
                        i, os.path.join(self._slot_dir(i), "worker.log")))
                if time.time() - t0 > timeout:
                    raise TimeoutError("trial workers did not get ready within {} s".format(timeout))
                time.sleep(0.02)

    def submit(self, slot, blob, tuner_id, log_dir=None):
        if self.procs[slot] is None or self.procs[slot].poll() is not None:
            self._start(slot)                  # replacement for a worker that died: a cold start
            self._wait_ready([slot], 300.0)
        self._seq += 1
        path = os.path.join(self._slot_dir(slot), "job{:06d}".format(self._seq))
        spec = {"blob": blob, "tuner_id": tuner_id,
                "log": os.path.join(log_dir or self._slot_dir(slot), tuner_id + ".log")}
        with open(path + ".tmp", "w") as f:
            json.dump(spec, f)
        os.rename(path + ".tmp", path + ".job")    # appears atomically
        return _PoolJob(self, slot, path)

    def close(self):
        for i, proc in enumerate(self.procs):
            if proc is None:
                continue
            try:
                open(os.path.join(self._slot_dir(i), "stop"), "w").close()
            except OSError:
                pass
        deadline = time.time() + 5
        for proc in self.procs:
            if proc is None:
                continue
            try:
                proc.wait(timeout=max(0.1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                proc.kill()
        self.procs = [None] * len(self.procs)
        if self._lease_owner:
            from ..launch import leases
            leases.release(self._lease_owner)
            self._lease_owner = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def _serve_main(argv):
    """Body of a pool worker: warm up once, then run the jobs that appear in the slot directory."""
    import glob
    d, parent = argv[0], int(argv[1])
    t0 = time.time()
    warm = {}
    try:                                        # everything a trial would pay for on its first step
        import torch
        import cloud_b200  # noqa: F401
        from cloud_b200 import keras as _keras  # noqa: F401
        if torch.cuda.is_available():
            torch.cuda.init()
            torch.zeros(1, device="cuda").add_(1).item()
            from cloud_b200.ops import _lib
            _lib.kernels()
            _lib.runtime()
            warm["cuda"] = True
    except Exception as e:                      # a broken environment shows up in the first job instead
        warm["error"] = "{}: {}".format(type(e).__name__, e)
    warm["seconds"] = time.time() - t0
    with open(os.path.join(d, "ready.tmp"), "w") as f:
        json.dump(warm, f)
    os.rename(os.path.join(d, "ready.tmp"), os.path.join(d, "ready"))
    while True:
        if os.path.exists(os.path.join(d, "stop")):
            return 0
        try:
            os.kill(parent, 0)                  # the pool's owner is gone: nobody will send jobs
        except OSError:
            return 0
        jobs = sorted(glob.glob(os.path.join(d, "*.job")))
        if not jobs:
            time.sleep(0.01)
            continue
        job = jobs[0]
        base = job[:-4]
        with open(job) as f:
            spec = json.load(f)
        os.rename(job, base + ".running")
        rc, result = 0, None
        os.environ["KERASTUNER_TUNER_ID"] = spec["tuner_id"]
        log = open(spec["log"], "ab", buffering=0)
        out, err = os.dup(1), os.dup(2)
        os.dup2(log.fileno(), 1)
        os.dup2(log.fileno(), 2)
        try:
            with open(spec["blob"], "rb") as f:
                target, args, kwargs = cloudpickle.loads(f.read())
            result = target(spec["tuner_id"], *args, **kwargs)
        except SystemExit as e:
            rc = int(e.code or 0) if not isinstance(e.code, str) else 1
        except BaseException:
            import traceback
            traceback.print_exc()
            rc = 1
        finally:
            sys.stdout.flush()
            sys.stderr.flush()
            os.dup2(out, 1)
            os.dup2(err, 2)
            os.close(out)
            os.close(err)
            log.close()
        try:                                    # leave the GPU as a fresh worker would find it
            import gc
            import torch
            gc.collect()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()
        except Exception:
            pass
        try:
            json.dumps(result)
        except TypeError:
            result = None
        with open(base + ".done.tmp", "w") as f:
            json.dump({"exit_code": rc, "result": result}, f)
        os.rename(base + ".done.tmp", base + ".done")


def _worker_main(argv):
    blob, tuner_id = argv
    with open(blob, "rb") as f:
        target, args, kwargs = cloudpickle.loads(f.read())
    result = target(tuner_id, *args, **kwargs)
    if result is not None:
        try:
            print("TRIAL_WORKER_RESULT " + json.dumps(result))
        except TypeError:
            pass
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--serve":
        sys.exit(_serve_main(sys.argv[2:]))
    sys.exit(_worker_main(sys.argv[1:]))
