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
            log_dir: str | None = None) -> list:
        """Runs ``target(tuner_id, *args, **kwargs)`` in ``num_workers`` processes, at most one per
        GPU slot at a time.  ``study=(project_id, region, full_study_id)`` enables crash clean-up
        in the optimizer store.  Returns one record per worker."""
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
            time.sleep(self.poll)
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
                            | {"gpus": slots[job["slot"]],
                               "log": os.path.join(log_dir, job["tuner_id"] + ".log")})
        done.sort(key=lambda r: r["index"])
        return done

    def _spawn(self, job, gpu_ids, blob, log_dir, env):
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
    sys.exit(_worker_main(sys.argv[1:]))
