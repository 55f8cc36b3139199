"""The pre-warmed trial worker pool (tuner/scheduler.py::WorkerPool): workers survive from one
search to the next, a crash costs one job and the slot gets a fresh worker, trials of a crashed
worker are marked infeasible.  CPU slots here; one GPU per worker on the B200 box.
(Reference pattern: tuner/tests/integration/tuner_integration_test.py:48,279-294 - a process pool of
tuners sharing one study.)"""
import os

from cloud_b200.tuner import TrialScheduler, WorkerPool, local_vizier


def _pid(tuner_id, tag):
    import os as _os
    return {"tuner": tuner_id, "pid": _os.getpid(), "tag": tag, "env_id": _os.environ.get("KERASTUNER_TUNER_ID")}


def _results(records):
    out = []
    for r in records:
        text = open(r["log"]).read() if os.path.exists(r["log"]) else ""
        out.append(text)
    return out


def test_pool_workers_are_reused_across_runs(tmp_path):
    import json
    with WorkerPool(num_workers=2, num_gpus=0, log_dir=str(tmp_path / "pool")) as pool:
        assert len(pool.slots) == 2 and pool.start_seconds < 120
        pids = [p.pid for p in pool.procs]
        sched = TrialScheduler(num_gpus=0)
        a = sched.run(_pid, args=("first",), num_workers=2, pool=pool, log_dir=str(tmp_path / "a"))
        b = sched.run(_pid, args=("second",), num_workers=4, pool=pool, log_dir=str(tmp_path / "b"))
        assert [r["exit_code"] for r in a] == [0, 0] and [r["exit_code"] for r in b] == [0, 0, 0, 0]
        assert [p.pid for p in pool.procs] == pids                 # nobody was restarted
        # the jobs really ran inside the two warm processes, each under its own tuner id
        seen = set()
        for d in ("slot0", "slot1"):
            for name in sorted(os.listdir(tmp_path / "pool" / d)):
                if name.endswith(".done"):
                    rec = json.load(open(tmp_path / "pool" / d / name))
                    assert rec["exit_code"] == 0
                    seen.add(rec["result"]["pid"])
                    assert rec["result"]["env_id"] == rec["result"]["tuner"]
        assert seen == set(pids)
    assert all(p is None for p in pool.procs)                       # closed: workers gone


def _crash_one(tuner_id, directory):
    import os as _os
    import cloud_b200 as tfc2
    from cloud_b200.tuner import HyperParameters

    def build(hp):
        m = tfc2.keras.Sequential([tfc2.keras.layers.Dense(4, activation="softmax", input_shape=(8,))])
        m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["acc"])
        return m
    hps = HyperParameters()
    hps.Float("learning_rate", min_value=1e-4, max_value=1e-2, sampling="log")
    t = tfc2.CloudTuner(build, project_id="p", region="r", objective="acc",
                        hyperparameters=hps, max_trials=4, study_id="poolcrash",
                        directory=_os.path.join(directory, tuner_id))
    t.oracle.create_trial(tuner_id)           # leaves an ACTIVE trial behind
    if tuner_id == "tuner0":
        _os._exit(7)
    return {"ok": tuner_id}


def test_a_crashing_job_costs_one_worker_and_its_trial_is_marked_infeasible(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_VIZIER_ROOT", str(tmp_path / "vizier"))
    with WorkerPool(num_workers=2, num_gpus=0, log_dir=str(tmp_path / "pool"),
                    env={"TFC_VIZIER_ROOT": str(tmp_path / "vizier")}) as pool:
        before = [p.pid for p in pool.procs]
        res = TrialScheduler(num_gpus=0, max_restarts=0).run(
            _crash_one, args=(str(tmp_path / "d"),), num_workers=2, pool=pool,
            study=("p", "r", "CloudTuner_study_poolcrash"), log_dir=str(tmp_path / "logs"))
        codes = {r["tuner_id"]: r["exit_code"] for r in res}
        assert codes["tuner0"] == 7 and codes["tuner1"] == 0
        # the slot of the dead worker is usable again (a fresh process), the other one kept its worker
        again = TrialScheduler(num_gpus=0).run(_pid, args=("after",), num_workers=2, pool=pool,
                                               log_dir=str(tmp_path / "after"))
        assert [r["exit_code"] for r in again] == [0, 0]
        after = [p.pid for p in pool.procs]
        assert sum(a == b for a, b in zip(before, after)) == 1
    trials = local_vizier.LocalVizier().list_trials(
        "projects/p/locations/r/studies/CloudTuner_study_poolcrash")["trials"]
    dead = [t for t in trials if t.get("clientId") == "tuner0"]
    assert dead and dead[0]["state"] == "COMPLETED" and dead[0]["trialInfeasible"] is True
