"""Port of tuner/tests/unit/tuner_test.py (autospecced _OptimizerClient = fake oracle backend) plus
the integration flows of tuner/tests/integration/tuner_integration_test.py on the local optimizer:
constructor matrix, create_trial (initial / below max / at max / after STOPPING), update_trial,
end_trial (ok / infeasible / unknown id / bad status), best-trial ordering, search with printed
summaries, distributed tuning with several processes on one study."""
import os
import re
from unittest import mock

import numpy as np
import pytest

import cloud_b200 as tfc
from cloud_b200.tuner import optimizer_client, tuner as tuner_mod, utils
from cloud_b200.tuner.engine import hyperparameters as hp_module
from cloud_b200.tuner.engine import oracle as oracle_module
from cloud_b200.tuner.engine import trial as trial_module

keras = tfc.keras

STUDY_CONFIG = {
    "algorithm": "ALGORITHM_UNSPECIFIED",
    "automatedStoppingConfig": {"decayCurveStoppingConfig": {"useElapsedTime": True}},
    "metrics": [{"metric": "val_acc", "goal": "MAXIMIZE"}],
    "parameters": [{"parameter": "learning_rate", "type": "DISCRETE",
                    "discrete_value_spec": {"values": [0.0001, 0.001, 0.01]}}],
}


def _hps():
    hps = hp_module.HyperParameters()
    hps.Choice("learning_rate", [1e-4, 1e-3, 1e-2])
    return hps


@pytest.fixture(autouse=True)
def _env(monkeypatch, tmp_path):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.setenv("TFC_VIZIER_ROOT", str(tmp_path / "vizier"))


@pytest.fixture
def fake_service():
    svc = mock.create_autospec(optimizer_client._OptimizerClient, instance=True, spec_set=True)
    with mock.patch.object(optimizer_client, "create_or_load_study", return_value=svc) as factory:
        yield svc, factory


def _oracle(**kw):
    args = dict(project_id="project", region="us-central1", objective="val_acc",
                hyperparameters=_hps(), max_trials=10, study_id="s1")
    args.update(kw)
    return tuner_mod.CloudOracle(**args)


def test_constructor_with_objective_and_hps(fake_service):
    svc, factory = fake_service
    o = _oracle()
    assert o.study_id == "CloudTuner_study_s1"
    assert o.objective == [oracle_module.Objective("val_acc", "max")]
    factory.assert_called_once_with("project", "us-central1", "CloudTuner_study_s1", STUDY_CONFIG)
    assert not o.allow_new_entries and not o.tune_new_entries


def test_constructor_with_study_config(fake_service):
    _svc, factory = fake_service
    o = _oracle(objective=None, hyperparameters=None, study_config=STUDY_CONFIG)
    assert o.hyperparameters.space[0].values == [0.0001, 0.001, 0.01]
    factory.assert_called_once_with("project", "us-central1", "CloudTuner_study_s1", STUDY_CONFIG)


def test_constructor_default_study_id_is_a_timestamp(fake_service):
    o = _oracle(study_id=None)
    assert re.fullmatch(r"CloudTuner_study_\d{8}_\d{6}", o.study_id)


@pytest.mark.parametrize("kw,regex", [
    (dict(study_config=STUDY_CONFIG), r"Please configure either study_config or"),
    (dict(objective=None), r"If study_config is not set, objective and hyperparameters must be set"),
    (dict(hyperparameters=None), r"If study_config is not set"),
    (dict(project_id=None), r'"project_id" is not found'),
    (dict(region=None), r'"region" is not found'),
])
def test_constructor_errors(fake_service, kw, regex):
    with pytest.raises(ValueError, match=regex):
        _oracle(**kw)


def _suggestion(tid="1", lr=0.001):
    return {"trials": [{"name": "projects/p/locations/r/studies/s/trials/" + tid, "state": "ACTIVE",
                        "parameters": [{"parameter": "learning_rate", "floatValue": lr}]}]}


def test_create_trial_initial_and_below_max(fake_service):
    svc, _ = fake_service
    o = _oracle()
    svc.list_trials.return_value = []
    svc.get_suggestions.return_value = _suggestion("1")
    t = o.create_trial("tuner_0")
    svc.get_suggestions.assert_called_once_with("tuner_0")
    assert t.trial_id == "1" and t.status == trial_module.TrialStatus.RUNNING
    assert t.hyperparameters.values == {"learning_rate": 0.001}
    assert o.ongoing_trials["tuner_0"] is t and o.trials["1"] is t
    svc.list_trials.return_value = [{"name": "a", "state": "ACTIVE"}] * 9
    svc.get_suggestions.return_value = _suggestion("10", 0.01)
    assert o.create_trial("tuner_1").trial_id == "10"


def test_create_trial_at_max_and_after_stopping(fake_service):
    svc, _ = fake_service
    o = _oracle()
    svc.list_trials.return_value = [{"name": "a", "state": "COMPLETED"}] * 10
    t = o.create_trial("tuner_0")
    assert t.status == trial_module.TrialStatus.STOPPED and t.hyperparameters.values is None
    svc.get_suggestions.assert_not_called()
    svc.list_trials.return_value = [{"name": "a", "state": "STOPPING"}]
    assert o.create_trial("tuner_0").status == trial_module.TrialStatus.STOPPED
    svc.list_trials.return_value = []
    svc.get_suggestions.return_value = {}
    assert o.create_trial("tuner_0").status == trial_module.TrialStatus.STOPPED


def test_update_trial_reports_and_polls(fake_service):
    svc, _ = fake_service
    o = _oracle()
    svc.list_trials.return_value = []
    svc.get_suggestions.return_value = _suggestion("1")
    with mock.patch("time.time", return_value=1000.0):
        o.create_trial("tuner_0")
    svc.should_trial_stop.return_value = False
    with mock.patch("time.time", return_value=1011.5):
        status = o.update_trial("1", {"val_acc": 0.8, "loss": 1.0}, step=3)
    svc.report_intermediate_objective_value.assert_called_once_with(
        3, 11.5, [{"metric": "val_acc", "value": 0.8}], "1")
    svc.should_trial_stop.assert_called_once_with("1")
    assert status == trial_module.TrialStatus.RUNNING
    svc.should_trial_stop.return_value = True
    assert o.update_trial("1", {"val_acc": 0.8}, step=4) == trial_module.TrialStatus.STOPPED


def test_end_trial_paths(fake_service):
    svc, _ = fake_service
    o = _oracle()
    svc.list_trials.return_value = []
    svc.get_suggestions.return_value = _suggestion("1")
    o.create_trial("tuner_0")
    svc.complete_trial.return_value = {"name": "x/trials/1", "state": "COMPLETED",
                                       "finalMeasurement": {"stepCount": 3, "metrics": [{"metric": "val_acc", "value": 0.9}]}}
    o.end_trial("1")
    svc.complete_trial.assert_called_once_with("1", False, None)
    assert o.trials["1"].score == 0.9 and o.trials["1"].best_step == 3 and not o.ongoing_trials
    svc.get_suggestions.return_value = _suggestion("2")
    o.create_trial("tuner_0")
    o.end_trial("2", "INVALID")
    svc.complete_trial.assert_called_with("2", True, "INVALID")
    with pytest.raises(ValueError, match=r"Ongoing trial with id: 9 not found"):
        o.end_trial("9")
    svc.get_suggestions.return_value = _suggestion("3")
    o.create_trial("tuner_0")
    with pytest.raises(ValueError, match=r"Unexpected status passed"):
        o.end_trial("3", "FOO")


def test_get_best_trials_ordering(fake_service):
    svc, _ = fake_service

    def tr(i, v, state="COMPLETED"):
        return {"name": "x/trials/%d" % i, "state": state,
                "parameters": [{"parameter": "learning_rate", "floatValue": 0.001 * i}],
                "finalMeasurement": {"stepCount": i, "metrics": [{"metric": "val_acc", "value": v}]}}
    svc.list_trials.return_value = [tr(1, 0.7), tr(2, 0.9), tr(3, 0.8), tr(4, 0.99, "ACTIVE")]
    best = _oracle().get_best_trials(2)
    assert [(t.trial_id, t.score, t.best_step) for t in best] == [("2", 0.9, 2), ("3", 0.8, 3)]
    assert all(t.status == trial_module.TrialStatus.COMPLETED for t in best)
    minimizing = _oracle(objective=oracle_module.Objective("val_acc", "min"))
    assert [t.trial_id for t in minimizing.get_best_trials(3)] == ["1", "3", "2"]
    svc.list_trials.return_value = []
    assert _oracle().get_best_trials(1) == []
    multi = _oracle(objective=["val_acc", "val_loss"])
    with pytest.raises(ValueError, match=r"multi-objective optimization is not supported"):
        multi.get_best_trials(1)


def test_cloud_tuner_wires_a_cloud_oracle(fake_service, tmp_path):
    t = tfc.CloudTuner(lambda hp: None, project_id="p", region="r", objective="val_acc",
                       hyperparameters=_hps(), max_trials=3, study_id="s", directory=str(tmp_path))
    assert isinstance(t.oracle, tfc.CloudOracle) and t.oracle.max_trials == 3
    assert t.tuner_id == "tuner0"
    t.tuner_id = "tuner7"
    assert t.tuner_id == "tuner7"


# ---- real local optimizer: the integration flows -------------------------------------------------
def _data(n=256):
    rng = np.random.RandomState(0)
    x = rng.rand(n, 8, 8).astype("float32")
    y = (x.reshape(n, -1)[:, :4].argmax(1)).astype("int64")
    return x, y


def _build_model(hp):
    m = keras.Sequential()
    m.add(keras.layers.Flatten(input_shape=(8, 8)))
    # CloudTuner does not support declaring hyperparameters inside build(): use hp.get
    for _ in range(hp.get("num_layers")):
        m.add(keras.layers.Dense(16, activation="relu"))
    m.add(keras.layers.Dense(4, activation="softmax"))
    m.compile(optimizer=keras.optimizers.Adam(lr=hp.get("learning_rate")),
              loss="sparse_categorical_crossentropy", metrics=["acc"])
    return m


def _search_space():
    hps = hp_module.HyperParameters()
    hps.Float("learning_rate", min_value=1e-4, max_value=1e-2, sampling="log")
    hps.Int("num_layers", 2, 4)
    return hps


def test_search_with_hps_prints_summaries_and_returns_best_models(tmp_path, capsys):
    x, y = _data()
    t = tfc.CloudTuner(_build_model, project_id="p", region="r", objective="acc",
                       hyperparameters=_search_space(), max_trials=3, study_id="it1",
                       directory=str(tmp_path / "t"))
    t.search_space_summary()
    out = capsys.readouterr().out
    assert re.search(r"Search space summary", out) and re.search(r"learning_rate \(Float\)", out)
    assert re.search(r"num_layers \(Int\)", out)
    t.search(x=x, y=y, epochs=2, steps_per_epoch=4, validation_data=(x, y), verbose=0)
    t.results_summary()
    out = capsys.readouterr().out
    assert re.search(r"Results summary", out) and re.search(r"Trial summary", out)
    assert re.search(r"Hyperparameters", out)
    best = t.oracle.get_best_trials(3)
    assert len(best) == 3 and best[0].score >= best[-1].score
    model = t.get_best_models(1)[0]
    assert model.evaluate(x, y, verbose=0)[1] == pytest.approx(best[0].score, abs=0.15)
    assert len(t.get_best_hyperparameters(2)) == 2
    assert os.path.exists(tmp_path / "t" / "untitled_project" / "oracle.json")


def test_search_with_study_config_and_dataset(tmp_path):
    x, y = _data()
    ds = tfc.data.Dataset.from_tensor_slices((x, y)).batch(64)
    cfg = utils.make_study_config("acc", _search_space())
    t = tfc.CloudTuner(_build_model, project_id="p", region="r", study_config=cfg, max_trials=2,
                       study_id="it2", directory=str(tmp_path / "t"))
    t.search(x=ds, epochs=1, verbose=0)
    assert len(t.oracle.get_best_trials(5)) == 2


def test_hyperparameters_cannot_be_declared_in_build(tmp_path):
    def build(hp):
        hp.Int("units", 1, 4)          # not declared up-front
        return _build_model(hp)
    t = tfc.CloudTuner(build, project_id="p", region="r", objective="acc",
                       hyperparameters=_search_space(), max_trials=1, study_id="it3",
                       directory=str(tmp_path / "t"))
    assert [p.name for p in t.oracle.get_space().space] == ["learning_rate", "num_layers"]


def _slow_build_model(hp):
    import time
    time.sleep(0.6)          # a trial long enough that late-starting workers still find work
    return _build_model(hp)


def _distributed_worker(tuner_id, directory, study_id):
    import cloud_b200 as tfc2
    x, y = _data(128)
    t = tfc2.CloudTuner(_slow_build_model, project_id="p", region="r", objective="acc",
                        hyperparameters=_search_space(), max_trials=6, study_id=study_id,
                        directory=os.path.join(directory, tuner_id))
    t.tuner_id = tuner_id
    t.search(x=x, y=y, epochs=1, steps_per_epoch=2, verbose=0)
    return {"tuner": tuner_id, "trials": sorted(t.oracle.trials)}


def test_distributed_tuning_shares_one_study(tmp_path):
    """Several tuner processes, one study id (tuner_integration_test.py:82-116,279-294), packed by
    the trial scheduler (CPU slots here; one GPU per worker on the B200 box)."""
    from cloud_b200.tuner import TrialScheduler, local_vizier
    sched = TrialScheduler(num_gpus=0)
    res = sched.run(_distributed_worker, args=(str(tmp_path / "dist"), "shared"), num_workers=3,
                    study=("p", "r", "CloudTuner_study_shared"), timeout=240,
                    log_dir=str(tmp_path / "logs"))
    assert [r["exit_code"] for r in res] == [0, 0, 0], [open(r["log"]).read()[-800:] for r in res]
    trials = local_vizier.LocalVizier().list_trials(
        "projects/p/locations/r/studies/CloudTuner_study_shared")["trials"]
    assert 6 <= len(trials) <= 8                      # max_trials, +workers-1 for the list/suggest race
    assert all(t["state"] == "COMPLETED" for t in trials)
    # the workers shared the study (a late starter may find every trial already taken)
    assert len({t["clientId"] for t in trials}) >= 2
    assert len({t["name"] for t in trials}) == len(trials)


def test_crashed_worker_trial_is_marked_infeasible(tmp_path):
    from cloud_b200.tuner import TrialScheduler, local_vizier

    def crashing(tuner_id, directory):
        import cloud_b200 as tfc2
        t = tfc2.CloudTuner(_build_model, project_id="p", region="r", objective="acc",
                            hyperparameters=_search_space(), max_trials=2, study_id="crash",
                            directory=os.path.join(directory, tuner_id))
        t.oracle.create_trial(tuner_id)       # leaves an ACTIVE trial behind
        os._exit(9)
    res = TrialScheduler(num_gpus=0, max_restarts=0).run(
        crashing, args=(str(tmp_path / "d"),), num_workers=1,
        study=("p", "r", "CloudTuner_study_crash"), log_dir=str(tmp_path / "logs"))
    assert res[0]["exit_code"] == 9
    trials = local_vizier.LocalVizier().list_trials(
        "projects/p/locations/r/studies/CloudTuner_study_crash")["trials"]
    assert trials[0]["state"] == "COMPLETED" and trials[0]["trialInfeasible"] is True


def test_distributing_cloud_tuner_trains_every_trial_as_a_launched_job(tmp_path, monkeypatch):
    """T5: each trial is shipped through cloud_fit to its own job; the tuner replays the job's
    per-epoch metrics into the oracle and can load the trained model back."""
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.setenv("TFC_VIZIER_ROOT", str(tmp_path / "vizier"))
    x, y = _data(128)
    directory = tmp_path / "dist_tuner"
    t = tfc.DistributingCloudTuner(_build_model, project_id="p", region="r", directory=str(directory),
                                   objective="acc", hyperparameters=_search_space(), max_trials=2,
                                   study_id="dct", replica_config=tfc.COMMON_MACHINE_CONFIGS["CPU"],
                                   replica_count=1)
    t.search(x=x, y=y, epochs=2, batch_size=32, verbose=0)
    trials = t.oracle.get_best_trials(5)
    assert len(trials) == 2
    for tr in trials:
        assert tr.status == "COMPLETED" and tr.score is not None
        tdir = directory / tr.trial_id
        assert (tdir / "logs" / "train" / "scalars.jsonl").exists()       # TensorBoard callback ran in the job
        assert (tdir / "output").exists()                                 # cloud_fit wrote the trained model
    from cloud_b200.tuner import local_vizier
    stored = local_vizier.LocalVizier().list_trials("projects/p/locations/r/studies/CloudTuner_study_dct")["trials"]
    assert len(stored) == 2
    assert all(len(s["measurements"]) == 2 for s in stored)               # one report per epoch of the job
    model = t.load_model(trials[0])
    assert model.predict(x[:4]).shape == (4, 4)
    # the jobs went through the launcher's registry
    from cloud_b200.launch import registry
    jobs = [j for j in os.listdir(tmp_path / "jobs" / "jobs")]
    assert len(jobs) == 2 and all(registry.status(j)["state"] == "SUCCEEDED" for j in jobs)


def test_distributing_cloud_tuner_fails_the_trial_when_its_job_fails(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.setenv("TFC_VIZIER_ROOT", str(tmp_path / "vizier"))
    monkeypatch.setenv("TFC_FAULT", "0:1")              # rank 0 dies at step 1 of every job
    x, y = _data(128)
    t = tfc.DistributingCloudTuner(_build_model, project_id="p", region="r", directory=str(tmp_path / "d"),
                                   objective="acc", hyperparameters=_search_space(), max_trials=1,
                                   study_id="dct_fail", replica_config=tfc.COMMON_MACHINE_CONFIGS["CPU"])
    with pytest.raises(RuntimeError, match="AIP Training job failed"):
        t.search(x=x, y=y, epochs=1, batch_size=32, verbose=0)
