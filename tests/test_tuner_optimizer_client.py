"""Port of tuner/tests/unit/optimizer_client_test.py against the REAL on-disk local optimizer
(upstream uses a MagicMock discovery resource): create / 409-load / 409-retry failure, suggestion
stickiness per client id, long-running operations, 429 after the trial cap, addMeasurement,
early-stop -> stop, complete, list, delete / 404, inactive study."""
import pytest

from cloud_b200.tuner import constants, local_vizier, optimizer_client

CONFIG = {
    "algorithm": "ALGORITHM_UNSPECIFIED",
    "automatedStoppingConfig": {"decayCurveStoppingConfig": {"useElapsedTime": True}},
    "metrics": [{"metric": "val_acc", "goal": "MAXIMIZE"}],
    "parameters": [
        {"parameter": "learning_rate", "type": "DOUBLE", "scale_type": "UNIT_LOG_SCALE",
         "double_value_spec": {"min_value": 1e-4, "max_value": 1e-2}},
        {"parameter": "units", "type": "INTEGER", "integer_value_spec": {"min_value": 2, "max_value": 10}},
        {"parameter": "act", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["relu", "elu"]}},
    ],
}


@pytest.fixture
def svc(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_VIZIER_ROOT", str(tmp_path / "vizier"))
    monkeypatch.setenv("TFC_STUDY_RETRY_SECS", "0")
    return local_vizier.LocalVizier()


def _client(svc, study="study-a"):
    return optimizer_client.create_or_load_study("proj", "us-central1", study, CONFIG, service_client=svc)


def test_create_then_load_existing_study(svc):
    c1 = _client(svc)
    c2 = _client(svc)            # 409 -> get
    assert c1.study_id == c2.study_id == "study-a"
    assert [s["name"] for s in c1.list_studies()] == ["projects/proj/locations/us-central1/studies/study-a"]
    assert svc.get_study(c1._make_study_name())["studyConfig"] == CONFIG


def test_create_409_then_get_keeps_failing(svc, monkeypatch):
    _client(svc)
    calls = {"n": 0}

    def broken(name):
        calls["n"] += 1
        raise local_vizier.HttpError(500, "backend down")
    monkeypatch.setattr(svc, "get_study", broken)
    with pytest.raises(RuntimeError, match=r"GetStudy wasn't successful after 3 tries"):
        _client(svc)
    assert calls["n"] == constants.NUM_TRIES_FOR_STUDIES


def test_create_other_http_error_propagates(svc, monkeypatch):
    monkeypatch.setattr(svc, "create_study",
                        lambda *a, **k: (_ for _ in ()).throw(local_vizier.HttpError(400, "bad")))
    with pytest.raises(local_vizier.HttpError):
        _client(svc)


def test_client_needs_study_id(svc):
    with pytest.raises(ValueError, match=r"Use create_or_load_study\(\) instead"):
        optimizer_client._OptimizerClient(svc, "p", "r", None)


def test_suggestions_are_sticky_per_client_and_lro_is_polled(svc):
    c = _client(svc)
    s1 = c.get_suggestions("tuner0")
    assert len(s1["trials"]) == constants.SUGGESTION_COUNT_PER_REQUEST
    t = s1["trials"][0]
    assert t["state"] == "ACTIVE" and t["name"].endswith("/trials/1")
    names = {p["parameter"] for p in t["parameters"]}
    assert names == {"learning_rate", "units", "act"}
    lr = [p for p in t["parameters"] if p["parameter"] == "learning_rate"][0]["floatValue"]
    assert 1e-4 <= lr <= 1e-2
    assert c.get_suggestions("tuner0")["trials"][0]["name"] == t["name"]      # same pending trial
    assert c.get_suggestions("tuner1")["trials"][0]["name"].endswith("/trials/2")
    # raw service answers "not done" first: the client really polls operations.get
    raw = svc.suggest_trials(c._make_study_name(), {"client_id": "x", "suggestion_count": 1})
    assert raw["done"] is False and svc.get_operation(raw["name"])["done"] is True


def test_suggest_429_returns_empty(svc, monkeypatch):
    c = _client(svc)
    monkeypatch.setattr(constants, "MAX_TRIALS_PER_STUDY", 2)
    for i in range(2):
        tid = c.get_suggestions("t%d" % i)["trials"][0]["name"].split("/")[-1]
        c.report_intermediate_objective_value(1, 1.0, [{"metric": "val_acc", "value": 0.5}], tid)
        c.complete_trial(tid, False)
    assert c.get_suggestions("t9") == {}


def test_suggest_on_inactive_study_raises(svc):
    c = _client(svc)
    svc.set_study_state(c._make_study_name(), "INACTIVE")
    with pytest.raises(optimizer_client.SuggestionInactiveError):
        c.get_suggestions("tuner0")


def test_measurements_complete_and_list(svc):
    c = _client(svc)
    tid = c.get_suggestions("tuner0")["trials"][0]["name"].split("/")[-1]
    c.report_intermediate_objective_value(1, 12.3, [{"metric": "val_acc", "value": 0.7}], tid)
    c.report_intermediate_objective_value(2, 20.0, [{"metric": "val_acc", "value": 0.8}], tid)
    done = c.complete_trial(tid, trial_infeasible=False)
    assert done["state"] == "COMPLETED"
    assert done["finalMeasurement"] == {"stepCount": 2, "elapsedTime": {"seconds": 20},
                                        "metrics": [{"metric": "val_acc", "value": 0.8}]}
    t2 = c.get_suggestions("tuner0")["trials"][0]["name"].split("/")[-1]
    assert t2 != tid
    bad = c.complete_trial(t2, trial_infeasible=True, infeasibility_reason="INVALID")
    assert bad["trialInfeasible"] is True and "finalMeasurement" not in bad
    states = [t["state"] for t in c.list_trials()]
    assert states == ["COMPLETED", "COMPLETED"]


def test_early_stopping_stops_the_trial(svc):
    c = _client(svc)
    for i, final in enumerate([0.9, 0.8, 0.85]):
        tid = c.get_suggestions("t%d" % i)["trials"][0]["name"].split("/")[-1]
        c.report_intermediate_objective_value(1, 1.0, [{"metric": "val_acc", "value": final}], tid)
        c.complete_trial(tid, False)
    good = c.get_suggestions("good")["trials"][0]["name"].split("/")[-1]
    c.report_intermediate_objective_value(1, 1.0, [{"metric": "val_acc", "value": 0.95}], good)
    assert c.should_trial_stop(good) is False
    poor = c.get_suggestions("poor")["trials"][0]["name"].split("/")[-1]
    c.report_intermediate_objective_value(1, 1.0, [{"metric": "val_acc", "value": 0.1}], poor)
    assert c.should_trial_stop(poor) is True
    assert [t["state"] for t in c.list_trials() if t["name"].endswith("/" + poor)] == ["STOPPING"]


def test_delete_study_and_404(svc):
    c = _client(svc)
    c.delete_study()
    with pytest.raises(ValueError, match=r"DeleteStudy failed. Study not found"):
        c.delete_study()
    with pytest.raises(local_vizier.HttpError):
        c.list_trials()


def test_lro_timeout(svc, monkeypatch):
    c = _client(svc)
    monkeypatch.setenv("TFC_LRO_SCALE", "0")
    monkeypatch.setattr(svc, "get_operation", lambda *a, **k: {"name": "op", "done": False})
    with pytest.raises(RuntimeError, match=r"GetLongRunningOperations timeout"):
        c._obtain_long_running_operation({"name": "op"})


def test_grid_search_enumerates_then_429(svc):
    cfg = dict(CONFIG, algorithm="GRID_SEARCH", parameters=[
        {"parameter": "units", "type": "INTEGER", "integer_value_spec": {"min_value": 1, "max_value": 2}},
        {"parameter": "act", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["relu", "elu"]}}])
    c = optimizer_client.create_or_load_study("proj", "r", "grid", cfg, service_client=svc)
    seen = []
    for i in range(4):
        t = c.get_suggestions("w%d" % i)["trials"][0]
        seen.append(tuple(p.get("intValue", p.get("stringValue")) for p in t["parameters"]))
    assert seen == [("1", "relu"), ("1", "elu"), ("2", "relu"), ("2", "elu")]
    assert c.get_suggestions("w9") == {}


def test_user_agent_is_stamped(svc):
    c = _client(svc)
    import json
    import os
    path = os.path.join(svc.root, "proj", "us-central1", "study-a", "study.json")
    assert json.load(open(path))["client"].startswith("cloud-b200/")
    assert c.project_id == "proj"


# -- the file-backed service under concurrent processes -------------------
def _hammer(args):
    root, worker, n_trials = args
    import os
    import random
    import time
    os.environ["TFC_VIZIER_ROOT"] = root
    os.environ["TFC_STUDY_RETRY_SECS"] = "0"
    from cloud_b200.tuner import local_vizier as lv, optimizer_client as oc
    time.sleep(random.random() * 0.05)                 # start jitter, as in the reference's integration test
    c = oc.create_or_load_study("proj", "r", "stress", CONFIG, service_client=lv.LocalVizier())
    mine = []
    for i in range(n_trials):
        sug = c.get_suggestions("w{}".format(worker))
        if not sug.get("trials"):
            break
        t = sug["trials"][0]
        tid = t["name"].split("/")[-1]
        again = c.get_suggestions("w{}".format(worker))["trials"][0]["name"]     # sticky while ACTIVE
        assert again == t["name"]
        c.report_intermediate_objective_value(step=1, elapsed_secs=0.1, metric_list=[{"acc": 0.1 * i}],
                                              trial_id=tid)
        c.complete_trial(tid, trial_infeasible=False)
        mine.append(t["name"])
    return mine


def test_file_locked_store_survives_concurrent_processes(tmp_path):
    import multiprocessing as mp
    root = str(tmp_path / "vizier")
    workers, per = 6, 8
    ctx = mp.get_context("spawn")
    with ctx.Pool(workers) as pool:
        results = pool.map(_hammer, [(root, w, per) for w in range(workers)])
    names = [n for r in results for n in r]
    assert len(names) == workers * per
    assert len(set(names)) == len(names), "two processes were handed the same trial"
    svc = local_vizier.LocalVizier(root)
    trials = svc.list_trials("projects/proj/locations/r/studies/stress")["trials"]
    assert len(trials) == workers * per
    assert sorted(int(t["name"].split("/")[-1]) for t in trials) == list(range(1, workers * per + 1))
    assert all(t["state"] == "COMPLETED" and t.get("finalMeasurement") for t in trials)
    assert {t["clientId"] for t in trials} == {"w{}".format(w) for w in range(workers)}
