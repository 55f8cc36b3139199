"""Suggestion and early-stopping algorithms of the local optimizer service.

The reference leaves both to Google's Vizier service and only names them in its wire contract
(tuner/api/ml_public_google_rest_v1.json: ``StudyConfig.algorithm``, ``automatedStoppingConfig``;
``make_study_config`` asks for ALGORITHM_UNSPECIFIED + decay-curve stopping, tuner/utils.py:58-68),
so these tests have no upstream counterpart: they check that the local stand-ins behave like the
algorithms they are named after."""
import json
import math
import os
import statistics

import numpy as np
import pytest

from cloud_b200.tuner import gp_bandit, local_vizier

PARENT = "projects/p/locations/r"
SPACE = [
    {"parameter": "lr", "type": "DOUBLE", "double_value_spec": {"min_value": 1e-5, "max_value": 1e-1},
     "scale_type": "UNIT_LOG_SCALE"},
    {"parameter": "units", "type": "INTEGER", "integer_value_spec": {"min_value": 8, "max_value": 128}},
    {"parameter": "act", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["relu", "tanh", "elu"]}},
    {"parameter": "depth", "type": "DISCRETE", "discrete_value_spec": {"values": [1, 2, 4, 8]}},
]


def _values(parameters):
    return {p["parameter"]: gp_bandit.param_value(p) for p in parameters}


def _objective(parameters):
    v = _values(parameters)
    return -(math.log10(v["lr"]) + 3) ** 2 - ((v["units"] - 96) / 40) ** 2 \
        - {"relu": 0.0, "tanh": 0.5, "elu": 0.2}[v["act"]] - 0.05 * abs(math.log2(v["depth"]) - 1)


@pytest.fixture
def svc(tmp_path):
    return local_vizier.LocalVizier(str(tmp_path / "vizier"))


def _study(svc, study_id, algorithm=None, seed=0, stopping=None, goal="MAXIMIZE"):
    cfg = {"metrics": [{"metric": "score", "goal": goal}], "parameters": SPACE}
    if algorithm:
        cfg["algorithm"] = algorithm
    if stopping:
        cfg["automatedStoppingConfig"] = stopping
    svc.create_study(PARENT, study_id, {"study_config": cfg})
    name = "{}/studies/{}".format(PARENT, study_id)
    path = os.path.join(svc._study_dir(name), "study.json")
    study = json.load(open(path))
    study["seed"] = seed                # the service draws one at random; pin it for the test
    json.dump(study, open(path, "w"))
    return name


def _suggest(svc, name, client="c"):
    op = svc.suggest_trials(name, {"client_id": client, "suggestion_count": 1})
    return svc.get_operation(op["name"])["response"]["trials"][0]


def _finish(svc, trial, curve):
    for step, value in enumerate(curve, 1):
        svc.add_measurement(trial["name"], {"measurement": {
            "stepCount": step, "elapsedTime": {"seconds": 10 * step},
            "metrics": [{"metric": "score", "value": value}]}})
    svc.complete_trial(trial["name"], {})


def _search(svc, name, n):
    best = -float("inf")
    for _ in range(n):
        t = _suggest(svc, name)
        value = _objective(t["parameters"])
        best = max(best, value)
        _finish(svc, t, [value])
    return best


def test_encoder_maps_every_parameter_kind_into_the_unit_cube():
    enc = gp_bandit.SpaceEncoder(SPACE)
    assert enc.dims == 1 + 1 + 3 + 1
    x = enc.encode([{"parameter": "lr", "floatValue": 1e-3}, {"parameter": "units", "intValue": "68"},
                    {"parameter": "act", "stringValue": "tanh"}, {"parameter": "depth", "floatValue": 8.0}])
    np.testing.assert_allclose(x, [0.5, 0.5, 0, 1, 0, 1.0], atol=1e-9)
    rev = gp_bandit.SpaceEncoder([dict(SPACE[0], scale_type="UNIT_REVERSE_LOG_SCALE")])
    lo, hi = rev.encode([{"parameter": "lr", "floatValue": 1e-5}]), rev.encode([{"parameter": "lr", "floatValue": 1e-1}])
    assert lo[0] == pytest.approx(0.0) and hi[0] == pytest.approx(1.0)


def test_gaussian_process_interpolates_and_is_uncertain_far_away():
    rng = np.random.default_rng(0)
    x = rng.random((24, 2))
    y = np.sin(4 * x[:, 0]) + x[:, 1] ** 2
    y = (y - y.mean()) / y.std()
    gp = gp_bandit.GaussianProcess(x, y)
    mean, std = gp.predict(x)
    np.testing.assert_allclose(mean, y, atol=0.05)
    assert std.max() < 0.2
    _m, far = gp.predict(np.array([[5.0, 5.0]]))
    assert far[0] > 0.9                                  # back to the prior away from the data
    ei = gp_bandit.expected_improvement(np.array([0.0, 1.0]), np.array([1.0, 1.0]), incumbent=0.5)
    assert ei[1] > ei[0] > 0


def test_gp_bandit_beats_random_search_on_a_smooth_objective(svc):
    random_best = [_search(svc, _study(svc, "rnd%d" % s, "RANDOM_SEARCH", seed=s), 24) for s in range(3)]
    gp_best = [_search(svc, _study(svc, "gp%d" % s, "GAUSSIAN_PROCESS_BANDIT", seed=s), 24) for s in range(3)]
    assert statistics.mean(gp_best) > statistics.mean(random_best)
    assert min(gp_best) > -0.2                           # close to the optimum (0) every time


def test_unspecified_algorithm_is_the_service_default(svc, monkeypatch):
    calls = []
    real = gp_bandit.choose
    monkeypatch.setattr(gp_bandit, "choose", lambda *a, **k: calls.append(1) or real(*a, **k))
    _search(svc, _study(svc, "default"), 10)
    assert calls, "ALGORITHM_UNSPECIFIED never reached the GP bandit"
    calls.clear()
    monkeypatch.setenv("TFC_VIZIER_DEFAULT_ALGORITHM", "RANDOM_SEARCH")
    _search(svc, _study(svc, "forced-random"), 10)
    assert not calls


def test_running_trials_spread_parallel_suggestions(svc):
    name = _study(svc, "parallel", "GAUSSIAN_PROCESS_BANDIT", seed=3)
    _search(svc, name, 12)
    batch = [_suggest(svc, name, client="w%d" % i) for i in range(8)]       # eight workers at once
    points = {json.dumps(t["parameters"], sort_keys=True) for t in batch}
    assert len(points) == 8
    for t in batch:
        v = _values(t["parameters"])
        assert 1e-5 <= v["lr"] <= 1e-1 and 8 <= v["units"] <= 128
        assert v["act"] in ("relu", "tanh", "elu") and v["depth"] in (1, 2, 4, 8)


def test_infeasible_trials_are_avoided_not_fatal(svc):
    name = _study(svc, "infeasible", "GAUSSIAN_PROCESS_BANDIT", seed=1)
    for i in range(14):
        t = _suggest(svc, name)
        if _values(t["parameters"])["act"] == "tanh":
            svc.complete_trial(t["name"], {"trial_infeasible": True, "infeasible_reason": "diverged"})
        else:
            _finish(svc, t, [_objective(t["parameters"])])
    later = [_values(_suggest(svc, name, client="z%d" % i)["parameters"])["act"] for i in range(6)]
    assert later.count("tanh") <= 2


def test_trial_ids_are_not_reused_after_delete(svc):
    name = _study(svc, "ids", "RANDOM_SEARCH")
    a, b = _suggest(svc, name, "a"), _suggest(svc, name, "b")
    svc.delete_trial(b["name"])
    c = _suggest(svc, name, "c")
    assert [t["name"].rpartition("/")[2] for t in (a, b, c)] == ["1", "2", "3"]


DECAY = {"decayCurveStoppingConfig": {"useElapsedTime": False}}
MEDIAN = {"medianAutomatedStoppingConfig": {"useElapsedTime": False}}


def _running(svc, name, curve, client):
    t = _suggest(svc, name, client)
    for step, value in enumerate(curve, 1):
        svc.add_measurement(t["name"], {"measurement": {"stepCount": step,
                                                        "metrics": [{"metric": "score", "value": value}]}})
    op = svc.check_early_stopping_state(t["name"])
    return svc.get_operation(op["name"])["response"]["shouldStop"]


def test_decay_curve_rule_extrapolates_before_it_stops(svc):
    name = _study(svc, "decay", "RANDOM_SEARCH", stopping=DECAY)
    for k in range(4):                  # finished trials climb 0.2 -> ~0.9 over ten steps
        _finish(svc, _suggest(svc, name, "f%d" % k),
                [0.9 - 0.7 * math.exp(-0.5 * s) - 0.01 * k for s in range(10)])
    # early and behind: the finished trials still gained a lot after step 2, so keep going
    assert _running(svc, name, [0.15, 0.35], "early") is False
    # late and far behind: even the optimistic gain cannot reach the best finished trial
    assert _running(svc, name, [0.1, 0.2, 0.25, 0.28, 0.3, 0.31, 0.31, 0.32], "late") is True
    # late and ahead of everything
    assert _running(svc, name, [0.3, 0.6, 0.8, 0.9, 0.93, 0.95, 0.95, 0.96], "ahead") is False


def test_median_rule_compares_at_equal_progress(svc):
    name = _study(svc, "median", "RANDOM_SEARCH", stopping=MEDIAN)
    for k, top in enumerate([0.9, 0.8, 0.7]):
        _finish(svc, _suggest(svc, name, "f%d" % k), [top * (1 - math.exp(-0.7 * s)) for s in range(1, 9)])
    assert _running(svc, name, [0.05, 0.1], "slow") is True          # below the median at step 2
    assert _running(svc, name, [0.5, 0.75], "fast") is False
    # the decay-curve rule would not have stopped the slow one that early
    other = _study(svc, "decay2", "RANDOM_SEARCH", stopping=DECAY)
    for k, top in enumerate([0.9, 0.8, 0.7]):
        _finish(svc, _suggest(svc, other, "f%d" % k), [top * (1 - math.exp(-0.7 * s)) for s in range(1, 9)])
    assert _running(svc, other, [0.25, 0.5], "slow") is False


def test_stopping_rules_respect_a_minimising_goal_and_elapsed_time(svc):
    name = _study(svc, "min", "RANDOM_SEARCH", goal="MINIMIZE",
                  stopping={"decayCurveStoppingConfig": {"useElapsedTime": True}})
    for k in range(3):
        _finish(svc, _suggest(svc, name, "f%d" % k), [2.0 * math.exp(-0.4 * s) + 0.1 * k for s in range(8)])
    t = _suggest(svc, name, "bad")
    for step, loss in enumerate([3.0, 2.9, 2.9, 2.8, 2.8, 2.8, 2.8], 1):
        svc.add_measurement(t["name"], {"measurement": {"stepCount": step, "elapsedTime": {"seconds": 10 * step},
                                                        "metrics": [{"metric": "score", "value": loss}]}})
    op = svc.check_early_stopping_state(t["name"])
    assert svc.get_operation(op["name"])["response"]["shouldStop"] is True
    fresh = _suggest(svc, name, "no-measurements")
    op = svc.check_early_stopping_state(fresh["name"])
    assert svc.get_operation(op["name"])["response"]["shouldStop"] is False


def test_no_stopping_config_never_stops(svc):
    name = _study(svc, "plain", "RANDOM_SEARCH")
    for k in range(4):
        _finish(svc, _suggest(svc, name, "f%d" % k), [0.9])
    assert _running(svc, name, [0.0, 0.0, 0.0], "hopeless") is False


# -- the same Gaussian-process model behind KerasTuner's BayesianOptimization tuner --------------------
def _drive(oracle, n):
    """create_trial / update_trial / end_trial like Tuner.search does, scoring with a known function."""
    best = -float("inf")
    for _ in range(n):
        trial = oracle.create_trial("tuner0")
        v = trial.hyperparameters.values
        score = -(math.log10(v["lr"]) + 3) ** 2 - ((v["units"] - 96) / 40) ** 2 - (0.0 if v["act"] == "relu" else 0.4) \
            - (0.1 if v["wide"] else 0.0)
        oracle.update_trial(trial.trial_id, {"score": score}, step=0)
        oracle.end_trial(trial.trial_id, "COMPLETED")
        best = max(best, score)
    return best


def _hp_space():
    from cloud_b200.tuner import HyperParameters
    hps = HyperParameters()
    hps.Float("lr", 1e-5, 1e-1, sampling="log", default=1e-2)
    hps.Int("units", 8, 128, step=8, default=32)
    hps.Choice("act", ["relu", "tanh"])
    hps.Boolean("wide", default=True)
    return hps


def test_bayesian_optimization_oracle_beats_random_search(tmp_path):
    from cloud_b200.tuner import Objective
    from cloud_b200.tuner.engine import oracle as oracle_module
    gp, rnd = [], []
    for seed in range(3):
        for cls, out in ((oracle_module.BayesianOptimizationOracle, gp), (oracle_module.RandomSearchOracle, rnd)):
            o = cls(Objective("score", "max"), max_trials=30, hyperparameters=_hp_space(), seed=seed + 1)
            o._set_project_dir(str(tmp_path / "{}{}".format(cls.__name__, seed)), "p", overwrite=True)
            out.append(_drive(o, 26))
    assert statistics.mean(gp) > statistics.mean(rnd) and statistics.median(gp) > -0.15
    # first trial = the declared defaults; values respect steps and choices
    o = oracle_module.BayesianOptimizationOracle(Objective("score", "max"), max_trials=4,
                                                 hyperparameters=_hp_space(), seed=7)
    o._set_project_dir(str(tmp_path / "first"), "p", overwrite=True)
    first = o.create_trial("t").hyperparameters.values
    assert first == {"lr": 1e-2, "units": 32, "act": "relu", "wide": True}


def test_bayesian_optimization_tuner_searches_a_real_model(tmp_path):
    import cloud_b200 as tfc
    from cloud_b200 import tuner as tuners
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    keras = tfc.keras
    rng = np.random.RandomState(0)
    centers = rng.randn(3, 10) * 3
    y = rng.randint(0, 3, 192)
    x = (centers[y] + rng.randn(192, 10)).astype("float32")

    def build(hp):
        m = keras.Sequential([keras.layers.Dense(hp.Int("units", 4, 32, step=4), activation="relu", input_shape=(10,)),
                              keras.layers.Dense(3, activation="softmax")])
        m.compile(keras.optimizers.Adam(hp.Float("lr", 1e-4, 1e-1, sampling="log")),
                  "sparse_categorical_crossentropy", metrics=["accuracy"])
        return m
    t = tuners.BayesianOptimization(build, objective="val_accuracy", max_trials=5, num_initial_points=2, seed=0,
                                    directory=str(tmp_path), project_name="bo")
    t.search(x=x, y=y, epochs=2, validation_data=(x, y), verbose=0)
    assert len(t.oracle.trials) == 5
    best = t.get_best_models(1)[0]
    assert best.evaluate(x, y, verbose=0)[1] >= max(tr.score for tr in t.oracle.trials.values()) - 0.2


# -- Hyperband ------------------------------------------------------------------------------------------
def test_hyperband_schedule_promotes_the_best_and_waits_for_unfinished_rounds(tmp_path):
    from cloud_b200.tuner import Objective
    from cloud_b200.tuner.engine import hyperband, trial as trial_lib
    o = hyperband.HyperbandOracle(Objective("score", "max"), max_epochs=9, factor=3, hyperparameters=_hp_space(), seed=3)
    o._set_project_dir(str(tmp_path), "hb", overwrite=True)
    assert o.num_brackets() == 3
    assert [[o.round_size(b, r) for r in range(o.num_rounds(b))] for b in (2, 1, 0)] == [[9, 3, 1], [5, 2], [3]]
    assert [[o.round_epochs(b, r) for r in range(o.num_rounds(b))] for b in (2, 1, 0)] == [[1, 3, 9], [3, 9], [9]]
    assert o.total_trials() == 23

    def finish(trial, score, status="COMPLETED"):
        if status == "COMPLETED":
            o.update_trial(trial.trial_id, {"score": score}, step=trial.hyperparameters.values["tuner/epochs"] - 1)
        o.end_trial(trial.trial_id, status)

    # bracket 2, round 0: nine fresh configurations on one epoch; two workers
    first = [o.create_trial("w%d" % i) for i in range(2)]
    assert first[0].hyperparameters.values["units"] == 32            # the defaults go first
    finish(first[0], 0.10)
    finish(first[1], 0.20)
    rest = []
    for i in range(7):
        t = o.create_trial("w0")
        assert t.hyperparameters.values["tuner/epochs"] == 1 and t.hyperparameters.values["tuner/round"] == 0
        rest.append(t)
        if i < 6:
            finish(t, 0.30 + 0.01 * i)
    # the round is handed out but its last trial still runs: a second worker has to wait
    assert o.create_trial("w1").status == trial_lib.TrialStatus.IDLE
    finish(rest[6], None, status="INVALID")                         # ... and then fails
    promoted = [o.create_trial("w%d" % i) for i in range(3)]
    parents = [t.hyperparameters.values["tuner/trial_id"] for t in promoted]
    assert parents == [rest[5].trial_id, rest[4].trial_id, rest[3].trial_id]      # best three, best first
    for t, parent in zip(promoted, rest[5:2:-1]):
        v = t.hyperparameters.values
        assert (v["tuner/epochs"], v["tuner/initial_epoch"], v["tuner/round"], v["tuner/bracket"]) == (3, 1, 1, 2)
        assert {k: v[k] for k in ("lr", "units", "act", "wide")} == \
            {k: parent.hyperparameters.values[k] for k in ("lr", "units", "act", "wide")}
    for i, t in enumerate(promoted):
        finish(t, 0.5 + 0.1 * i)
    final = o.create_trial("w0")
    assert final.hyperparameters.values["tuner/trial_id"] == promoted[2].trial_id
    assert final.hyperparameters.values["tuner/epochs"] == 9 and final.hyperparameters.values["tuner/initial_epoch"] == 3
    finish(final, 0.9)
    # the remaining brackets run to the end, then the oracle stops
    count = 13
    while True:
        t = o.create_trial("w0")
        if t.status == trial_lib.TrialStatus.STOPPED:
            break
        finish(t, 0.1)
        count += 1
    assert count == o.total_trials()
    assert o.get_best_trials(1)[0].trial_id == final.trial_id


def test_hyperband_tuner_continues_training_from_the_parent_checkpoint(tmp_path):
    import cloud_b200 as tfc
    from cloud_b200 import tuner as tuners
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    keras = tfc.keras
    rng = np.random.RandomState(0)
    centers = rng.randn(3, 10) * 2
    y = rng.randint(0, 3, 192)
    x = (centers[y] + rng.randn(192, 10)).astype("float32")
    histories, built = {}, []

    class Record(keras.callbacks.Callback):
        def on_epoch_end(self, epoch, logs=None):
            histories.setdefault(self.model._build_index, []).append((epoch, logs["loss"]))

    def build(hp):
        m = keras.Sequential([keras.layers.Dense(hp.Int("units", 8, 32, step=8), activation="relu", input_shape=(10,)),
                              keras.layers.Dense(3, activation="softmax")])
        m.compile(keras.optimizers.SGD(hp.Choice("lr", [0.01, 0.03])), "sparse_categorical_crossentropy",
                  metrics=["accuracy"])
        m._build_index = len(built)                 # object ids can be reused once a model is freed
        built.append(m._build_index)
        return m
    t = tuners.Hyperband(build, objective="loss", max_epochs=4, factor=3, seed=1, directory=str(tmp_path),
                         project_name="hb")
    assert t.oracle.total_trials() == 10
    t.search(x=x, y=y, batch_size=32, verbose=0, callbacks=[Record()])
    trials = list(t.oracle.trials.values())
    assert len(trials) == 10
    promoted = [tr for tr in trials if "tuner/trial_id" in tr.hyperparameters.values]
    assert len(promoted) == 2 and all(tr.hyperparameters.values["tuner/initial_epoch"] == 2 for tr in promoted)
    runs = sorted(histories.values(), key=lambda h: h[0][0])
    fresh = [h for h in runs if h[0][0] == 0]
    resumed = [h for h in runs if h[0][0] == 2]
    assert len(fresh) == 8 and len(resumed) == 2
    assert all([e for e, _l in h] == [2, 3] for h in resumed)           # two more epochs, not four
    # a resumed run starts where its parent stopped: below the parent's own first epoch and close to
    # its last one (models are built once for the search space, then once per trial, in trial order)
    by_trial = {tr.trial_id: histories[built[i + 1]] for i, tr in enumerate(trials)}
    for tr in promoted:
        parent = by_trial[tr.hyperparameters.values["tuner/trial_id"]]
        mine = by_trial[tr.trial_id]
        assert mine[0][1] < parent[0][1] and mine[0][1] < parent[-1][1] + 0.05
    best = t.get_best_models(1)[0]
    assert best.evaluate(x, y, verbose=0)[0] < t.oracle.get_best_trials(1)[0].score + 0.05
