"""Port of tuner/tests/unit/utils_test.py - including the eight convert_study_config_* cases that
never run upstream (missing `test_` prefix, :137-222), with automatedStoppingConfig in the goldens."""
import pytest

from cloud_b200.tuner import utils
from cloud_b200.tuner.engine import hyperparameters as hp_module
from cloud_b200.tuner.engine import oracle as oracle_module

STOP = {"decayCurveStoppingConfig": {"useElapsedTime": True}}


def _cfg(params, metric="val_accuracy", goal="MAXIMIZE"):
    return {"algorithm": "ALGORITHM_UNSPECIFIED", "automatedStoppingConfig": STOP,
            "metrics": [{"metric": metric, "goal": goal}], "parameters": params}


def test_make_study_config_float_linear_and_log_and_discrete():
    hps = hp_module.HyperParameters()
    hps.Float("learning_rate", min_value=1e-4, max_value=1e-1, sampling="log")
    hps.Float("dropout", 0.0, 0.5, sampling="linear")
    hps.Float("beta", 0.1, 0.4, step=0.1)
    got = utils.make_study_config("val_accuracy", hps)
    assert got["parameters"][0] == {"parameter": "learning_rate", "type": "DOUBLE",
                                    "double_value_spec": {"min_value": 1e-4, "max_value": 1e-1},
                                    "scale_type": "UNIT_LOG_SCALE"}
    assert got["parameters"][1]["scale_type"] == "UNIT_LINEAR_SCALE"
    assert got["parameters"][2]["type"] == "DISCRETE"
    assert got["parameters"][2]["discrete_value_spec"]["values"] == pytest.approx([0.1, 0.2, 0.3, 0.4][:len(got["parameters"][2]["discrete_value_spec"]["values"])])
    assert got["algorithm"] == "ALGORITHM_UNSPECIFIED" and got["automatedStoppingConfig"] == STOP
    assert got["metrics"] == [{"metric": "val_accuracy", "goal": "MAXIMIZE"}]


def test_make_study_config_int_choice_boolean_fixed():
    hps = hp_module.HyperParameters()
    hps.Int("units", 2, 10)
    hps.Int("stride", 8, 40, step=8)
    hps.Choice("opt", ["sgd", "adam"])
    hps.Choice("width", [64, 128, 256])
    hps.Boolean("has_beta")
    hps.Fixed("beta", 0.1)
    hps.Fixed("type", "WIDE_AND_DEEP")
    hps.Fixed("condition", True)
    got = utils.make_study_config(oracle_module.Objective("loss", "min"), hps)
    assert got == _cfg([
        {"parameter": "units", "type": "INTEGER", "integer_value_spec": {"min_value": 2, "max_value": 10}},
        {"parameter": "stride", "type": "DISCRETE", "discrete_value_spec": {"values": [8, 16, 24, 32]}},
        {"parameter": "opt", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["sgd", "adam"]}},
        {"parameter": "width", "type": "DISCRETE", "discrete_value_spec": {"values": [64, 128, 256]}},
        {"parameter": "has_beta", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["True", "False"]}},
        {"parameter": "beta", "type": "DISCRETE", "discrete_value_spec": {"values": [0.1]}},
        {"parameter": "type", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["WIDE_AND_DEEP"]}},
        {"parameter": "condition", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["True"]}},
    ], metric="loss", goal="MINIMIZE")


def test_convert_study_config_round_trip():
    cfg = _cfg([
        {"parameter": "learning_rate", "type": "DOUBLE", "double_value_spec": {"min_value": 1e-4, "max_value": 1e-2},
         "scale_type": "UNIT_LOG_SCALE"},
        {"parameter": "num_layers", "type": "INTEGER", "integer_value_spec": {"min_value": 2, "max_value": 10}},
        {"parameter": "act", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["relu", "elu"]}},
        {"parameter": "width", "type": "DISCRETE", "discrete_value_spec": {"values": [64, 128]}},
    ])
    hps = utils.convert_study_config_to_hps(cfg)
    space = {p.name: p for p in hps.space}
    assert isinstance(space["learning_rate"], hp_module.Float) and space["learning_rate"].sampling == "log"
    assert isinstance(space["num_layers"], hp_module.Int) and space["num_layers"].max_value == 10
    assert space["act"].values == ["relu", "elu"] and space["width"].values == [64, 128]
    obj = utils.convert_study_config_to_objective(cfg)
    assert obj == [oracle_module.Objective("val_accuracy", "max")]
    again = utils.make_study_config(obj[0], hps)
    assert again == cfg


@pytest.mark.parametrize("param,regex", [
    ({"type": "DOUBLE"}, r'"parameter" \(name\) is not specified'),
    ({"parameter": "x"}, r"type is not specified"),
    ({"parameter": "x", "type": "DISCRETE"}, r"missing discrete_value_spec"),
    ({"parameter": "x", "type": "DISCRETE", "discrete_value_spec": {"v": 1}}, r'missing "values"'),
    ({"parameter": "x", "type": "CATEGORICAL"}, r"missing categorical_value_spec"),
    ({"parameter": "x", "type": "DOUBLE"}, r"missing double_value_spec"),
    ({"parameter": "x", "type": "DOUBLE", "double_value_spec": {"min_value": 1, "max_value": 2}},
     r'requires both "min_value" and "max_value"'),
    ({"parameter": "x", "type": "INTEGER", "integer_value_spec": {"min_value": 1.0, "max_value": 2}},
     r'requires both "min_value" and "max_value"'),
    ({"parameter": "x", "type": "COMPLEX"}, r"Unknown parameter type"),
])
def test_invalid_study_config_parameters(param, regex):
    with pytest.raises(ValueError, match=regex):
        utils.convert_study_config_to_hps(_cfg([param]))


def test_study_config_without_metrics_or_parameters():
    with pytest.raises(ValueError, match=r'"metrics" not found'):
        utils.convert_study_config_to_objective({"parameters": []})
    with pytest.raises(ValueError):
        utils.convert_study_config_to_hps({"metrics": [{"metric": "a", "goal": "MAXIMIZE"}]})


def test_convert_optimizer_trial_to_hps():
    hps = hp_module.HyperParameters()
    hps.Choice("learning_rate", [1e-4, 1e-3, 1e-2])
    hps.Int("units", 1, 8)
    hps.Choice("act", ["relu", "elu"])
    trial = {"name": "projects/p/locations/r/studies/s/trials/7",
             "parameters": [{"parameter": "learning_rate", "floatValue": 0.0001},
                            {"parameter": "units", "intValue": "4"},
                            {"parameter": "act", "stringValue": "elu"}]}
    got = utils.convert_optimizer_trial_to_hps(hps, trial)
    assert got.values == {"learning_rate": 0.0001, "units": 4, "act": "elu"}
    assert utils.get_trial_id(trial) == "7"


def test_format_objective_and_goal():
    assert utils.format_objective("accuracy") == [oracle_module.Objective("accuracy", "max")]
    assert utils.format_objective("val_loss") == [oracle_module.Objective("val_loss", "min")]
    assert utils.format_objective(oracle_module.Objective("val_acc", "max")) == \
        [oracle_module.Objective("val_acc", "max")]
    assert utils.format_objective(["accuracy", "loss"]) == \
        [oracle_module.Objective("accuracy", "max"), oracle_module.Objective("loss", "min")]
    assert utils.format_objective("custom", direction="min") == [oracle_module.Objective("custom", "min")]
    with pytest.raises(TypeError):
        utils.format_objective(3)
    assert utils.format_goal("max") == "MAXIMIZE" and utils.format_goal("MINIMIZE") == "min"
    assert utils.format_goal("sideways") == "GOAL_TYPE_UNSPECIFIED"
