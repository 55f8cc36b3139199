"""Search space <-> StudyConfig codec
(reference: src/python/tensorflow_cloud/tuner/utils.py:47-399).

The dictionaries produced here are the wire contract of the optimizer service (exact dict equality
is tested upstream: tuner/tests/unit/tuner_test.py:60-73), so key names - including the snake_case
``*_value_spec`` / ``scale_type`` spellings - are kept as they are.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Text

from .engine import hyperparameters as hp_module
from .engine import metrics_tracking
from .engine import oracle as oracle_module

# CAIP Optimizer constants
_DISCRETE = "DISCRETE"
_CATEGORICAL = "CATEGORICAL"
_DOUBLE = "DOUBLE"
_INTEGER = "INTEGER"
_LINEAR_SCALE = "UNIT_LINEAR_SCALE"
_LOG_SCALE = "UNIT_LOG_SCALE"
_REVERSE_LOG_SCALE = "UNIT_REVERSE_LOG_SCALE"
_SCALE_TYPE_UNSPECIFIED = "SCALE_TYPE_UNSPECIFIED"
_SAMPLING_LINEAR = "linear"
_SAMPLING_LOG = "log"
_SAMPLING_REVERSE_LOG = "reverse_log"
_GOAL_MAXIMIZE = "MAXIMIZE"
_GOAL_MINIMIZE = "MINIMIZE"
_DIRECTION_MAX = "max"
_DIRECTION_MIN = "min"


def make_study_config(objective, hyperparams) -> Dict[Text, Any]:
    """StudyConfig from an objective and a search space (utils.py:47-81)."""
    study_config = {}
    study_config["algorithm"] = "ALGORITHM_UNSPECIFIED"   # the service default
    study_config["automatedStoppingConfig"] = {"decayCurveStoppingConfig": {"useElapsedTime": True}}
    study_config["metrics"] = []
    for obj in format_objective(objective):
        study_config["metrics"].append({"metric": obj.name, "goal": format_goal(obj.direction)})
    study_config["parameters"] = _convert_hyperparams_to_optimizer_params(hyperparams)
    return study_config


def convert_study_config_to_objective(study_config):
    """StudyConfig metrics -> list of Objective (utils.py:84-98)."""
    if not study_config.get("metrics"):
        raise ValueError('"metrics" not found in study_config {}'.format(study_config))
    if not isinstance(study_config["metrics"], list):
        raise ValueError('study_config["metrics"] should be a list of {"metric": ..., "goal": ...}')
    if len(study_config["metrics"]) == 1:
        m = study_config["metrics"][0]
        return [oracle_module.Objective(name=m["metric"], direction=format_goal(m["goal"]))]
    return [oracle_module.Objective(name=m["metric"], direction=format_goal(m["goal"]))
            for m in study_config["metrics"]]


def convert_study_config_to_hps(study_config) -> hp_module.HyperParameters:
    """StudyConfig parameters -> HyperParameters (utils.py:101-158)."""
    if not study_config.get("parameters"):
        raise ValueError("Parameters are not found in the study_config: ", study_config)
    if not isinstance(study_config["parameters"], list):
        raise ValueError("Parameters should be a list of parameter with at least 1 parameter, "
                         "found ", study_config["parameters"])
    hps = hp_module.HyperParameters()
    for param in study_config["parameters"]:
        _is_parameter_valid(param)
        name = param["parameter"]
        if param["type"] == _DISCRETE:
            hps.Choice(name, param["discrete_value_spec"]["values"])
        elif param["type"] == _CATEGORICAL:
            hps.Choice(name, param["categorical_value_spec"]["values"])
        elif param["type"] == _DOUBLE:
            spec = param["double_value_spec"]
            hps.Float(name, min_value=spec["min_value"], max_value=spec["max_value"],
                      sampling=_format_sampling(param.get("scale_type")))
        elif param["type"] == _INTEGER:
            spec = param["integer_value_spec"]
            hps.Int(name, min_value=spec["min_value"], max_value=spec["max_value"],
                    sampling=_format_sampling(param.get("scale_type")))
    return hps


def _is_parameter_valid(param):
    if not param.get("parameter"):
        raise ValueError('"parameter" (name) is not specified.')
    if not param.get("type"):
        raise ValueError("Parameter {} type is not specified.".format(param))
    kind = param["type"]
    if kind == _DISCRETE:
        if not param.get("discrete_value_spec"):
            raise ValueError("Parameter {} is missing discrete_value_spec.".format(param))
        if not isinstance(param["discrete_value_spec"].get("values"), list):
            raise ValueError('Parameter spec {} is missing "values".'.format(param["discrete_value_spec"]))
    elif kind == _CATEGORICAL:
        if not param.get("categorical_value_spec"):
            raise ValueError("Parameter {} is missing categorical_value_spec.".format(param))
        if not isinstance(param["categorical_value_spec"].get("values"), list):
            raise ValueError('Parameter spec {} is missing "values".'.format(param["categorical_value_spec"]))
    elif kind == _DOUBLE:
        if not param.get("double_value_spec"):
            raise ValueError("Parameter {} is missing double_value_spec.".format(param))
        spec = param["double_value_spec"]
        if not (isinstance(spec.get("min_value"), float) and isinstance(spec.get("max_value"), float)):
            raise ValueError('Parameter spec {} requires both "min_value" and "max_value".'.format(spec))
    elif kind == _INTEGER:
        if not param.get("integer_value_spec"):
            raise ValueError("Parameter {} is missing integer_value_spec.".format(param))
        spec = param["integer_value_spec"]
        if not (isinstance(spec.get("min_value"), int) and isinstance(spec.get("max_value"), int)):
            raise ValueError('Parameter spec {} requires both "min_value" and "max_value".'.format(spec))
    else:
        raise ValueError("Unknown parameter type: {}.".format(kind))


def _convert_hyperparams_to_optimizer_params(hyperparams) -> List[Any]:
    """HyperParameters -> list of ParameterSpec (utils.py:220-282)."""
    out = []
    for hp in hyperparams.space:
        param = {"parameter": hp.name}
        if isinstance(hp, hp_module.Choice):
            values = hp.values
            if isinstance(values[0], str):
                param["type"] = _CATEGORICAL
                param["categorical_value_spec"] = {"values": values}
            else:
                param["type"] = _DISCRETE
                param["discrete_value_spec"] = {"values": values}
        elif isinstance(hp, hp_module.Int):
            if hp.step is not None and hp.step != 1:
                param["type"] = _DISCRETE
                param["discrete_value_spec"] = {"values": list(range(hp.min_value, hp.max_value, hp.step))}
            else:
                param["type"] = _INTEGER
                param["integer_value_spec"] = {"min_value": hp.min_value, "max_value": hp.max_value}
                if hp.sampling is not None:
                    param.update(_get_scale_type(hp.sampling))
        elif isinstance(hp, hp_module.Float):
            if hp.step is not None:
                values, val = [], hp.min_value
                while val < hp.max_value:
                    values.append(val)
                    val += hp.step
                param["type"] = _DISCRETE
                param["discrete_value_spec"] = {"values": values}
            else:
                param["type"] = _DOUBLE
                param["double_value_spec"] = {"min_value": hp.min_value, "max_value": hp.max_value}
                if hp.sampling is not None:
                    param.update(_get_scale_type(hp.sampling))
        elif isinstance(hp, hp_module.Boolean):
            param["type"] = _CATEGORICAL
            param["categorical_value_spec"] = {"values": ["True", "False"]}
        elif isinstance(hp, hp_module.Fixed):
            if isinstance(hp.value, (str, bool)):
                param["type"] = _CATEGORICAL
                param["categorical_value_spec"] = {"values": [str(hp.value)]}
            else:
                param["type"] = _DISCRETE
                param["discrete_value_spec"] = {"values": [float(hp.value)]}
        else:
            raise ValueError("`HyperParameter` type not recognized: {}".format(hp))
        out.append(param)
    return out


def format_objective(objective, direction: Text = None) -> Optional[List[oracle_module.Objective]]:
    """str / Objective / list of either -> list of Objective (utils.py:285-323)."""
    if isinstance(objective, oracle_module.Objective):
        return [objective]
    if isinstance(objective, str):
        if direction:
            return [oracle_module.Objective(name=objective, direction=direction)]
        return [oracle_module.Objective(
            name=objective, direction=metrics_tracking.infer_metric_direction(objective))]
    if isinstance(objective, list):
        if all(isinstance(o, oracle_module.Objective) for o in objective):
            return objective
        if all(isinstance(o, str) for o in objective):
            return [oracle_module.Objective(name=o, direction=metrics_tracking.infer_metric_direction(o))
                    for o in objective]
    raise TypeError("Objective should be either string or oracle_module.Objective, found {}"
                    .format(objective))


def format_goal(metric_direction: Text) -> Text:
    """'max' <-> 'MAXIMIZE', 'min' <-> 'MINIMIZE' (utils.py:326-346)."""
    table = {_DIRECTION_MAX: _GOAL_MAXIMIZE, _DIRECTION_MIN: _GOAL_MINIMIZE,
             _GOAL_MAXIMIZE: _DIRECTION_MAX, _GOAL_MINIMIZE: _DIRECTION_MIN}
    if metric_direction not in table:
        return "GOAL_TYPE_UNSPECIFIED"
    return table[metric_direction]


def _get_scale_type(sampling):
    if sampling == _SAMPLING_LINEAR:
        return {"scale_type": _LINEAR_SCALE}
    if sampling == _SAMPLING_LOG:
        return {"scale_type": _LOG_SCALE}
    if sampling == _SAMPLING_REVERSE_LOG:
        return {"scale_type": _REVERSE_LOG_SCALE}
    return {"scale_type": _SCALE_TYPE_UNSPECIFIED}


def get_trial_id(optimizer_trial: Dict[Text, Any]) -> Text:
    """Last path segment of ``projects/.../studies/<study>/trials/<id>`` (utils.py:360-371)."""
    return optimizer_trial["name"].split("/")[-1]


def convert_optimizer_trial_to_hps(hps, optimizer_trial) -> hp_module.HyperParameters:
    """Optimizer trial parameters -> HyperParameters values (utils.py:374-388)."""
    hps = hp_module.HyperParameters.from_config(hps.get_config())
    hps.values = {}
    for param in optimizer_trial["parameters"]:
        if "floatValue" in param:
            hps.values[param["parameter"]] = float(param["floatValue"])
        if "intValue" in param:
            hps.values[param["parameter"]] = int(param["intValue"])
        if "stringValue" in param:
            hps.values[param["parameter"]] = str(param["stringValue"])
    return hps


def _format_sampling(scale_type: Text) -> Optional[Text]:
    if scale_type == _LINEAR_SCALE:
        return _SAMPLING_LINEAR
    if scale_type == _LOG_SCALE:
        return _SAMPLING_LOG
    if scale_type == _REVERSE_LOG_SCALE:
        return _SAMPLING_REVERSE_LOG
    return None
