"""Search space <-> StudyConfig codec
(reference behaviour: src/python/tensorflow_cloud/tuner/utils.py:47-399).

The dictionaries produced here are the wire contract of the optimizer service (exact dict equality
is tested upstream: tuner/tests/unit/tuner_test.py:60-73), so key names - including the snake_case
``*_value_spec`` / ``scale_type`` spellings - are kept as they are.

Organisation: the four parameter kinds of the service are rows of one table (``_KINDS``) that says
where a kind keeps its specification, how that specification is checked and how it becomes a
hyperparameter; encoding goes through one small function per hyperparameter class
(``_ENCODERS``).  Both directions of the sampling / goal vocabularies are derived from a single
mapping each.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, NamedTuple, Optional, Text

from .engine import hyperparameters as hp_module
from .engine import metrics_tracking
from .engine import oracle as oracle_module

# ----------------------------------------------------------------------------------------------
# vocabularies (KerasTuner word -> service word; the reverse maps are derived)
# ----------------------------------------------------------------------------------------------
_SCALE_OF_SAMPLING = {"linear": "UNIT_LINEAR_SCALE", "log": "UNIT_LOG_SCALE",
                      "reverse_log": "UNIT_REVERSE_LOG_SCALE"}
_SAMPLING_OF_SCALE = {v: k for k, v in _SCALE_OF_SAMPLING.items()}
_GOAL_OF_DIRECTION = {"max": "MAXIMIZE", "min": "MINIMIZE"}
_DIRECTION_OF_GOAL = {v: k for k, v in _GOAL_OF_DIRECTION.items()}


def format_goal(metric_direction: Text) -> Text:
    """'max' <-> 'MAXIMIZE', 'min' <-> 'MINIMIZE'; anything else is the service's "unspecified"."""
    for table in (_GOAL_OF_DIRECTION, _DIRECTION_OF_GOAL):
        if metric_direction in table:
            return table[metric_direction]
    return "GOAL_TYPE_UNSPECIFIED"


def _get_scale_type(sampling) -> Dict[Text, Text]:
    return {"scale_type": _SCALE_OF_SAMPLING.get(sampling, "SCALE_TYPE_UNSPECIFIED")}


def _format_sampling(scale_type: Optional[Text]) -> Optional[Text]:
    return _SAMPLING_OF_SCALE.get(scale_type)


# ----------------------------------------------------------------------------------------------
# objectives
# ----------------------------------------------------------------------------------------------
def _objective(name, direction=None):
    return oracle_module.Objective(
        name=name, direction=direction or metrics_tracking.infer_metric_direction(name))


def format_objective(objective, direction: Text = None) -> Optional[List[oracle_module.Objective]]:
    """A metric name, an ``Objective``, or a homogeneous list of either -> list of ``Objective``."""
    if isinstance(objective, oracle_module.Objective):
        return [objective]
    if isinstance(objective, str):
        return [_objective(objective, direction)]
    if isinstance(objective, list) and objective:
        kinds = {type(o) for o in objective}
        if kinds == {oracle_module.Objective}:
            return objective
        if kinds == {str}:
            return [_objective(o) for o in objective]
    if isinstance(objective, list) and not objective:
        return objective
    raise TypeError("Objective should be either string or oracle_module.Objective, found {}"
                    .format(objective))


def convert_study_config_to_objective(study_config):
    """``study_config["metrics"]`` -> list of ``Objective``."""
    metrics = study_config.get("metrics")
    if not metrics:
        raise ValueError('"metrics" not found in study_config {}'.format(study_config))
    if not isinstance(metrics, list):
        raise ValueError('study_config["metrics"] should be a list of {"metric": ..., "goal": ...}')
    return [oracle_module.Objective(name=m["metric"], direction=format_goal(m["goal"])) for m in metrics]


# ----------------------------------------------------------------------------------------------
# service parameter specs -> hyperparameters
# ----------------------------------------------------------------------------------------------
class _Kind(NamedTuple):
    spec_key: Text                                  # where the ParameterSpec keeps this kind's details
    check: Callable[[Dict[Text, Any]], Optional[Text]]     # spec -> complaint (None = fine)
    declare: Callable[[hp_module.HyperParameters, Text, Dict[Text, Any], Optional[Text]], Any]


def _needs_values(spec):
    if not isinstance(spec.get("values"), list):
        return 'Parameter spec {} is missing "values".'.format(spec)
    return None


def _needs_range(number_type):
    def check(spec):
        lo, hi = spec.get("min_value"), spec.get("max_value")
        if not (isinstance(lo, number_type) and isinstance(hi, number_type)):
            return 'Parameter spec {} requires both "min_value" and "max_value".'.format(spec)
        return None
    return check


def _declare_choice(hps, name, spec, _sampling):
    return hps.Choice(name, spec["values"])


def _declare_float(hps, name, spec, sampling):
    return hps.Float(name, min_value=spec["min_value"], max_value=spec["max_value"], sampling=sampling)


def _declare_int(hps, name, spec, sampling):
    return hps.Int(name, min_value=spec["min_value"], max_value=spec["max_value"], sampling=sampling)


_KINDS: Dict[Text, _Kind] = {
    "DISCRETE": _Kind("discrete_value_spec", _needs_values, _declare_choice),
    "CATEGORICAL": _Kind("categorical_value_spec", _needs_values, _declare_choice),
    "DOUBLE": _Kind("double_value_spec", _needs_range(float), _declare_float),
    "INTEGER": _Kind("integer_value_spec", _needs_range(int), _declare_int),
}


def _is_parameter_valid(param) -> _Kind:
    """Raises ``ValueError`` for a malformed ParameterSpec; returns its table row otherwise."""
    if not param.get("parameter"):
        raise ValueError('"parameter" (name) is not specified.')
    if not param.get("type"):
        raise ValueError("Parameter {} type is not specified.".format(param))
    kind = _KINDS.get(param["type"])
    if kind is None:
        raise ValueError("Unknown parameter type: {}.".format(param["type"]))
    if not param.get(kind.spec_key):
        raise ValueError("Parameter {} is missing {}.".format(param, kind.spec_key))
    complaint = kind.check(param[kind.spec_key])
    if complaint:
        raise ValueError(complaint)
    return kind


def convert_study_config_to_hps(study_config) -> hp_module.HyperParameters:
    """``study_config["parameters"]`` -> a ``HyperParameters`` search space."""
    params = study_config.get("parameters")
    if not params:
        raise ValueError("Parameters are not found in the study_config: ", study_config)
    if not isinstance(params, list):
        raise ValueError("Parameters should be a list of parameter with at least 1 parameter, "
                         "found ", params)
    hps = hp_module.HyperParameters()
    for param in params:
        kind = _is_parameter_valid(param)
        kind.declare(hps, param["parameter"], param[kind.spec_key], _format_sampling(param.get("scale_type")))
    return hps


# ----------------------------------------------------------------------------------------------
# hyperparameters -> service parameter specs
# ----------------------------------------------------------------------------------------------
def _spec(kind_name, payload, sampling=None):
    out = {"type": kind_name, _KINDS[kind_name].spec_key: payload}
    if sampling is not None:
        out.update(_get_scale_type(sampling))
    return out


def _values_spec(values):
    """A list of values is CATEGORICAL when it holds strings, DISCRETE (numbers) otherwise."""
    return _spec("CATEGORICAL" if isinstance(values[0], str) else "DISCRETE", {"values": values})


def _encode_choice(hp):
    return _values_spec(hp.values)


def _encode_int(hp):
    if hp.step not in (None, 1):              # a stepped range is enumerated
        return _spec("DISCRETE", {"values": list(range(hp.min_value, hp.max_value, hp.step))})
    return _spec("INTEGER", {"min_value": hp.min_value, "max_value": hp.max_value}, hp.sampling)


def _encode_float(hp):
    if hp.step is None:
        return _spec("DOUBLE", {"min_value": hp.min_value, "max_value": hp.max_value}, hp.sampling)
    grid, v = [], hp.min_value                # repeated addition, like the values the tuner samples
    while v < hp.max_value:
        grid.append(v)
        v += hp.step
    return _spec("DISCRETE", {"values": grid})


def _encode_boolean(_hp):
    return _spec("CATEGORICAL", {"values": ["True", "False"]})


def _encode_fixed(hp):
    if isinstance(hp.value, (str, bool)):
        return _spec("CATEGORICAL", {"values": [str(hp.value)]})
    return _spec("DISCRETE", {"values": [float(hp.value)]})


_ENCODERS = ((hp_module.Choice, _encode_choice), (hp_module.Int, _encode_int),
             (hp_module.Float, _encode_float), (hp_module.Boolean, _encode_boolean),
             (hp_module.Fixed, _encode_fixed))


def _convert_hyperparams_to_optimizer_params(hyperparams) -> List[Any]:
    """Every hyperparameter of the search space as a ParameterSpec dictionary."""
    specs = []
    for hp in hyperparams.space:
        for cls, encode in _ENCODERS:
            if isinstance(hp, cls):
                specs.append({"parameter": hp.name, **encode(hp)})
                break
        else:
            raise ValueError("`HyperParameter` type not recognized: {}".format(hp))
    return specs


def make_study_config(objective, hyperparams) -> Dict[Text, Any]:
    """The StudyConfig the optimizer service is created with: service-default algorithm, decay-curve
    automated stopping on elapsed time, one metric per objective, one ParameterSpec per hyperparameter."""
    return {
        "algorithm": "ALGORITHM_UNSPECIFIED",
        "automatedStoppingConfig": {"decayCurveStoppingConfig": {"useElapsedTime": True}},
        "metrics": [{"metric": o.name, "goal": format_goal(o.direction)} for o in format_objective(objective)],
        "parameters": _convert_hyperparams_to_optimizer_params(hyperparams),
    }


# ----------------------------------------------------------------------------------------------
# trials
# ----------------------------------------------------------------------------------------------
def get_trial_id(optimizer_trial: Dict[Text, Any]) -> Text:
    """``projects/.../studies/<study>/trials/<id>`` -> ``<id>``."""
    return optimizer_trial["name"].rsplit("/", 1)[-1]


_VALUE_FIELDS = (("floatValue", float), ("intValue", int), ("stringValue", str))


def convert_optimizer_trial_to_hps(hps, optimizer_trial) -> hp_module.HyperParameters:
    """A copy of the search space ``hps`` whose values are the trial's suggested parameters."""
    out = hp_module.HyperParameters.from_config(hps.get_config())
    out.values = {}
    for param in optimizer_trial["parameters"]:
        for field, cast in _VALUE_FIELDS:
            if field in param:
                out.values[param["parameter"]] = cast(param[field])
    return out
