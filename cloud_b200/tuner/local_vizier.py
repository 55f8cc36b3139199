"""A local, file-backed optimizer service with the verbs and JSON shapes of the CAIP Optimizer
("Vizier") REST API the reference talks to (tuner/optimizer_client.py; wire contract in
tuner/api/ml_public_google_rest_v1.json:846-1366 - here ``api/local_optimizer_v1.json``).

It is a library plus an on-disk store, not a server: every verb takes an exclusive ``flock`` on
the study directory, reads ``study.json`` / ``trials.json``, mutates and atomically rewrites them,
so any number of tuner processes (one per GPU, see ``scheduler.py``) can share a study - the
situation ``create_or_load_study`` was designed for (optimizer_client.py:378-381).

Service semantics reproduced: ``studies.create`` on an existing id ->
409; ``trials.suggest`` / ``checkEarlyStoppingState`` return long-running operations that are
polled with ``operations.get``; the same ``client_id`` gets the same trial back while it is still
ACTIVE; more than ``MAX_TRIALS_PER_STUDY`` trials or an exhausted grid -> 429; ``complete`` without
a final measurement uses the last reported one; trial names end in a numeric id.

Algorithms: ``RANDOM_SEARCH`` -> seeded random search honouring
``scale_type``; ``GRID_SEARCH`` -> lexicographic grid over DISCRETE / CATEGORICAL / INTEGER;
``GAUSSIAN_PROCESS_BANDIT`` - and ``ALGORITHM_UNSPECIFIED``, whose meaning is "the service's
default", unless ``TFC_VIZIER_DEFAULT_ALGORITHM`` says otherwise - -> Gaussian-process regression
with expected improvement (``gp_bandit.py``) after a few random trials.  Early stopping:
``medianAutomatedStoppingConfig`` -> median rule at equal progress; ``decayCurveStoppingConfig``
-> optimistic extrapolation of the learning curve against the best finished trial; progress is
the step count, or the elapsed time with ``useElapsedTime``.
"""
from __future__ import annotations

import contextlib
import fcntl
import itertools
import json
import math
import os
import random
import time
import uuid

from . import constants, gp_bandit
from ..utils import api_client


class _Resp:
    def __init__(self, status):
        self.status = status
        self.reason = {404: "Not Found", 409: "Conflict", 429: "Too Many Requests",
                       400: "Bad Request"}.get(status, "Error")


class HttpError(Exception):
    """Shape-compatible with ``googleapiclient.errors.HttpError`` (``e.resp.status``)."""

    def __init__(self, status, message=""):
        super().__init__("<HttpError {} \"{}\">".format(status, message))
        self.resp = _Resp(status)
        self.status_code = status
        self.content = message.encode()


MIN_TRIALS_FOR_STOPPING = 3
DECAY_CURVE_OPTIMISM = 2.0
GP_RANDOM_CANDIDATES = 384
GP_ANCHORS = 4
GP_PERTURBATIONS_PER_ANCHOR = 32


def default_root() -> str:
    root = os.environ.get("TFC_VIZIER_ROOT")
    if not root:
        from ..core.containerize import job_root
        root = os.path.join(job_root(), "vizier")
    os.makedirs(root, exist_ok=True)
    return root


def _now():
    return time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())


class LocalVizier:
    def __init__(self, root: str | None = None):
        self.root = root or default_root()

    # -- storage ------------------------------------------------------------------------------------
    def _study_dir(self, study_name: str) -> str:
        # projects/<p>/locations/<r>/studies/<s>
        parts = study_name.strip("/").split("/")
        if len(parts) < 6 or parts[0] != "projects" or parts[2] != "locations" or parts[4] != "studies":
            raise HttpError(400, "malformed study name: " + study_name)
        return os.path.join(self.root, parts[1], parts[3], parts[5])

    @staticmethod
    def _split_trial(trial_name: str):
        study, _, trial_id = trial_name.rpartition("/trials/")
        if not study:
            raise HttpError(400, "malformed trial name: " + trial_name)
        return study, trial_id

    @contextlib.contextmanager
    def _locked(self, directory: str, create: bool = False):
        if create:
            os.makedirs(directory, exist_ok=True)
        elif not os.path.isdir(directory):
            raise HttpError(404, "study not found")
        fd = os.open(os.path.join(directory, ".lock"), os.O_CREAT | os.O_RDWR, 0o644)
        try:
            fcntl.flock(fd, fcntl.LOCK_EX)
            yield
        finally:
            fcntl.flock(fd, fcntl.LOCK_UN)
            os.close(fd)

    @staticmethod
    def _read(path, default):
        if not os.path.exists(path):
            return default
        with open(path) as f:
            return json.load(f)

    @staticmethod
    def _write(path, obj):
        tmp = "{}.{}.tmp".format(path, os.getpid())
        with open(tmp, "w") as f:
            json.dump(obj, f)
        os.replace(tmp, path)

    def _load(self, d):
        study = self._read(os.path.join(d, "study.json"), None)
        if study is None:
            raise HttpError(404, "study not found")
        return study, self._read(os.path.join(d, "trials.json"), [])

    # -- studies ------------------------------------------------------------------------------------
    def create_study(self, parent: str, study_id: str, body: dict) -> dict:
        name = "{}/studies/{}".format(parent.rstrip("/"), study_id)
        d = self._study_dir(name)
        with self._locked(d, create=True):
            path = os.path.join(d, "study.json")
            if os.path.exists(path):
                raise HttpError(409, "study {} already exists".format(study_id))
            study = {"name": name, "studyConfig": body.get("study_config") or body.get("studyConfig") or {},
                     "state": "ACTIVE", "createTime": _now(), "client": api_client.user_agent(),
                     "seed": random.randint(0, 2**31 - 1), "grid_cursor": 0}
            self._write(path, study)
            self._write(os.path.join(d, "trials.json"), [])
            return self._public_study(study)

    @staticmethod
    def _public_study(study):
        return {k: study[k] for k in ("name", "studyConfig", "state", "createTime")}

    def get_study(self, name: str) -> dict:
        d = self._study_dir(name)
        with self._locked(d):
            return self._public_study(self._load(d)[0])

    def list_studies(self, parent: str) -> dict:
        parts = parent.strip("/").split("/")
        base = os.path.join(self.root, parts[1], parts[3]) if len(parts) >= 4 else self.root
        out = []
        if os.path.isdir(base):
            for sid in sorted(os.listdir(base)):
                path = os.path.join(base, sid, "study.json")
                if os.path.exists(path):
                    out.append(self._public_study(self._read(path, {})))
        return {"studies": out}

    def delete_study(self, name: str) -> dict:
        d = self._study_dir(name)
        if not os.path.exists(os.path.join(d, "study.json")):
            raise HttpError(404, "study not found")
        with self._locked(d):
            for f in ("study.json", "trials.json", "operations.json"):
                p = os.path.join(d, f)
                if os.path.exists(p):
                    os.remove(p)
        return {}

    def set_study_state(self, name: str, state: str) -> None:
        d = self._study_dir(name)
        with self._locked(d):
            study, _ = self._load(d)
            study["state"] = state
            self._write(os.path.join(d, "study.json"), study)

    # -- operations -----------------------------------------------------------------------------------
    def _new_operation(self, d, study_name, response):
        ops = self._read(os.path.join(d, "operations.json"), {})
        op_name = "{}/operations/{}".format(study_name.split("/studies/")[0], uuid.uuid4().hex[:16])
        ops[op_name] = {"name": op_name, "done": True, "response": response}
        if len(ops) > 256:       # keep the file small
            for k in list(ops)[:-128]:
                ops.pop(k)
        self._write(os.path.join(d, "operations.json"), ops)
        self._op_index = getattr(self, "_op_index", {})
        self._op_index[op_name] = d
        return {"name": op_name, "done": False}     # first answer is "still running", like the service

    def get_operation(self, name: str, study_name: str | None = None) -> dict:
        d = getattr(self, "_op_index", {}).get(name)
        if d is None and study_name is not None:
            d = self._study_dir(study_name)
        if d is None:
            raise HttpError(404, "operation not found")
        with self._locked(d):
            ops = self._read(os.path.join(d, "operations.json"), {})
        if name not in ops:
            raise HttpError(404, "operation not found")
        return ops[name]

    # -- trials ---------------------------------------------------------------------------------------
    def suggest_trials(self, parent: str, body: dict) -> dict:
        client_id = body.get("client_id") or body.get("clientId") or ""
        count = int(body.get("suggestion_count") or body.get("suggestionCount") or 1)
        d = self._study_dir(parent)
        with self._locked(d):
            study, trials = self._load(d)
            if study["state"] != "ACTIVE":
                return self._new_operation(d, parent, {"studyState": study["state"]})
            pending = [t for t in trials if t["state"] == "ACTIVE" and t.get("clientId") == client_id]
            out = pending[:count]
            while len(out) < count:
                if len(trials) >= constants.MAX_TRIALS_PER_STUDY:
                    if out:
                        break
                    raise HttpError(429, "trial limit of the study reached")
                params = self._next_parameters(study, trials)
                if params is None:
                    if out:
                        break
                    raise HttpError(429, "search space exhausted")
                tid = self._next_trial_id(study, trials)
                trial = {"name": "{}/trials/{}".format(parent, tid), "state": "ACTIVE",
                         "parameters": params, "clientId": client_id, "startTime": _now(),
                         "measurements": []}
                trials.append(trial)
                out.append(trial)
            self._write(os.path.join(d, "study.json"), study)
            self._write(os.path.join(d, "trials.json"), trials)
            return self._new_operation(d, parent, {"trials": out, "studyState": study["state"]})

    def create_trial(self, parent: str, body: dict) -> dict:
        d = self._study_dir(parent)
        with self._locked(d):
            study, trials = self._load(d)
            tid = self._next_trial_id(study, trials)
            self._write(os.path.join(d, "study.json"), study)
            trial = dict(body)
            trial.update({"name": "{}/trials/{}".format(parent, tid), "startTime": _now()})
            trial.setdefault("state", "REQUESTED")
            trial.setdefault("measurements", [])
            trials.append(trial)
            self._write(os.path.join(d, "trials.json"), trials)
            return trial

    @staticmethod
    def _next_trial_id(study, trials) -> str:
        """Trial ids are never reused, also not after ``trials.delete``."""
        last = max([int(study.get("last_trial_id", 0))] +
                   [int(t["name"].rpartition("/")[2]) for t in trials if t["name"].rpartition("/")[2].isdigit()])
        study["last_trial_id"] = last + 1
        return str(last + 1)

    def list_trials(self, parent: str) -> dict:
        d = self._study_dir(parent)
        with self._locked(d):
            return {"trials": self._load(d)[1]}

    def _mutate_trial(self, trial_name, fn):
        study_name, _tid = self._split_trial(trial_name)
        d = self._study_dir(study_name)
        with self._locked(d):
            study, trials = self._load(d)
            for t in trials:
                if t["name"] == trial_name:
                    result = fn(study, trials, t)
                    self._write(os.path.join(d, "trials.json"), trials)
                    return result if result is not None else t
        raise HttpError(404, "trial not found")

    def get_trial(self, name: str) -> dict:
        return self._mutate_trial(name, lambda s, ts, t: t)

    def delete_trial(self, name: str) -> dict:
        study_name, _ = self._split_trial(name)
        d = self._study_dir(study_name)
        with self._locked(d):
            _study, trials = self._load(d)
            kept = [t for t in trials if t["name"] != name]
            if len(kept) == len(trials):
                raise HttpError(404, "trial not found")
            self._write(os.path.join(d, "trials.json"), kept)
        return {}

    def add_measurement(self, name: str, body: dict) -> dict:
        m = body.get("measurement") or {}

        def fn(_study, _trials, t):
            t.setdefault("measurements", []).append(m)
        return self._mutate_trial(name, fn)

    def complete_trial(self, name: str, body: dict) -> dict:
        infeasible = bool(body.get("trial_infeasible") or body.get("trialInfeasible"))
        reason = body.get("infeasible_reason") or body.get("infeasibleReason")
        final = body.get("final_measurement") or body.get("finalMeasurement")

        def fn(study, _trials, t):
            t["state"] = "COMPLETED"
            t["endTime"] = _now()
            if infeasible:
                t["trialInfeasible"] = True
                if reason:
                    t["infeasibleReason"] = reason
                return
            if final:
                t["finalMeasurement"] = final
            elif t.get("measurements"):
                t["finalMeasurement"] = self._best_measurement(study, t)
            else:
                raise HttpError(400, "trial has no measurement to complete with")
        return self._mutate_trial(name, fn)

    @staticmethod
    def _goal(study):
        metrics = study.get("studyConfig", {}).get("metrics") or []
        if metrics:
            return metrics[0].get("metric"), metrics[0].get("goal", "MAXIMIZE")
        return None, "MAXIMIZE"

    def _best_measurement(self, study, t):
        """The service completes with the last measurement; ties in spirit with Keras' best-step
        bookkeeping are left to the caller."""
        return t["measurements"][-1]

    def stop_trial(self, name: str) -> dict:
        def fn(_s, _ts, t):
            if t["state"] == "ACTIVE":
                t["state"] = "STOPPING"
        return self._mutate_trial(name, fn)

    def check_early_stopping_state(self, name: str) -> dict:
        study_name, _ = self._split_trial(name)
        d = self._study_dir(study_name)
        with self._locked(d):
            study, trials = self._load(d)
            cur = next((t for t in trials if t["name"] == name), None)
            if cur is None:
                raise HttpError(404, "trial not found")
            should_stop = self._should_stop(study, trials, cur)
            return self._new_operation(d, study_name, {"shouldStop": should_stop})

    # -- automated stopping -----------------------------------------------------------------------------
    @staticmethod
    def _progress(measurement, use_elapsed_time: bool) -> float:
        """Where on its learning curve a measurement sits: step count, or seconds since the trial
        started when the config says ``useElapsedTime`` (what make_study_config asks for,
        tuner/utils.py:62-68)."""
        if use_elapsed_time:
            elapsed = measurement.get("elapsedTime")
            if isinstance(elapsed, dict):
                return float(elapsed.get("seconds", 0))
            if isinstance(elapsed, str):
                return float(elapsed.rstrip("s") or 0)
        return float(measurement.get("stepCount", 0))

    def _curve(self, study, trial, use_elapsed_time):
        """[(progress, objective value)] of a trial, in the order reported."""
        metric, _goal = self._goal(study)
        out = []
        for m in trial.get("measurements", []):
            for mv in m.get("metrics", []):
                if metric is None or mv.get("metric") == metric:
                    out.append((self._progress(m, use_elapsed_time), float(mv["value"])))
                    break
        return out

    @staticmethod
    def _best_until(curve, progress, maximize):
        values = [v for p, v in curve if p <= progress]
        if not values:
            return None
        return max(values) if maximize else min(values)

    def _should_stop(self, study, trials, cur) -> bool:
        cfg = study.get("studyConfig", {}).get("automatedStoppingConfig")
        if not cfg or not cur.get("measurements"):
            return False
        decay = cfg.get("decayCurveStoppingConfig")
        median = cfg.get("medianAutomatedStoppingConfig")
        rule_cfg = decay if decay is not None else (median or {})
        use_elapsed = bool(rule_cfg.get("useElapsedTime"))
        _metric, goal = self._goal(study)
        maximize = goal == "MAXIMIZE"
        mine = self._curve(study, cur, use_elapsed)
        if not mine:
            return False
        # a trial whose measurements carry no elapsed time falls back to step counts
        if use_elapsed and all(p == 0 for p, _v in mine):
            use_elapsed = False
            mine = self._curve(study, cur, False)
        now = mine[-1][0]
        finished = [self._curve(study, t, use_elapsed) for t in trials
                    if t["name"] != cur["name"] and t["state"] == "COMPLETED"
                    and not t.get("trialInfeasible")]
        finished = [c for c in finished if c]
        if len(finished) < MIN_TRIALS_FOR_STOPPING or len(mine) < 2 and now < 1:
            return False
        if decay is not None:
            return self._decay_curve_rule(mine, finished, now, maximize)
        return self._median_rule(mine, finished, now, maximize)

    def _median_rule(self, mine, finished, now, maximize) -> bool:
        """Stop when the trial's best value so far is worse than the median of what the finished
        trials had reached at the same point."""
        current = self._best_until(mine, now, maximize)
        others = sorted(v for v in (self._best_until(c, now, maximize) for c in finished) if v is not None)
        if current is None or len(others) < MIN_TRIALS_FOR_STOPPING:
            return False
        middle = others[len(others) // 2]
        return current < middle if maximize else current > middle

    def _decay_curve_rule(self, mine, finished, now, maximize) -> bool:
        """Stop when an optimistic extrapolation of the trial's curve ends below the best finished
        trial.  The extrapolation adds to the current value ``DECAY_CURVE_OPTIMISM`` times the
        upper-quartile gain that finished trials still made after the same point of their own
        curves - learning curves flatten, so late in a trial that gain is small and a hopeless trial
        is cut, while early on it is large and only trials far behind are stopped."""
        sign = 1.0 if maximize else -1.0
        current = self._best_until(mine, now, maximize)
        gains, finals = [], []
        for curve in finished:
            final = self._best_until(curve, float("inf"), maximize)
            at_now = self._best_until(curve, now, maximize)
            finals.append(sign * final)
            if at_now is not None:      # a curve that ended before `now` gained nothing after it
                gains.append(sign * (final - at_now))
        if current is None or len(gains) < MIN_TRIALS_FOR_STOPPING:
            return False
        gains.sort()
        optimistic_gain = DECAY_CURVE_OPTIMISM * max(0.0, gains[min(len(gains) - 1, (3 * len(gains)) // 4)])
        return sign * current + optimistic_gain < max(finals)

    # -- suggestion algorithms ----------------------------------------------------------------------------
    def _next_parameters(self, study, trials):
        cfg = study.get("studyConfig", {})
        algo = cfg.get("algorithm", "ALGORITHM_UNSPECIFIED")
        if algo == "ALGORITHM_UNSPECIFIED":     # the service's default is its GP bandit
            algo = os.environ.get("TFC_VIZIER_DEFAULT_ALGORITHM", "GAUSSIAN_PROCESS_BANDIT")
        specs = cfg.get("parameters", [])
        rng = random.Random(study["seed"] + 7919 * len(trials))
        if algo == "GRID_SEARCH":
            return self._grid_point(study, specs)
        if algo == "GAUSSIAN_PROCESS_BANDIT":
            params = self._gp_bandit_point(study, specs, trials, rng)
            if params is not None:
                return params
        return [self._sample(spec, rng) for spec in specs]

    def _objective_of(self, study, trial):
        metric, _goal = self._goal(study)
        for mv in (trial.get("finalMeasurement") or {}).get("metrics", []):
            if metric is None or mv.get("metric") == metric:
                return float(mv["value"])
        return None

    def _gp_bandit_point(self, study, specs, trials, rng):
        """Expected-improvement maximiser over random candidates and perturbations of the best
        trials (``gp_bandit.choose``); None until enough trials have finished."""
        _metric, goal = self._goal(study)
        observed, pending = [], []
        for t in trials:
            if t["state"] == "COMPLETED":
                value = None if t.get("trialInfeasible") else self._objective_of(study, t)
                if value is not None or t.get("trialInfeasible"):
                    observed.append((t["parameters"], value))
            elif t["state"] in ("ACTIVE", "STOPPING", "REQUESTED"):
                pending.append(t["parameters"])
        feasible = [(p, v) for p, v in observed if v is not None]
        feasible.sort(key=lambda pv: pv[1], reverse=goal == "MAXIMIZE")
        candidates = [[self._sample(spec, rng) for spec in specs] for _ in range(GP_RANDOM_CANDIDATES)]
        for anchor, _v in feasible[:GP_ANCHORS]:
            candidates += [self._perturb(specs, anchor, rng) for _ in range(GP_PERTURBATIONS_PER_ANCHOR)]
        seen = {json.dumps(p, sort_keys=True) for p, _v in observed} | \
            {json.dumps(p, sort_keys=True) for p in pending}
        fresh = [c for c in candidates if json.dumps(c, sort_keys=True) not in seen]
        candidates = fresh or candidates
        pick = gp_bandit.choose(specs, observed, pending, candidates, goal == "MAXIMIZE")
        return None if pick is None else candidates[pick]

    @staticmethod
    def _value_entry(spec, v):
        kind = spec["type"]
        if kind == "INTEGER":
            return {"parameter": spec["parameter"], "intValue": str(int(v))}
        if kind == "CATEGORICAL":
            return {"parameter": spec["parameter"], "stringValue": str(v)}
        return {"parameter": spec["parameter"], "floatValue": float(v)}

    def _sample(self, spec, rng):
        kind = spec["type"]
        scale = spec.get("scale_type") or spec.get("scaleType")
        if kind in ("DOUBLE", "INTEGER"):
            vs = spec.get("double_value_spec") or spec.get("doubleValueSpec") or \
                spec.get("integer_value_spec") or spec.get("integerValueSpec")
            lo, hi = float(vs.get("min_value", vs.get("minValue"))), float(vs.get("max_value", vs.get("maxValue")))
            u = rng.random()
            if scale == "UNIT_LOG_SCALE" and lo > 0:
                v = math.exp(math.log(lo) + u * (math.log(hi) - math.log(lo)))
            elif scale == "UNIT_REVERSE_LOG_SCALE" and lo > 0:
                v = hi + lo - math.exp(math.log(lo) + u * (math.log(hi) - math.log(lo)))
            else:
                v = lo + u * (hi - lo)
            if kind == "INTEGER":
                v = int(min(max(round(v), lo), hi))
            return self._value_entry(spec, v)
        vs = spec.get("discrete_value_spec") or spec.get("discreteValueSpec") or \
            spec.get("categorical_value_spec") or spec.get("categoricalValueSpec")
        return self._value_entry(spec, rng.choice(vs["values"]))

    def _perturb(self, specs, anchor_params, rng):
        """A neighbour of ``anchor_params``: numeric coordinates move by a Gaussian step of 10 % of
        their (scaled) range, discrete ones mostly to an adjacent value, categorical ones mostly stay."""
        anchor = {p["parameter"]: gp_bandit.param_value(p) for p in anchor_params}
        out = []
        for spec in specs:
            kind = spec["type"]
            a = anchor.get(spec["parameter"])
            if a is None:
                out.append(self._sample(spec, rng))
            elif kind == "CATEGORICAL":
                out.append(self._value_entry(spec, a) if rng.random() < 0.75 else self._sample(spec, rng))
            elif kind == "DISCRETE":
                vs = spec.get("discrete_value_spec") or spec.get("discreteValueSpec")
                values = sorted(vs["values"])
                near = min(range(len(values)), key=lambda i: abs(float(values[i]) - float(a)))
                near = min(max(near + rng.choice((-1, 0, 0, 1)), 0), len(values) - 1)
                out.append(self._value_entry(spec, values[near]))
            else:
                vs = spec.get("double_value_spec") or spec.get("doubleValueSpec") or \
                    spec.get("integer_value_spec") or spec.get("integerValueSpec")
                lo = float(vs.get("min_value", vs.get("minValue")))
                hi = float(vs.get("max_value", vs.get("maxValue")))
                log = (spec.get("scale_type") or spec.get("scaleType")) == "UNIT_LOG_SCALE" and lo > 0
                x = math.log(float(a)) if log else float(a)
                l2, h2 = (math.log(lo), math.log(hi)) if log else (lo, hi)
                x = min(max(rng.gauss(x, 0.1 * (h2 - l2)), l2), h2)
                v = math.exp(x) if log else x
                if kind == "INTEGER":
                    v = int(min(max(round(v), lo), hi))
                out.append(self._value_entry(spec, v))
        return out

    def _grid_point(self, study, specs):
        axes = []
        for spec in specs:
            kind = spec["type"]
            if kind == "DOUBLE":
                raise HttpError(400, "GRID_SEARCH does not support DOUBLE parameters")
            if kind == "INTEGER":
                vs = spec.get("integer_value_spec") or spec.get("integerValueSpec")
                axes.append(list(range(int(vs["min_value"]), int(vs["max_value"]) + 1)))
            else:
                vs = spec.get("discrete_value_spec") or spec.get("categorical_value_spec")
                axes.append(list(vs["values"]))
        total = 1
        for a in axes:
            total *= len(a)
        cursor = study.get("grid_cursor", 0)
        if cursor >= total:
            return None
        study["grid_cursor"] = cursor + 1
        point = next(itertools.islice(itertools.product(*axes), cursor, None))
        return [self._value_entry(s, v) for s, v in zip(specs, point)]
