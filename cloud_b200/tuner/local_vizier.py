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

Algorithms: ``ALGORITHM_UNSPECIFIED`` and ``RANDOM_SEARCH`` -> seeded random search honouring
``scale_type``; ``GRID_SEARCH`` -> lexicographic grid over DISCRETE / CATEGORICAL / INTEGER;
``GAUSSIAN_PROCESS_BANDIT`` -> random exploration mixed with Gaussian perturbations of the best
trials so far (a bandit-flavoured stand-in, not a GP).  Early stopping: median rule at equal step
for both ``decayCurveStoppingConfig`` and ``medianAutomatedStoppingConfig``.
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

from . import constants
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
                tid = str(len(trials) + 1)
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
            _study, trials = self._load(d)
            tid = str(len(trials) + 1)
            trial = dict(body)
            trial.update({"name": "{}/trials/{}".format(parent, tid), "startTime": _now()})
            trial.setdefault("state", "REQUESTED")
            trial.setdefault("measurements", [])
            trials.append(trial)
            self._write(os.path.join(d, "trials.json"), trials)
            return trial

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
            should_stop = self._median_rule(study, trials, cur)
            return self._new_operation(d, study_name, {"shouldStop": should_stop})

    def _median_rule(self, study, trials, cur) -> bool:
        cfg = study.get("studyConfig", {}).get("automatedStoppingConfig")
        if not cfg or not cur.get("measurements"):
            return False
        metric, goal = self._goal(study)

        def value_at(t, step):
            best = None
            for m in t.get("measurements", []):
                if int(m.get("stepCount", 0)) > step:
                    continue
                for mv in m.get("metrics", []):
                    if metric is None or mv.get("metric") == metric:
                        v = float(mv["value"])
                        if best is None or (v > best if goal == "MAXIMIZE" else v < best):
                            best = v
            return best
        step = int(cur["measurements"][-1].get("stepCount", 0))
        mine = value_at(cur, step)
        others = [value_at(t, step) for t in trials
                  if t["name"] != cur["name"] and t["state"] == "COMPLETED" and not t.get("trialInfeasible")]
        others = sorted(v for v in others if v is not None)
        if mine is None or len(others) < 3 or step < 1:
            return False
        median = others[len(others) // 2]
        return mine < median if goal == "MAXIMIZE" else mine > median

    # -- suggestion algorithms ----------------------------------------------------------------------------
    def _next_parameters(self, study, trials):
        cfg = study.get("studyConfig", {})
        algo = cfg.get("algorithm", "ALGORITHM_UNSPECIFIED")
        specs = cfg.get("parameters", [])
        rng = random.Random(study["seed"] + 7919 * len(trials))
        if algo == "GRID_SEARCH":
            return self._grid_point(study, specs)
        if algo == "GAUSSIAN_PROCESS_BANDIT":
            done = [t for t in trials if t["state"] == "COMPLETED" and t.get("finalMeasurement")
                    and not t.get("trialInfeasible")]
            if len(done) >= 4 and rng.random() < 0.6:
                _metric, goal = self._goal(study)
                done.sort(key=lambda t: t["finalMeasurement"]["metrics"][0]["value"],
                          reverse=goal == "MAXIMIZE")
                anchor = rng.choice(done[:max(1, len(done) // 4)])
                return self._perturb(specs, anchor["parameters"], rng)
        return [self._sample(spec, rng) for spec in specs]

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
        anchor = {}
        for p in anchor_params:
            anchor[p["parameter"]] = p.get("floatValue", p.get("intValue", p.get("stringValue")))
        out = []
        for spec in specs:
            kind = spec["type"]
            a = anchor.get(spec["parameter"])
            if a is None or kind in ("CATEGORICAL", "DISCRETE"):
                if a is not None and rng.random() < 0.7:
                    out.append(self._value_entry(spec, a))
                else:
                    out.append(self._sample(spec, rng))
                continue
            vs = spec.get("double_value_spec") or spec.get("integer_value_spec") or {}
            lo, hi = float(vs["min_value"]), float(vs["max_value"])
            log = (spec.get("scale_type") == "UNIT_LOG_SCALE") and lo > 0
            x = math.log(float(a)) if log else float(a)
            l2, h2 = (math.log(lo), math.log(hi)) if log else (lo, hi)
            x = min(max(rng.gauss(x, 0.15 * (h2 - l2)), l2), h2)
            v = math.exp(x) if log else x
            if kind == "INTEGER":
                v = int(round(v))
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
