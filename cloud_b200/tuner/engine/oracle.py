"""``Objective`` and the ``Oracle`` base class (KerasTuner engine; CloudOracle subclasses it,
tuner/tuner.py:32-119).  State is persisted as ``<dir>/oracle.json`` + ``trial_<id>/trial.json``."""
from __future__ import annotations

import collections
import json
import os
import random

from . import hyperparameters as hp_module
from . import metrics_tracking
from . import trial as trial_lib

Objective = collections.namedtuple("Objective", "name direction")


class Oracle:
    def __init__(self, objective, max_trials=None, hyperparameters=None, allow_new_entries=True,
                 tune_new_entries=True, seed=None):
        self.objective = _format_objective(objective)
        self.max_trials = max_trials
        if not hyperparameters:
            if not tune_new_entries:
                raise ValueError("If you set `tune_new_entries=False`, you must specify the search "
                                 "space via the `hyperparameters` argument.")
            if not allow_new_entries:
                raise ValueError("If you set `allow_new_entries=False`, you must specify the search "
                                 "space via the `hyperparameters` argument.")
            self.hyperparameters = hp_module.HyperParameters()
        else:
            self.hyperparameters = hyperparameters
        self.allow_new_entries = allow_new_entries
        self.tune_new_entries = tune_new_entries
        self.trials = {}
        self.ongoing_trials = {}
        self.seed = seed or random.randint(1, int(1e4))
        self._seed_state = self.seed
        self._directory = None
        self._project_name = None

    # -- to be overridden ---------------------------------------------------------------------------
    def _populate_space(self, trial_id):
        raise NotImplementedError

    def _score_trial(self, trial):
        pass

    # -- tuner-facing API ---------------------------------------------------------------------------
    def get_space(self):
        return self.hyperparameters.copy()

    def update_space(self, hyperparameters):
        """Adds new hyperparameters to the search space (no-op when new entries are not allowed)."""
        if not self.allow_new_entries:
            return
        known = {p.name for p in self.hyperparameters.space}
        for hp in hyperparameters.space:
            if hp.name not in known:
                self.hyperparameters.register(hp)

    def create_trial(self, tuner_id):
        if tuner_id in self.ongoing_trials:
            return self.ongoing_trials[tuner_id]
        trial_id = trial_lib.generate_trial_id()
        if self.max_trials and len(self.trials) >= self.max_trials:
            status, values = trial_lib.TrialStatus.STOPPED, None
        else:
            response = self._populate_space(trial_id)
            status, values = response["status"], response.get("values")
        hps = self.hyperparameters.copy()
        hps.values = values
        trial = trial_lib.Trial(hyperparameters=hps, trial_id=trial_id, status=status)
        if status == trial_lib.TrialStatus.RUNNING:
            self.ongoing_trials[tuner_id] = trial
            self.trials[trial_id] = trial
            self._save_trial(trial)
            self.save()
        return trial

    def update_trial(self, trial_id, metrics, step=0):
        trial = self.trials[trial_id]
        self._check_objective_found(metrics)
        for name, value in metrics.items():
            if not trial.metrics.exists(name):
                direction = _maybe_infer_direction_from_objective(self.objective, name)
                trial.metrics.register(name, direction=direction)
            trial.metrics.update(name, value, step=step)
        self._save_trial(trial)
        return trial.status

    def end_trial(self, trial_id, status="COMPLETED"):
        trial = None
        for tuner_id, ongoing in self.ongoing_trials.items():
            if ongoing.trial_id == trial_id:
                trial = self.ongoing_trials.pop(tuner_id)
                break
        if not trial:
            raise ValueError("Ongoing trial with id: {} not found.".format(trial_id))
        trial.status = status
        if status == trial_lib.TrialStatus.COMPLETED:
            self._score_trial(trial)
        self._save_trial(trial)
        self.save()

    def get_trial(self, trial_id):
        return self.trials[trial_id]

    def get_best_trials(self, num_trials=1):
        trials = [t for t in self.trials.values() if t.status == trial_lib.TrialStatus.COMPLETED
                  and t.score is not None]
        objective = self.objective[0] if isinstance(self.objective, list) else self.objective
        trials.sort(key=lambda t: t.score, reverse=objective.direction == "max")
        return trials[:num_trials]

    def remaining_trials(self):
        return self.max_trials - len(self.trials) if self.max_trials else None

    # -- persistence ----------------------------------------------------------------------------------
    def _set_project_dir(self, directory, project_name, overwrite=False):
        self._directory = directory
        self._project_name = project_name
        if not overwrite and os.path.exists(self._get_oracle_fname()):
            self.reload()

    @property
    def _project_dir(self):
        dirname = os.path.join(str(self._directory), self._project_name)
        os.makedirs(dirname, exist_ok=True)
        return dirname

    def _get_oracle_fname(self):
        return os.path.join(self._project_dir, "oracle.json")

    def _save_trial(self, trial):
        if self._directory is None:
            return
        trial.save(os.path.join(self._project_dir, "trial_" + str(trial.trial_id), "trial.json"))

    def get_state(self):
        return {"ongoing_trials": {tid: t.trial_id for tid, t in self.ongoing_trials.items()},
                "hyperparameters": self.hyperparameters.get_config(), "seed": self.seed,
                "seed_state": self._seed_state, "tried_so_far": []}

    def set_state(self, state):
        self.hyperparameters = hp_module.HyperParameters.from_config(state["hyperparameters"])
        self.seed = state.get("seed", self.seed)
        self._seed_state = state.get("seed_state", self._seed_state)
        self._ongoing_ids = state.get("ongoing_trials", {})

    def save(self):
        if self._directory is None:
            return
        fname = self._get_oracle_fname()
        tmp = "{}.{}.tmp".format(fname, os.getpid())
        with open(tmp, "w") as f:
            json.dump(self.get_state(), f)
        os.replace(tmp, fname)

    def reload(self):
        with open(self._get_oracle_fname()) as f:
            self.set_state(json.load(f))
        base = self._project_dir
        for name in sorted(os.listdir(base)):
            path = os.path.join(base, name, "trial.json")
            if name.startswith("trial_") and os.path.exists(path):
                t = trial_lib.Trial.load(path)
                self.trials[t.trial_id] = t
        for tuner_id, trial_id in getattr(self, "_ongoing_ids", {}).items():
            if trial_id in self.trials:
                self.ongoing_trials[tuner_id] = self.trials[trial_id]

    # -- helpers ---------------------------------------------------------------------------------------
    def _check_objective_found(self, metrics):
        if isinstance(self.objective, list):
            names = [o.name for o in self.objective]
        else:
            names = [self.objective.name]
        for name in names:
            if name not in metrics:
                raise ValueError("Objective value missing in metrics reported to the Oracle, "
                                 "expected: {}, found: {}".format(names, list(metrics.keys())))

    def _random_values(self):
        """One random point of the search space (seeded, advancing the oracle's seed state)."""
        values = {}
        for hp in self.hyperparameters.space:
            values[hp.name] = hp.random_sample(self._seed_state)
            self._seed_state += 1
        return values


def _format_objective(objective):
    if isinstance(objective, list):
        return [_format_objective(o) for o in objective]
    if isinstance(objective, Objective):
        return objective
    if isinstance(objective, str):
        direction = metrics_tracking.infer_metric_direction(objective)
        if direction is None:
            raise ValueError('Could not infer optimization direction ("min" or "max") for unknown '
                             'metric "{}". Please specify the objective as a `Objective`.'.format(objective))
        return Objective(name=objective, direction=direction)
    raise ValueError("`objective` not understood, expected str or `Objective` object, found: "
                     "{}".format(objective))


def _maybe_infer_direction_from_objective(objective, metric_name):
    objs = objective if isinstance(objective, list) else [objective]
    for o in objs:
        if o.name == metric_name:
            return o.direction
    return None


class RandomSearchOracle(Oracle):
    """Plain random search (KerasTuner's RandomSearch oracle) - the fully local tuner path."""

    def _populate_space(self, trial_id):
        # like KerasTuner, the first trial evaluates the declared defaults: a sensible baseline
        # that every later random point has to beat
        if not self.trials:
            values = {hp.name: hp.default for hp in self.hyperparameters.space}
        else:
            values = self._random_values()
        return {"status": trial_lib.TrialStatus.RUNNING, "values": values}

    def _score_trial(self, trial):
        obj = self.objective[0] if isinstance(self.objective, list) else self.objective
        if trial.metrics.exists(obj.name):
            trial.score = trial.metrics.get_best_value(obj.name)
            trial.best_step = trial.metrics.get_best_step(obj.name)


class BayesianOptimizationOracle(RandomSearchOracle):
    """KerasTuner's ``BayesianOptimization`` oracle on this framework's Gaussian-process bandit
    (``tuner/gp_bandit.py``, the model the local optimizer service uses too): the first trial
    evaluates the defaults, the next ``num_initial_points - 1`` are random, after that every trial is
    the random candidate with the largest expected improvement given the finished trials; running
    trials count as observations at their predicted value, so parallel workers spread out."""

    def __init__(self, objective, max_trials=None, num_initial_points=None, num_candidates=256,
                 hyperparameters=None, allow_new_entries=True, tune_new_entries=True, seed=None):
        super().__init__(objective, max_trials=max_trials, hyperparameters=hyperparameters,
                         allow_new_entries=allow_new_entries, tune_new_entries=tune_new_entries, seed=seed)
        self.num_initial_points = num_initial_points
        self.num_candidates = int(num_candidates)

    @staticmethod
    def _entries(specs, values):
        """A value dict in the optimizer service's parameter-list shape (what gp_bandit encodes)."""
        out = []
        for spec in specs:
            name = spec["parameter"]
            if name not in values:
                continue
            v = values[name]
            if spec["type"] == "CATEGORICAL":
                out.append({"parameter": name, "stringValue": str(v)})
            elif spec["type"] == "INTEGER":
                out.append({"parameter": name, "intValue": str(int(v))})
            else:
                out.append({"parameter": name, "floatValue": float(v)})
        return out

    def _neighbour(self, values):
        """A point near ``values``: numeric hyperparameters take a Gaussian step of 10 % of their
        (log-)range and are snapped to their step, the others mostly keep their value."""
        import math
        rng = random.Random(self._seed_state)
        self._seed_state += 1
        out = {}
        for hp in self.hyperparameters.space:
            v = values.get(hp.name, hp.default)
            if isinstance(hp, (hp_module.Int, hp_module.Float)):
                lo, hi = float(hp.min_value), float(hp.max_value)
                log = getattr(hp, "sampling", None) == "log" and lo > 0
                a, b = (math.log(lo), math.log(hi)) if log else (lo, hi)
                x = rng.gauss(math.log(float(v)) if log else float(v), 0.1 * (b - a))
                x = min(max(x, a), b)
                x = math.exp(x) if log else x
                step = getattr(hp, "step", None)
                if isinstance(hp, hp_module.Int):
                    step = step or 1
                if step:
                    x = lo + round((x - lo) / step) * step
                x = min(max(x, lo), hi)
                out[hp.name] = int(round(x)) if isinstance(hp, hp_module.Int) else x
            elif isinstance(hp, hp_module.Fixed):
                out[hp.name] = hp.value
            else:
                out[hp.name] = v if rng.random() < 0.75 else hp.random_sample(rng.randrange(1 << 30))
        return out

    def _populate_space(self, trial_id):
        from .. import gp_bandit, utils
        obj = self.objective[0] if isinstance(self.objective, list) else self.objective
        specs = utils._convert_hyperparams_to_optimizer_params(self.hyperparameters)
        done, running = [], []
        for t in self.trials.values():
            if t.status == trial_lib.TrialStatus.COMPLETED and t.score is not None:
                done.append((self._entries(specs, t.hyperparameters.values), float(t.score)))
            elif t.status == trial_lib.TrialStatus.INVALID:
                done.append((self._entries(specs, t.hyperparameters.values), None))
            elif t.status == trial_lib.TrialStatus.RUNNING:
                running.append(self._entries(specs, t.hyperparameters.values))
        n_init = self.num_initial_points or gp_bandit.min_initial_trials(gp_bandit.SpaceEncoder(specs).dims)
        if len(done) < n_init:
            return super()._populate_space(trial_id)
        candidates = [self._random_values() for _ in range(self.num_candidates)]
        ranked = sorted((d for d in done if d[1] is not None), key=lambda d: d[1],
                        reverse=obj.direction == "max")
        best_values = [t.hyperparameters.values for t in self.trials.values()
                       if t.score is not None and any(t.score == r[1] for r in ranked[:3])]
        for anchor in best_values[:3]:                   # exploit: neighbours of the best trials
            candidates += [self._neighbour(anchor) for _ in range(self.num_candidates // 8)]
        pick = gp_bandit.choose(specs, done, running, [self._entries(specs, c) for c in candidates],
                                maximize=obj.direction == "max")
        values = candidates[pick] if pick is not None else self._random_values()
        return {"status": trial_lib.TrialStatus.RUNNING, "values": values}
