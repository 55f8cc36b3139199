"""``Tuner``: the search loop (KerasTuner engine ``Tuner``; CloudTuner subclasses it,
tuner/tuner.py:325-376).

``search`` = ``on_search_begin``; loop { ``oracle.create_trial(tuner_id)``; STOPPED -> break;
IDLE -> continue; ``on_trial_begin``; ``run_trial``; ``on_trial_end`` }; ``on_search_end``.
``run_trial`` builds the model from the trial's hyperparameters and calls ``fit`` with a
``TunerCallback`` (reports every epoch to the oracle and stops training when the oracle says
STOPPED) and a best-only weights checkpoint under
``<directory>/<project_name>/trial_<id>/checkpoints/epoch_<n>/``.
"""
from __future__ import annotations

import copy
import json
import os

from . import hypermodel as hm_module
from . import trial as trial_lib
from .oracle import RandomSearchOracle


class TunerCallback:
    """Keras callback wiring ``fit`` to the oracle.  Built lazily on the Keras Callback base."""

    def __new__(cls, tuner, trial):
        from ...keras.callbacks import Callback

        class _TunerCallback(Callback):
            def __init__(self, tuner, trial):
                super().__init__()
                self.tuner, self.trial = tuner, trial

            def on_epoch_begin(self, epoch, logs=None):
                self.tuner.on_epoch_begin(self.trial, self.model, epoch, logs=logs)

            def on_epoch_end(self, epoch, logs=None):
                self.tuner.on_epoch_end(self.trial, self.model, epoch, logs=logs)

        return _TunerCallback(tuner, trial)


class Tuner:
    def __init__(self, oracle, hypermodel, max_model_size=None, optimizer=None, loss=None,
                 metrics=None, distribution_strategy=None, directory=None, project_name=None,
                 logger=None, tuner_id=None, overwrite=False, executions_per_trial=1):
        self.oracle = oracle
        self.hypermodel = hm_module.get_hypermodel(hypermodel)
        self.directory = str(directory) if directory is not None else "."
        self.project_name = project_name or "untitled_project"
        self.oracle._set_project_dir(self.directory, self.project_name, overwrite=overwrite)
        self.tuner_id = tuner_id or os.environ.get("KERASTUNER_TUNER_ID", "tuner0")
        self.optimizer, self.loss, self.metrics = optimizer, loss, metrics
        self.distribution_strategy = distribution_strategy
        self.executions_per_trial = executions_per_trial
        self.max_model_size = max_model_size
        self._display_trials = 0
        self._populate_initial_space()
        if not overwrite and os.path.exists(self._get_tuner_fname()):
            self.reload()

    # -- search space -------------------------------------------------------------------------------
    def _populate_initial_space(self):
        """Builds the hypermodel once so dynamically declared hyperparameters get registered
        (only when the oracle allows new entries - CloudOracle does not, tuner.py:87-93)."""
        if not getattr(self.oracle, "allow_new_entries", True):
            return
        hp = self.oracle.get_space()
        try:
            self.hypermodel.build(hp)
            self.oracle.update_space(hp)
        except Exception:
            pass

    def search_space_summary(self, extended=False):
        print("Search space summary")
        hp = self.oracle.get_space()
        print("|-Default search space size: %d" % len(hp.space))
        for p in hp.space:
            config = p.get_config()
            name = config.pop("name")
            print("%s (%s)" % (name, p.__class__.__name__))
            for k, v in config.items():
                print("|-%s: %s" % (k, v))

    # -- the loop -------------------------------------------------------------------------------------
    def search(self, *fit_args, **fit_kwargs):
        self.on_search_begin()
        while True:
            trial = self.oracle.create_trial(self.tuner_id)
            if trial.status == trial_lib.TrialStatus.STOPPED:
                break
            if trial.status == trial_lib.TrialStatus.IDLE:
                continue
            self.on_trial_begin(trial)
            self.run_trial(trial, *fit_args, **fit_kwargs)
            self.on_trial_end(trial)
        self.on_search_end()

    def _build_model(self, hp):
        model = self.hypermodel.build(hp)
        if self.optimizer or self.loss or self.metrics:
            model.compile(optimizer=self.optimizer or model.optimizer,
                          loss=self.loss or model.loss, metrics=self.metrics or model._compiled_with["metrics"])
        return model

    def run_trial(self, trial, *fit_args, **fit_kwargs):
        from ...keras.callbacks import ModelCheckpoint
        original_callbacks = fit_kwargs.pop("callbacks", [])
        for execution in range(self.executions_per_trial):
            kwargs = dict(fit_kwargs)
            callbacks = self._deepcopy_callbacks(original_callbacks)
            callbacks.append(TunerCallback(self, trial))
            objective = self.oracle.objective[0] if isinstance(self.oracle.objective, list) \
                else self.oracle.objective
            callbacks.append(ModelCheckpoint(
                filepath=os.path.join(self._get_checkpoint_dir(trial.trial_id, "{epoch}"), "checkpoint"),
                monitor=objective.name, mode=objective.direction, save_best_only=True,
                save_weights_only=True))
            kwargs["callbacks"] = callbacks
            model = self._build_model(trial.hyperparameters)
            model.fit(*fit_args, **kwargs)

    def _deepcopy_callbacks(self, callbacks):
        try:
            return copy.deepcopy(list(callbacks))
        except Exception:
            raise ValueError("All callbacks used during a search should be deep-copyable (since "
                             "they are reused across trials). It is not possible to do "
                             "`copy.deepcopy(%s)`" % (callbacks,))

    # -- hooks ------------------------------------------------------------------------------------------
    def on_search_begin(self):
        pass

    def on_search_end(self):
        pass

    def on_trial_begin(self, trial):
        self._display_trials += 1

    def on_epoch_begin(self, trial, model, epoch, logs=None):
        pass

    def on_epoch_end(self, trial, model, epoch, logs=None):
        self.save_model(trial.trial_id, model, step=epoch)
        status = self.oracle.update_trial(trial.trial_id, metrics=logs, step=epoch)
        trial.status = status
        if trial.status == trial_lib.TrialStatus.STOPPED:
            model.stop_training = True

    def on_trial_end(self, trial):
        self.oracle.end_trial(trial.trial_id, trial_lib.TrialStatus.COMPLETED)
        self.oracle.update_space(trial.hyperparameters)
        self.save()

    # -- results ----------------------------------------------------------------------------------------
    def get_best_hyperparameters(self, num_trials=1):
        return [t.hyperparameters for t in self.oracle.get_best_trials(num_trials)]

    def get_best_models(self, num_models=1):
        return [self.load_model(t) for t in self.oracle.get_best_trials(num_models)]

    def results_summary(self, num_trials=10):
        print("Results summary")
        print("|-Results in %s" % self.project_dir)
        best_trials = self.oracle.get_best_trials(num_trials)
        obj = self.oracle.objective
        print("|-Showing %d best trials" % num_trials)
        print("|-%s" % (obj if not isinstance(obj, list) else obj[0],))
        for trial in best_trials:
            trial.summary()

    # -- persistence --------------------------------------------------------------------------------------
    @property
    def project_dir(self):
        dirname = os.path.join(self.directory, self.project_name)
        os.makedirs(dirname, exist_ok=True)
        return dirname

    def get_trial_dir(self, trial_id):
        dirname = os.path.join(self.project_dir, "trial_" + str(trial_id))
        os.makedirs(dirname, exist_ok=True)
        return dirname

    def _get_tuner_fname(self):
        return os.path.join(self.project_dir, str(self.tuner_id) + ".json")

    def _get_checkpoint_dir(self, trial_id, epoch):
        return os.path.join(self.get_trial_dir(trial_id), "checkpoints", "epoch_" + str(epoch))

    def save_model(self, trial_id, model, step=0):
        """Per-epoch hook kept for API parity; the best-only ModelCheckpoint does the writing."""

    def load_model(self, trial):
        model = self._build_model(trial.hyperparameters)
        best = trial.best_step if trial.best_step is not None else 0
        base = os.path.join(self.get_trial_dir(trial.trial_id), "checkpoints")
        candidates = []
        if os.path.isdir(base):
            for name in os.listdir(base):
                if name.startswith("epoch_"):
                    try:
                        candidates.append(int(name.split("_", 1)[1]))
                    except ValueError:
                        pass
        if not candidates:
            return model
        # checkpoints are written as epoch (1-based) = step + 1; take the one at/just before best
        want = best + 1
        chosen = max([c for c in candidates if c <= want] or [min(candidates)])
        path = os.path.join(base, "epoch_%d" % chosen, "checkpoint")
        if not model.built:
            raise RuntimeError("hypermodel.build() must return a built model (give the first layer "
                               "an input_shape) for get_best_models() to restore weights")
        model.load_weights(path)
        return model

    def get_state(self):
        return {"tuner_id": self.tuner_id, "display_trials": self._display_trials}

    def set_state(self, state):
        self._display_trials = state.get("display_trials", 0)

    def save(self):
        tmp = self._get_tuner_fname() + ".tmp"
        with open(tmp, "w") as f:
            json.dump(self.get_state(), f)
        os.replace(tmp, self._get_tuner_fname())

    def reload(self):
        with open(self._get_tuner_fname()) as f:
            self.set_state(json.load(f))


class RandomSearch(Tuner):
    """``kerastuner.tuners.RandomSearch`` (used by core/tests/testdata/keras_tuner_cifar_example.py:92-99)."""

    def __init__(self, hypermodel, objective, max_trials, seed=None, hyperparameters=None,
                 tune_new_entries=True, allow_new_entries=True, **kwargs):
        oracle = RandomSearchOracle(objective=objective, max_trials=max_trials, seed=seed,
                                    hyperparameters=hyperparameters,
                                    tune_new_entries=tune_new_entries,
                                    allow_new_entries=allow_new_entries)
        super().__init__(oracle, hypermodel, **kwargs)


class BayesianOptimization(Tuner):
    """``kerastuner.tuners.BayesianOptimization``: Gaussian-process regression + expected improvement
    over the declared search space (``oracle.BayesianOptimizationOracle``)."""

    def __init__(self, hypermodel, objective, max_trials, num_initial_points=None, seed=None,
                 hyperparameters=None, tune_new_entries=True, allow_new_entries=True, **kwargs):
        from .oracle import BayesianOptimizationOracle
        oracle = BayesianOptimizationOracle(objective=objective, max_trials=max_trials,
                                            num_initial_points=num_initial_points, seed=seed,
                                            hyperparameters=hyperparameters,
                                            tune_new_entries=tune_new_entries,
                                            allow_new_entries=allow_new_entries)
        super().__init__(oracle, hypermodel, **kwargs)
