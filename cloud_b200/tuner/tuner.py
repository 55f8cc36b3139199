"""``CloudOracle``, ``CloudTuner`` and ``DistributingCloudTuner``
(reference: src/python/tensorflow_cloud/tuner/tuner.py:32-376; DistributingCloudTuner is named by
BASELINE.json but absent from the reference snapshot - SURVEY.md section 2.1 row T5 gives its
upstream contract, which is followed here).

The oracle keeps the reference's behaviour verbatim at the API level: exactly one of
``study_config`` xor (``objective`` + ``hyperparameters``); the search space must be declared
up-front (``allow_new_entries=False``); study id ``CloudTuner_study_<id|timestamp>``;
``create_trial`` stops at ``max_trials`` or when any trial is STOPPING; ``update_trial`` reports
the objective and polls the early-stop decision; ``end_trial`` completes (or marks infeasible);
``get_best_trials`` ranks COMPLETED trials by their final measurement.  ``project_id`` / ``region``
are accepted as local namespace labels of the on-disk optimizer store.
"""
from __future__ import annotations

import datetime
import logging
import os
import time
from typing import Any, Dict, List, Mapping, Optional, Text, Union

from . import optimizer_client
from . import utils
from .engine import hypermodel as hypermodel_module
from .engine import hyperparameters as hp_module
from .engine import oracle as oracle_module
from .engine import trial as trial_module
from .engine import tuner as tuner_module

logger = logging.getLogger(__name__)


class CloudOracle(oracle_module.Oracle):
    """Oracle backed by the (local) optimizer service."""

    def __init__(self, project_id: Text, region: Text,
                 objective: Union[Text, oracle_module.Objective] = None,
                 hyperparameters: hp_module.HyperParameters = None,
                 study_config: Optional[Dict[Text, Any]] = None, max_trials: int = None,
                 study_id: Optional[Text] = None):
        if not project_id:
            raise ValueError('"project_id" is not found.')
        if not region:
            raise ValueError('"region" is not found.')
        objective, hyperparameters, self.study_config = self._search_space(
            objective, hyperparameters, study_config)
        super(CloudOracle, self).__init__(objective=objective, hyperparameters=hyperparameters,
                                          max_trials=max_trials, allow_new_entries=False,
                                          tune_new_entries=False)
        self.project_id, self.region = project_id, region
        self.objective = utils.format_objective(objective)
        self.hyperparameters = hyperparameters
        self.max_trials = max_trials
        suffix = study_id or datetime.datetime.now().strftime("%Y%m%d_%H%M%S")
        self.study_id = "CloudTuner_study_{}".format(suffix)
        self.service = optimizer_client.create_or_load_study(
            self.project_id, self.region, self.study_id, self.study_config)
        self.trials = {}
        self._start_time = None

    @staticmethod
    def _search_space(objective, hyperparameters, study_config):
        """The search space comes either from a ready StudyConfig or from (objective,
        hyperparameters) - never from both; returns (objective, hyperparameters, study_config)."""
        if study_config:
            if objective or hyperparameters:
                raise ValueError('Please configure either study_config or '
                                 '"objective, and hyperparameters".')
            return (utils.convert_study_config_to_objective(study_config),
                    utils.convert_study_config_to_hps(study_config), study_config)
        if not (objective and hyperparameters):
            raise ValueError("If study_config is not set, objective and hyperparameters must be set.")
        return objective, hyperparameters, utils.make_study_config(objective, hyperparameters)

    # -- service trial <-> engine trial ---------------------------------------------------------
    def _as_trial(self, optimizer_trial, status) -> trial_module.Trial:
        """An engine trial for a trial of the optimizer service (same id, suggested values)."""
        hps = utils.convert_optimizer_trial_to_hps(self.hyperparameters.copy(), optimizer_trial)
        trial = trial_module.Trial(hyperparameters=hps, trial_id=utils.get_trial_id(optimizer_trial),
                                   status=status)
        final = optimizer_trial.get("finalMeasurement")
        if final:
            trial.best_step = int(final.get("stepCount", 1))
            trial.score = final["metrics"][0]["value"]
        return trial

    def _no_more_trials(self) -> trial_module.Trial:
        hps = self.hyperparameters.copy()
        hps.values = None
        return trial_module.Trial(hyperparameters=hps, trial_id="n", status=trial_module.TrialStatus.STOPPED)

    def create_trial(self, tuner_id: Text) -> trial_module.Trial:
        """Next trial for ``tuner_id``: none once the study holds ``max_trials`` trials or the
        service is stopping one; otherwise the service's suggestion for this client."""
        known = self.service.list_trials()
        exhausted = bool(self.max_trials) and len(known) >= self.max_trials
        if exhausted or any(t["state"] == "STOPPING" for t in known):
            return self._no_more_trials()
        suggestions = self.service.get_suggestions(tuner_id)
        if "trials" not in suggestions:
            return trial_module.Trial(hyperparameters={}, status=trial_module.TrialStatus.STOPPED)
        trial = self._as_trial(suggestions["trials"][0], trial_module.TrialStatus.RUNNING)
        logger.info("Hyperparameters requested by tuner (%s): %s ", tuner_id, trial.hyperparameters.values)
        self._start_time = time.time()
        self.trials[trial.trial_id] = trial
        self.ongoing_trials[tuner_id] = trial
        self._save_trial(trial)
        self.save()
        return trial

    def update_trial(self, trial_id: Text, metrics: Mapping[Text, Union[int, float]], step: int = 0):
        """Reports the objective values of ``metrics`` at ``step`` and asks the service whether
        the trial should stop early; returns the trial's (possibly STOPPED) status."""
        elapsed_secs = time.time() - self._start_time
        if elapsed_secs < 0 or step < 0:
            raise ValueError("Both elapsed_secs and step must be non-negative.")
        if elapsed_secs == 0 and step == 0:
            raise ValueError("At least one of {elapsed_secs, step} must be positive")
        wanted = [ob.name for ob in self.objective]
        for name in wanted:
            if name not in metrics:
                logger.info('Objective "%s" is not found in metrics.', name)
        reported = [{"metric": n, "value": float(metrics[n])} for n in wanted if n in metrics]
        self.service.report_intermediate_objective_value(step, elapsed_secs, reported, trial_id)
        trial = self.trials[trial_id]
        logger.info("UpdateTrial: polls the stop decision.")
        if self.service.should_trial_stop(trial_id):
            trial.status = trial_module.TrialStatus.STOPPED
        return trial.status

    # status -> (trial_infeasible, infeasibility_reason) for the service's complete call
    _COMPLETION = {trial_module.TrialStatus.COMPLETED: (False, None),
                   trial_module.TrialStatus.INVALID: (True, trial_module.TrialStatus.INVALID)}

    def end_trial(self, trial_id: Text, status: Text = "COMPLETED"):
        """Completes the ongoing trial ``trial_id`` in the service (as infeasible for INVALID) and
        records its final score / best step."""
        owner = next((tid for tid, t in self.ongoing_trials.items() if t.trial_id == trial_id), None)
        if owner is None:
            raise ValueError("Ongoing trial with id: {} not found.".format(trial_id))
        logger.info("End trial requested by tuner (%s)", owner)
        trial = self.ongoing_trials.pop(owner)
        trial.status = status
        if status not in self._COMPLETION:
            raise ValueError('Unexpected status passed. Expected "COMPLETED" or '
                             '"INVALID", found {}'.format(status))
        infeasible, reason = self._COMPLETION[status]
        finished = self.service.complete_trial(trial_id, infeasible, reason)
        if not infeasible:
            final = finished["finalMeasurement"]
            trial.best_step = int(final.get("stepCount", 1))
            trial.score = final["metrics"][0]["value"]
        self._save_trial(trial)
        self.save()

    def get_best_trials(self, num_trials: int = 1) -> List[trial_module.Trial]:
        """COMPLETED trials of the study ranked by their final measurement (single objective)."""
        if len(self.objective) > 1:
            raise ValueError("Getting the best trials for multi-objective optimization "
                             "is not supported. ")
        finished = [t for t in self.service.list_trials()
                    if t["state"] == "COMPLETED" and "finalMeasurement" in t]
        finished.sort(key=lambda t: t["finalMeasurement"]["metrics"][0]["value"],
                      reverse=utils.format_goal(self.objective[0].direction) == "MAXIMIZE")
        return [self._as_trial(t, trial_module.TrialStatus.COMPLETED) for t in finished[:num_trials]]


class CloudTuner(tuner_module.Tuner):
    """Tuner whose oracle is a ``CloudOracle`` (tuner.py:325-376)."""

    def __init__(self, hypermodel: Union[hypermodel_module.HyperModel, Any], project_id: Text,
                 region: Text, objective: Union[Text, oracle_module.Objective] = None,
                 hyperparameters: hp_module.HyperParameters = None,
                 study_config: Optional[Dict[Text, Any]] = None, max_trials: int = None,
                 study_id: Optional[Text] = None, **kwargs):
        oracle = CloudOracle(project_id=project_id, region=region, objective=objective,
                             hyperparameters=hyperparameters, study_config=study_config,
                             max_trials=max_trials, study_id=study_id)
        super(CloudTuner, self).__init__(oracle=oracle, hypermodel=hypermodel, **kwargs)


class DistributingCloudTuner(CloudTuner):
    """``CloudTuner`` whose every trial is trained by a separately launched job
    (``cloud_fit``) on ``replica_count`` rank groups of ``replica_config`` GPUs each.

    ``run_trial`` builds the model for the trial's hyperparameters, adds TensorBoard and
    ModelCheckpoint callbacks rooted at ``<directory>/<trial_id>/{logs,checkpoint}``, ships model +
    data through ``cloud_fit`` into ``<directory>/<trial_id>``, blocks until the job ends, replays
    the per-epoch metrics from the job's training log into ``oracle.update_trial`` and fails the
    trial if the job failed.  ``load_model`` restores from the trial's output directory.
    ``directory`` must be storage every rank can reach (any local path on the single box);
    ``container_uri`` is accepted and ignored.
    """

    def __init__(self, hypermodel, project_id: Text, region: Text, directory: Text,
                 objective=None, hyperparameters=None, study_config=None, max_trials: int = None,
                 study_id: Optional[Text] = None, container_uri: Optional[Text] = None,
                 replica_config="auto", replica_count: int = 1, gpu_ids=None, **kwargs):
        self._study_id = study_id
        super(DistributingCloudTuner, self).__init__(
            hypermodel=hypermodel, project_id=project_id, region=region, objective=objective,
            hyperparameters=hyperparameters, study_config=study_config, max_trials=max_trials,
            study_id=study_id, directory=directory, **kwargs)
        from ..core import machine_config
        if replica_config == "auto":
            replica_config = machine_config.COMMON_MACHINE_CONFIGS["T4_1X"]
        self.container_uri = container_uri
        self.replica_config = replica_config
        self.replica_count = int(replica_count)
        self.gpu_ids = gpu_ids
        self.directory = str(directory)

    def _trial_dir(self, trial_id):
        return os.path.join(self.directory, str(trial_id))

    def run_trial(self, trial, *fit_args, **fit_kwargs):
        from ..experimental.cloud_fit import client as cloud_fit_client
        from ..keras import callbacks as cb
        from ..launch import registry
        remote_dir = self._trial_dir(trial.trial_id)
        callbacks = list(fit_kwargs.pop("callbacks", []))
        callbacks.append(cb.TensorBoard(log_dir=os.path.join(remote_dir, "logs")))
        callbacks.append(cb.ModelCheckpoint(os.path.join(remote_dir, "checkpoint", "cp"),
                                            save_weights_only=True))
        fit_kwargs["callbacks"] = callbacks
        model = self._build_model(trial.hyperparameters)
        if fit_args:
            names = ["x", "y"]
            for name, value in zip(names, fit_args):
                fit_kwargs[name] = value
        job_id = "{}_{}".format(self.oracle.study_id, trial.trial_id)
        job_spec = {"chief_config": self.replica_config,
                    "worker_count": max(self.replica_count - 1, 0),
                    "worker_config": self.replica_config, "gpu_ids": self.gpu_ids}
        launched = cloud_fit_client.cloud_fit(
            model, remote_dir=remote_dir, region=self.oracle.region,
            project_id=self.oracle.project_id, image_uri=self.container_uri,
            distribution_strategy="MultiWorkerMirroredStrategy" if self.replica_count > 1 or
            self.replica_config.accelerator_count > 1 else "MirroredStrategy",
            job_spec=job_spec, job_id=job_id, **fit_kwargs)
        status = registry.wait(launched)
        if status.get("state") != "SUCCEEDED":
            raise RuntimeError("AIP Training job failed: trial {} job {} ended in state {}: {}".format(
                trial.trial_id, launched, status.get("state"), status.get("error_tail", "")[-500:]))
        # replay per-epoch metrics written by the job (TensorBoard JSONL) into the oracle
        self._report_from_logs(trial, remote_dir)

    def _report_from_logs(self, trial, remote_dir):
        import json
        merged = {}
        for split, prefix in (("train", ""), ("validation", "val_")):
            path = os.path.join(remote_dir, "logs", split, "scalars.jsonl")
            if not os.path.exists(path):
                continue
            with open(path) as f:
                for line in f:
                    rec = json.loads(line)
                    step = int(rec.pop("step"))
                    rec.pop("wall_time", None)
                    merged.setdefault(step, {}).update({prefix + k: v for k, v in rec.items()})
        for step in sorted(merged):
            status = self.oracle.update_trial(trial.trial_id, metrics=merged[step], step=step)
            trial.status = status

    def load_model(self, trial):
        from ..keras import models
        return models.load_model(os.path.join(self._trial_dir(trial.trial_id), "output"))
