"""Client of the optimizer service for one study
(reference: src/python/tensorflow_cloud/tuner/optimizer_client.py:33-443).

Same class, methods and failure semantics as the reference's ``_OptimizerClient`` - suggestions
with 429 -> ``{}``, long-running-operation polling with bounded exponential back-off,
``SuggestionInactiveError``, early-stop polling that stops the trial, ``complete_trial``,
``list_trials`` / ``list_studies`` / ``delete_study`` (404 -> ``ValueError``) and
``create_or_load_study`` with its 409 -> load (x ``NUM_TRIES_FOR_STUDIES``) race handling - but the
``service_client`` is the local file-backed service (``local_vizier.LocalVizier``) instead of a
googleapiclient discovery resource.
"""
from __future__ import annotations

import logging
import time
from typing import Any, Dict, List, Mapping, Optional, Text, Union

from . import constants
from . import local_vizier
from .local_vizier import HttpError

logger = logging.getLogger(__name__)

_MAX_POLLING_ATTEMPTS = 30


class SuggestionInactiveError(Exception):
    """Indicates that a suggestion was requested from an inactive study."""


class _OptimizerClient(object):
    """One instance talks to one study."""

    def __init__(self, service_client, project_id: Text, region: Text, study_id: Text = None):
        self.service_client = service_client
        self.project_id = project_id
        self.region = region
        if not study_id:
            raise ValueError("Use create_or_load_study() instead of constructing the"
                             "OptimizerClient class directly")
        self.study_id = study_id

    # -- suggestions ---------------------------------------------------------------------------------
    def get_suggestions(self, client_id: Text):
        """A dict with a ``"trials"`` list, or ``{}`` when the study cannot give more trials."""
        try:
            resp = self.service_client.suggest_trials(
                self._make_study_name(),
                {"client_id": client_id,
                 "suggestion_count": constants.SUGGESTION_COUNT_PER_REQUEST})
        except HttpError as e:
            if e.resp.status == 429:
                logger.info("Reached max number of trials.")
                return {}
            logger.info("SuggestTrial failed.")
            raise e
        logger.info("CreateTrial: polls the suggestions.")
        operation = self._obtain_long_running_operation(resp)
        suggestions = operation["response"]
        if "trials" not in suggestions:
            if operation["response"].get("studyState") == "INACTIVE":
                raise SuggestionInactiveError("The study is stopped due to an internal error.")
        return suggestions

    # -- measurements ---------------------------------------------------------------------------------
    def report_intermediate_objective_value(self, step: int, elapsed_secs: float,
                                            metric_list: List[Mapping[Text, Union[int, float]]],
                                            trial_id: Text) -> None:
        measurement = {"stepCount": step, "elapsedTime": {"seconds": int(elapsed_secs)},
                       "metrics": metric_list}
        try:
            self.service_client.add_measurement(self._make_trial_name(trial_id),
                                                {"measurement": measurement})
        except HttpError as e:
            logger.info("AddMeasurement failed.")
            raise e

    def should_trial_stop(self, trial_id: Text) -> bool:
        """Polls the early-stopping decision; a positive answer also stops the trial."""
        trial_name = self._make_trial_name(trial_id)
        try:
            resp = self.service_client.check_early_stopping_state(trial_name)
        except HttpError as e:
            logger.info("CheckEarlyStoppingState failed.")
            raise e
        operation = self._obtain_long_running_operation(resp)
        if operation["response"].get("shouldStop"):
            try:
                self.service_client.stop_trial(trial_name)
            except HttpError as e:
                logger.info("StopTrial failed.")
                raise e
            return True
        return False

    def complete_trial(self, trial_id: Text, trial_infeasible: bool,
                       infeasibility_reason: Text = None) -> Dict[Text, Any]:
        try:
            return self.service_client.complete_trial(
                self._make_trial_name(trial_id),
                {"trial_infeasible": trial_infeasible, "infeasible_reason": infeasibility_reason})
        except HttpError as e:
            logger.info("CompleteTrial failed.")
            raise e

    # -- listing / deleting -----------------------------------------------------------------------------
    def list_trials(self) -> List[Mapping[Text, Any]]:
        try:
            resp = self.service_client.list_trials(self._make_study_name())
        except HttpError as e:
            logger.info("ListTrials failed.")
            raise e
        return resp.get("trials", [])

    def list_studies(self) -> List[Dict[Text, Any]]:
        parent = self._make_study_parent()
        try:
            resp = self.service_client.list_studies(parent)
        except HttpError:
            logger.info("List studies failed.")
            raise
        return resp.get("studies", [])

    def delete_study(self, study_name: Text = None) -> None:
        if study_name is None:
            study_name = self._make_study_name()
        try:
            self.service_client.delete_study(study_name)
        except HttpError as e:
            if e.resp.status == 404:
                raise ValueError("DeleteStudy failed. Study not found: {}.".format(study_name))
            logger.info("DeleteStudy failed.")
            raise
        logger.info("Study deleted: {}.".format(study_name))

    # -- long-running operations ------------------------------------------------------------------------
    def _obtain_long_running_operation(self, resp):
        """Polls ``operations.get`` until ``done``: back-off ``max(1, 0.3) * 1.41^min(n, 9)``
        seconds upstream (optimizer_client.py:312-348); the local service answers on the first
        poll, so the same schedule is kept but scaled down by ``TFC_LRO_SCALE`` (default 0.001)."""
        op_id = resp["name"]
        if "done" in resp and resp["done"]:
            return resp
        import os
        scale = float(os.environ.get("TFC_LRO_SCALE", "0.001"))
        num_attempts = 0
        polling_secs = 0.3
        while num_attempts < _MAX_POLLING_ATTEMPTS:
            try:
                resp = self.service_client.get_operation(op_id, self._make_study_name())
            except HttpError as e:
                logger.info("GetLongRunningOperations failed.")
                raise e
            if "done" in resp and resp["done"]:
                return resp
            polling_secs = max(polling_secs, 1.0) * 1.41 ** min(num_attempts, 9)
            time.sleep(polling_secs * scale)
            num_attempts += 1
        raise RuntimeError("GetLongRunningOperations timeout.")

    # -- names ------------------------------------------------------------------------------------------
    def _make_study_name(self):
        return "projects/{}/locations/{}/studies/{}".format(self.project_id, self.region, self.study_id)

    def _make_trial_name(self, trial_id):
        return "projects/{}/locations/{}/studies/{}/trials/{}".format(
            self.project_id, self.region, self.study_id, trial_id)

    def _make_study_parent(self):
        return "projects/{}/locations/{}".format(self.project_id, self.region)


def create_or_load_study(project_id: Text, region: Text, study_id: Text,
                         study_config: Optional[Dict[Text, Any]] = None,
                         service_client=None) -> _OptimizerClient:
    """Creates the study, or loads it if another worker created it first.

    Designed for many tuner workers racing on one study id (optimizer_client.py:378-381): a 409 on
    create turns into up to ``NUM_TRIES_FOR_STUDIES`` ``get`` attempts one second apart."""
    service_client = service_client if service_client is not None else local_vizier.LocalVizier()
    study_parent = "projects/{}/locations/{}".format(project_id, region)
    request = {"study_config": study_config}
    try:
        logger.info("Creating a new study: %s/studies/%s", study_parent, study_id)
        service_client.create_study(study_parent, study_id, request)
    except HttpError as create_error:
        if create_error.resp.status != 409:
            raise
        # the study exists: load it
        study_name = "{}/studies/{}".format(study_parent, study_id)
        import os
        num_tries = 0
        while True:
            try:
                service_client.get_study(study_name)
                break
            except HttpError as err:
                num_tries += 1
                if num_tries >= constants.NUM_TRIES_FOR_STUDIES:
                    raise RuntimeError("GetStudy wasn't successful after {0} tries: {1!s}".format(
                        constants.NUM_TRIES_FOR_STUDIES, err))
                time.sleep(float(os.environ.get("TFC_STUDY_RETRY_SECS", "1")))
    return _OptimizerClient(service_client, project_id, region, study_id)
