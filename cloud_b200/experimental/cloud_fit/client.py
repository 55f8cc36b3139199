"""cloud_fit client (reference: experimental/cloud_fit/client.py:45-286).

``cloud_fit(model, remote_dir, ..., **fit_kwargs)`` serialises the model and every ``fit``
argument into ``remote_dir`` and launches a job that runs ``model.fit`` there under the requested
strategy.  Layout (same as the reference): ``<remote_dir>/training_assets`` (datasets, cloudpickled
callbacks and the other fit kwargs), ``<remote_dir>/model`` (the saved model) and, written by the
job, ``<remote_dir>/output``.  The AI Platform job submission becomes a launcher job on local GPUs.
"""
from __future__ import annotations

import datetime
import logging
import os
import types
from typing import Any, Dict, Optional, Text

import cloudpickle

from . import utils
from ...core import deploy, machine_config
from ...launch import plan as plan_mod
from ...launch import topology

logger = logging.getLogger(__name__)

MULTI_WORKER_MIRRORED_STRATEGY_NAME = utils.MULTI_WORKER_MIRRORED_STRATEGY_NAME
MIRRORED_STRATEGY_NAME = utils.MIRRORED_STRATEGY_NAME
SUPPORTED_DISTRIBUTION_STRATEGIES = utils.SUPPORTED_DISTRIBUTION_STRATEGIES


def cloud_fit(model, remote_dir: Text, region: Text = None, project_id: Text = None,
              image_uri: Text = None,
              distribution_strategy: Text = MULTI_WORKER_MIRRORED_STRATEGY_NAME,
              job_spec: Dict[str, Any] = None, job_id: Text = None, **fit_kwargs) -> Text:
    """Launches ``model.fit(**fit_kwargs)`` as a job; returns the job id."""
    if distribution_strategy not in SUPPORTED_DISTRIBUTION_STRATEGIES:
        raise ValueError("{} is not supported. Must be one of {}".format(
            distribution_strategy, list(SUPPORTED_DISTRIBUTION_STRATEGIES.keys())))
    if utils.is_tf_v1():
        raise RuntimeError("TF 1.X is not supported.")
    os.makedirs(remote_dir, exist_ok=True)
    # serialise first: a bad argument must fail before anything is launched
    _serialize_assets(remote_dir, model, **fit_kwargs)

    args = ["--remote_dir", remote_dir, "--distribution_strategy", distribution_strategy]
    plan, labels = _plan_from_job_spec(job_spec)
    job_id = job_id or _default_job_id()
    image = {"workdir": remote_dir,
             "entrypoint": ["python", "-m", "cloud_b200.experimental.cloud_fit.remote"],
             "image_uri": image_uri}
    return _submit_job(plan, image, args, job_id, labels, region)


def _default_job_id():
    return "cloud_fit_{}".format(datetime.datetime.now().strftime("%Y%m%d%H%M%S"))


def _serialize_assets(remote_dir: Text, model, **fit_kwargs) -> None:
    """Writes ``training_assets`` and ``model`` under ``remote_dir`` (client.py:138-192)."""
    assets_dir = os.path.join(remote_dir, "training_assets")
    os.makedirs(assets_dir, exist_ok=True)
    kwargs = dict(fit_kwargs)
    for key in ("x", "validation_data"):
        value = kwargs.get(key)
        if isinstance(value, types.GeneratorType) or (
                value is not None and hasattr(value, "__next__") and not hasattr(value, "__len__")):
            raise NotImplementedError("Generators are not supported for `{}`; pass arrays or a "
                                      "cloud_b200.data.Dataset.".format(key))
    callbacks = kwargs.pop("callbacks", None)
    if callbacks:
        with open(os.path.join(assets_dir, "callbacks.pkl"), "wb") as f:
            f.write(cloudpickle.dumps(callbacks))
    with open(os.path.join(assets_dir, "fit_kwargs.pkl"), "wb") as f:
        f.write(cloudpickle.dumps(kwargs))
    with open(os.path.join(assets_dir, "MANIFEST"), "w") as f:
        f.write("\n".join(sorted(list(kwargs.keys()) + (["callbacks"] if callbacks else []))))
    model.save(os.path.join(remote_dir, "model"))


def _default_job_spec(region=None, image_uri=None, entry_point_args=None) -> Dict[str, Any]:
    """The reference's default: CUSTOM tier, one n1-standard-4 master + 1 worker (client.py:195-224).
    Locally: one chief and one worker, one GPU each when at least two GPUs are visible, else the
    chief alone on one GPU (or the CPU)."""
    topo = topology.discover()
    if topo.gpu_count >= 2:
        one = machine_config.COMMON_MACHINE_CONFIGS["T4_1X"]
        return {"chief_config": one, "worker_count": 1, "worker_config": one}
    if topo.gpu_count == 1:
        return {"chief_config": machine_config.COMMON_MACHINE_CONFIGS["T4_1X"], "worker_count": 0,
                "worker_config": None}
    cpu = machine_config.COMMON_MACHINE_CONFIGS["CPU"]
    return {"chief_config": cpu, "worker_count": 1, "worker_config": cpu}


def _plan_from_job_spec(job_spec: Optional[Dict[str, Any]]):
    spec = dict(job_spec) if job_spec else _default_job_spec()
    labels = spec.pop("labels", None)
    if "trainingInput" in spec:          # a CAIP-style spec: count replicas, one device each
        ti = spec["trainingInput"]
        n_workers = int(ti.get("workerCount", 0) or 0)
        cfg = _default_job_spec()["chief_config"]
        spec = {"chief_config": cfg, "worker_count": n_workers, "worker_config": cfg}
    topo = topology.discover()
    gpu_ids = spec.get("gpu_ids")
    if gpu_ids is None:
        gpu_ids = list(range(topo.gpu_count))
    chief = spec["chief_config"]
    if topo.gpu_count == 0 and chief.accelerator_type != machine_config.AcceleratorType.NO_ACCELERATOR:
        cpu = machine_config.COMMON_MACHINE_CONFIGS["CPU"]     # GPU-less box: run the ranks on CPU
        spec = {"chief_config": cpu, "worker_count": spec.get("worker_count", 0), "worker_config": cpu}
        gpu_ids = []
    plan = plan_mod.plan_ranks(spec["chief_config"], spec.get("worker_count", 0),
                               spec.get("worker_config"), gpu_ids=gpu_ids)
    return plan, labels


def _submit_job(plan, image, args, job_id, labels, region) -> Text:
    launched = deploy.submit_plan(plan, image, args=args, job_id=job_id, labels=labels, region=region)
    logger.info("Job submitted successfully.")
    logger.info("Your job ID is: %s", launched)
    return launched
