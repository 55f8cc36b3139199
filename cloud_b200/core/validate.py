"""Input validation for ``run`` (reference: src/python/tensorflow_cloud/core/validate.py:23-218).

Error-message prefixes are part of the contract (the reference's tests match them by regex,
core/tests/unit/validate_test.py:116-361) and are kept verbatim.  Differences: ``docker_base_image``
is accepted but unused (no containers on the local box), and a TPU worker is accepted by
validation exactly like in the reference - the launcher is what reports "no TPU on this box".
"""
from __future__ import annotations

import os

from . import machine_config
from . import platform as gcp
from ..utils import framework_utils as tf_utils


def validate(entry_point, requirements_txt, distribution_strategy, chief_config, worker_config,
             worker_count, region, args, stream_logs, docker_image_bucket_name,
             called_from_notebook, job_labels=None, docker_base_image=None):
    _validate_files(entry_point, requirements_txt)
    _validate_distribution_strategy(distribution_strategy)
    _validate_cluster_config(chief_config, worker_count, worker_config, docker_base_image)
    _validate_job_labels(job_labels or {})
    _validate_other_args(region, args, stream_logs, docker_image_bucket_name, called_from_notebook)


def _validate_files(entry_point, requirements_txt):
    cwd = os.getcwd()
    if entry_point is not None and not os.path.isfile(os.path.join(cwd, entry_point)):
        raise ValueError("Invalid `entry_point`. Expected a relative path in the current "
                         "directory tree. Received: {}".format(entry_point))
    if requirements_txt is not None and not os.path.isfile(os.path.join(cwd, requirements_txt)):
        raise ValueError("Invalid `requirements_txt`. Expected a relative path in the current "
                         "directory tree. Received: {}".format(requirements_txt))
    if entry_point is not None and not (entry_point.endswith("py") or entry_point.endswith("ipynb")):
        raise ValueError("Invalid `entry_point`. Expected a python file or an iPython notebook. "
                         "Received: {}".format(entry_point))


def _validate_distribution_strategy(distribution_strategy):
    if distribution_strategy not in ["auto", None]:
        raise ValueError('Invalid `distribution_strategy` input. Expected "auto" or None. '
                         "Received {}.".format(distribution_strategy))


def _validate_cluster_config(chief_config, worker_count, worker_config, docker_base_image):
    if not isinstance(chief_config, machine_config.MachineConfig):
        raise ValueError('Invalid `chief_config` input. Expected "auto" or `MachineConfig` '
                         "instance. Received {}.".format(chief_config))
    if worker_count < 0:
        raise ValueError("Invalid `worker_count` input. Expected a postive integer value. "
                         "Received {}.".format(worker_count))
    if worker_count > 0 and not isinstance(worker_config, machine_config.MachineConfig):
        raise ValueError('Invalid `worker_config` input. Expected "auto" or `MachineConfig` '
                         "instance. Received {}.".format(worker_config))
    if machine_config.is_tpu_config(chief_config):
        raise ValueError("Invalid `chief_config` input. `chief_config` cannot be a TPU config. "
                         "Received {}.".format(chief_config))
    if machine_config.is_tpu_config(worker_config):
        if worker_count != 1:
            raise ValueError("Invalid `worker_count` input. Expected worker_count=1 for TPU "
                             "`worker_config`. Received {}.".format(worker_count))
        elif docker_base_image is None:
            version = tf_utils.get_version()
            if version is not None and version not in gcp.get_cloud_tpu_supported_tf_versions():
                raise NotImplementedError("TPUs are only supported for TF version <= 2.1.0")


def _validate_job_labels(job_labels):
    gcp.validate_job_labels(job_labels)


def _validate_other_args(region, args, stream_logs, docker_image_bucket_name, called_from_notebook):
    if not isinstance(region, str):
        raise ValueError("Invalid `region` input. Expected None or a string value. "
                         "Received {}.".format(str(region)))
    if args is not None and not isinstance(args, list):
        raise ValueError("Invalid `entry_point_args` input. Expected None or a list. "
                         "Received {}.".format(str(args)))
    if not isinstance(stream_logs, bool):
        raise ValueError("Invalid `stream_logs` input. Expected a boolean. "
                         "Received {}.".format(str(stream_logs)))
    if called_from_notebook and docker_image_bucket_name is None:
        raise ValueError("Invalid `docker_image_bucket_name` input. When `run` API is used within "
                         "a python notebook, `docker_image_bucket_name` is expected to be specifed. "
                         "It names the staging area of the job (a directory under the job root on "
                         "this box). Received {}.".format(str(docker_image_bucket_name)))
