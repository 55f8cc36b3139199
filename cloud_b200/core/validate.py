"""Input validation for ``run`` (reference behaviour: src/python/tensorflow_cloud/core/validate.py:23-218).

The beginnings of the error messages ("Invalid `entry_point`", "Invalid `worker_count` input", ...)
are part of the contract - the reference's tests match them by regex
(core/tests/unit/validate_test.py:116-361) - so they are kept.  Differences: ``docker_base_image`` is
accepted but unused (no containers on the local box), and a TPU worker passes validation exactly
like upstream - the launcher is what reports "no TPU on this box".

Organisation: every rule is a row ``(argument, broken?, what was expected)`` evaluated in order by
``_first_violation``; the two rules that are not plain value checks (TPU framework version, job
labels) follow the table.
"""
from __future__ import annotations

import os

from . import machine_config
from . import platform as gcp
from ..utils import framework_utils as tf_utils


def _is_machine(cfg) -> bool:
    return isinstance(cfg, machine_config.MachineConfig)


def _missing_file(path) -> bool:
    return path is not None and not os.path.isfile(os.path.join(os.getcwd(), path))


def _complain(argument, expected, received, with_input=True):
    head = "Invalid `{}`{}. ".format(argument, " input" if with_input else "")
    return ValueError("{}{} Received{} {}{}".format(head, expected, "" if with_input else ":", received,
                                                    "." if with_input else ""))


def _first_violation(rules):
    for argument, broken, expected, received, with_input in rules:
        if broken:
            raise _complain(argument, expected, received, with_input)


def validate(entry_point, requirements_txt, distribution_strategy, chief_config, worker_config,
             worker_count, region, args, stream_logs, docker_image_bucket_name,
             called_from_notebook, job_labels=None, docker_base_image=None):
    """Raises ``ValueError`` (``NotImplementedError`` for an unsupported TPU set-up) on the first
    argument of ``run`` that cannot be used."""
    relative = "Expected a relative path in the current directory tree."
    machine = 'Expected "auto" or `MachineConfig` instance.'
    tpu_worker = machine_config.is_tpu_config(worker_config)
    _first_violation([
        # files
        ("entry_point", _missing_file(entry_point), relative, entry_point, False),
        ("requirements_txt", _missing_file(requirements_txt), relative, requirements_txt, False),
        ("entry_point", entry_point is not None and not entry_point.endswith(("py", "ipynb")),
         "Expected a python file or an iPython notebook.", entry_point, False),
        # strategy
        ("distribution_strategy", distribution_strategy not in ("auto", None),
         'Expected "auto" or None.', distribution_strategy, True),
        # cluster
        ("chief_config", not _is_machine(chief_config), machine, chief_config, True),
        ("worker_count", worker_count < 0, "Expected a postive integer value.", worker_count, True),
        ("worker_config", worker_count > 0 and not _is_machine(worker_config), machine, worker_config, True),
        ("chief_config", _is_machine(chief_config) and machine_config.is_tpu_config(chief_config),
         "`chief_config` cannot be a TPU config.", chief_config, True),
        ("worker_count", tpu_worker and worker_count != 1,
         "Expected worker_count=1 for TPU `worker_config`.", worker_count, True),
    ])
    if tpu_worker and docker_base_image is None:
        version = tf_utils.get_version()
        if version is not None and version not in gcp.get_cloud_tpu_supported_tf_versions():
            raise NotImplementedError("TPUs are only supported for TF version <= 2.1.0")
    gcp.validate_job_labels(job_labels or {})
    _first_violation([
        ("region", not isinstance(region, str), "Expected None or a string value.", str(region), True),
        ("entry_point_args", args is not None and not isinstance(args, list), "Expected None or a list.",
         str(args), True),
        ("stream_logs", not isinstance(stream_logs, bool), "Expected a boolean.", str(stream_logs), True),
        ("docker_image_bucket_name", bool(called_from_notebook) and docker_image_bucket_name is None,
         "When `run` API is used within a python notebook, `docker_image_bucket_name` is expected to be "
         "specifed. It names the staging area of the job (a directory under the job root on this box).",
         str(docker_image_bucket_name), True),
    ])
