"""``run`` - train your script on the B200 box with one call
(reference: src/python/tensorflow_cloud/core/run.py:31-263).

Same signature, defaults and control flow as the reference: no-op when already running inside a
launched job, unknown kwargs -> ``TypeError``, 'auto' machine configs -> ``T4_1X`` (one GPU),
validation, strategy-injecting preprocessing iff ``distribution_strategy == 'auto'`` or the entry
point is a notebook or ``None``, staging (the containerize replacement), deployment (the local
launcher) and ``sys.exit(0)`` so the calling script stops after submission (unless in a notebook).
"""
from __future__ import annotations

import os
import sys

from . import containerize
from . import deploy
from . import machine_config
from . import platform as gcp
from . import preprocess
from . import validate


def remote():
    """True when the code runs inside a job launched by ``run`` (run.py:31-33)."""
    return bool(os.environ.get("TF_KERAS_RUNNING_REMOTELY"))


def run(entry_point=None, requirements_txt=None, distribution_strategy="auto",
        docker_base_image=None, chief_config="auto", worker_config="auto", worker_count=0,
        entry_point_args=None, stream_logs=False, docker_image_bucket_name=None, job_labels=None,
        **kwargs):
    """Runs your training code on this box's B200 GPUs (see module docstring).

    ``chief_config`` / ``worker_config`` / ``worker_count`` map to GPU counts on the local box:
    the chief takes ``chief_config.accelerator_count`` GPUs, each worker
    ``worker_config.accelerator_count`` more.  ``docker_base_image`` is accepted and ignored;
    ``docker_image_bucket_name`` names the staging area (required from notebooks, as upstream).
    """
    if remote():
        return                                             # run.py:134-135
    if kwargs:
        raise TypeError("Unknown keyword arguments: %s" % (kwargs.keys(),))   # run.py:137-145
    job_labels = {} if job_labels is None else job_labels

    if chief_config == "auto":
        chief_config = machine_config.COMMON_MACHINE_CONFIGS["T4_1X"]
    if worker_config == "auto":
        worker_config = machine_config.COMMON_MACHINE_CONFIGS["T4_1X"]
    region = gcp.get_region()
    destination_dir = "/app/"
    if not isinstance(worker_count, int):
        worker_count = int(worker_count)
    called_from_notebook = _called_from_notebook()

    validate.validate(entry_point, requirements_txt, distribution_strategy, chief_config,
                      worker_config, worker_count, region, entry_point_args, stream_logs,
                      docker_image_bucket_name, called_from_notebook, job_labels=job_labels,
                      docker_base_image=docker_base_image)

    project_id = gcp.get_project_name()
    docker_registry = "local/{}".format(project_id)

    # Run preprocessing if (a) the strategy is to be injected, (b) the entry point is a notebook,
    # or (c) the calling script itself is the entry point.  (The reference evaluates
    # `entry_point.endswith` before the None check, run.py:188-192; fixed here.)
    preprocessed_entry_point = None
    if (distribution_strategy == "auto" or entry_point is None
            or entry_point.endswith("ipynb")):
        preprocessed_entry_point = preprocess.get_preprocessed_entry_point(
            entry_point, chief_config, worker_config, worker_count, distribution_strategy,
            called_from_notebook=called_from_notebook)

    cb_args = (entry_point, preprocessed_entry_point, chief_config, worker_config, docker_registry,
               project_id)
    cb_kwargs = {"requirements_txt": requirements_txt, "destination_dir": destination_dir,
                 "docker_base_image": docker_base_image,
                 "docker_image_bucket_name": docker_image_bucket_name,
                 "called_from_notebook": called_from_notebook}
    if docker_image_bucket_name is None:
        container_builder = containerize.LocalContainerBuilder(*cb_args, **cb_kwargs)
    else:
        container_builder = containerize.CloudContainerBuilder(*cb_args, **cb_kwargs)
    docker_img_uri = container_builder.get_docker_image()

    for f in container_builder.get_generated_files():      # run.py:226-229
        if f is not None and os.path.exists(f):
            os.remove(f)
    if preprocessed_entry_point is not None and os.path.exists(preprocessed_entry_point):
        os.remove(preprocessed_entry_point)

    job_id = deploy.deploy_job(region, docker_img_uri, chief_config, worker_count, worker_config,
                               entry_point_args, stream_logs, job_labels=job_labels)

    # Stop the calling script so it does not also train locally (run.py:243-246).
    if not called_from_notebook:
        sys.exit(0)
    return job_id


def _called_from_notebook():
    """Detects a Jupyter / Colab kernel (run.py:249-263), or the headless notebook runner
    (``cloud_b200.utils.notebook``), which names the running notebook in ``TFC_NOTEBOOK_PATH``."""
    if os.environ.get("TFC_NOTEBOOK_PATH"):
        return True
    try:
        import IPython  # noqa: F401
    except ImportError:
        return False
    try:
        shell = IPython.get_ipython().__class__.__name__
        return "Shell" in shell and shell != "TerminalInteractiveShell"
    except Exception:
        return False
