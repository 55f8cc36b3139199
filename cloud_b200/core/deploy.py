"""Job deployment - the local replacement for "submit to AI Platform Training"
(reference: src/python/tensorflow_cloud/core/deploy.py:28-218).

``deploy_job`` keeps the reference's signature and printed messages (deploy.py:170-184, pinned by
core/tests/unit/deploy_test.py:136-144) but the "request" is a rank plan over the GPUs of this box
and the "service" is a detached supervisor process (``cloud_b200.launch.supervisor``).  Like the
remote service, submission returns immediately; ``enable_stream_logs=True`` tails the ranks' logs
until the job ends (the ``gcloud ai-platform jobs stream-logs`` replacement, deploy.py:187-211).
"""
from __future__ import annotations

import os
import subprocess
import sys
import time
import uuid

from . import containerize
from . import platform as gcp
from ..launch import plan as plan_mod
from ..launch import registry, topology
from ..utils import api_client


def deploy_job(region, image_uri, chief_config, worker_count, worker_config, entry_point_args,
               enable_stream_logs, job_labels=None, gpu_ids=None, env=None, wait=False):
    """Launches the staged image on the local GPUs.  Returns the job id."""
    job_id = _generate_job_id()
    project_id = gcp.get_project_name()
    request_dict = _create_request_dict(job_id, region, image_uri, chief_config, worker_count,
                                        worker_config, entry_point_args, job_labels=job_labels or {},
                                        gpu_ids=gpu_ids, env=env)
    try:
        registry.create(job_id, request_dict)
        _start_supervisor(job_id)
        _print_logs_info(job_id, project_id)
        if enable_stream_logs:
            _stream_logs(job_id)
        elif wait:
            registry.wait(job_id)
    except OSError as err:
        print("There was an error submitting the job.")
        raise err
    return job_id


def _create_request_dict(job_id, region, image_uri, chief_config, worker_count, worker_config,
                         entry_point_args, job_labels, gpu_ids=None, env=None):
    """The job spec stored in the registry.  ``trainingInput`` mirrors the CAIP request body
    (deploy.py:98-167) so tools written against it keep working; ``plan`` is what actually runs."""
    visible = gpu_ids if gpu_ids is not None else list(range(topology.gpu_count()))
    plan = plan_mod.plan_ranks(chief_config, worker_count, worker_config, gpu_ids=visible)
    training_input = {
        "region": region,
        "scaleTier": "custom",
        "masterType": gcp.get_machine_type(chief_config.cpu_cores, chief_config.memory,
                                           chief_config.accelerator_type),
        "masterConfig": {
            "imageUri": image_uri,
            "acceleratorConfig": {
                "count": str(chief_config.accelerator_count),
                "type": gcp.get_accelerator_type(chief_config.accelerator_type.value)},
        },
        "workerCount": str(worker_count),
        "use_chief_in_tf_config": True,
    }
    if worker_count > 0:
        training_input["workerType"] = gcp.get_machine_type(
            worker_config.cpu_cores, worker_config.memory, worker_config.accelerator_type)
        training_input["workerConfig"] = {
            "imageUri": image_uri,
            "acceleratorConfig": {
                "count": str(worker_config.accelerator_count),
                "type": gcp.get_accelerator_type(worker_config.accelerator_type.value)},
        }
    if entry_point_args is not None:
        training_input["args"] = entry_point_args
    request = {
        "jobId": job_id,
        "trainingInput": training_input,
        "plan": plan,
        "image": containerize.resolve_image(image_uri),
        "args": list(entry_point_args or []),
        "python": sys.executable,
        "env": dict(env or {}),
        "client": api_client.user_agent(),
        "submitted": time.time(),
    }
    if job_labels:
        request["labels"] = job_labels
    return request


def submit_plan(plan, image, args=None, job_id=None, labels=None, env=None, region=None):
    """Lower-level entry used by cloud_fit and the trial scheduler: launch an explicit rank plan
    with an explicit image manifest ``{"workdir", "entrypoint"}``.  Returns the job id."""
    job_id = job_id or _generate_job_id()
    request = {"jobId": job_id, "trainingInput": {"region": region or gcp.get_region(),
                                                   "scaleTier": "custom", "args": list(args or [])},
               "plan": plan, "image": image, "args": list(args or []), "python": sys.executable,
               "env": dict(env or {}), "client": api_client.user_agent(), "submitted": time.time()}
    if labels:
        request["labels"] = labels
    registry.create(job_id, request)
    _start_supervisor(job_id)
    return job_id


def _start_supervisor(job_id):
    jdir = registry.job_dir(job_id)
    log = open(os.path.join(jdir, "logs", "supervisor.log"), "ab", buffering=0)
    env = dict(os.environ)
    repo_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    env["PYTHONPATH"] = os.pathsep.join([repo_root, env.get("PYTHONPATH", "")])
    proc = subprocess.Popen([sys.executable, "-m", "cloud_b200.launch.supervisor", job_id], env=env,
                            stdout=log, stderr=subprocess.STDOUT, start_new_session=True, cwd=jdir)
    # known from the first moment: a job can be cancelled while its supervisor is still starting
    registry.set_status(job_id, "QUEUED", supervisor_pid=proc.pid)


def _print_logs_info(job_id, project_id):
    print("Job submitted successfully.")
    print("Your job ID is: ", job_id)
    print("Please access your job logs at the following URL:")
    print("file://{}".format(os.path.join(registry.job_dir(job_id), "logs")))


def _stream_logs(job_id):
    """Tails every rank's log to stdout (prefixed) until the job reaches a terminal state."""
    try:
        print("Streaming job logs: ")
        offsets = {}
        while True:
            state = registry.status(job_id).get("state")
            for path in registry.log_paths(job_id):
                name = os.path.basename(path)
                if name == "supervisor.log":
                    continue
                with open(path, "rb") as f:
                    f.seek(offsets.get(path, 0))
                    chunk = f.read()
                    offsets[path] = f.tell()
                for line in chunk.decode(errors="replace").replace("\x08", "").splitlines():
                    print("[{}] {}".format(name[:-4], line))
            if state in registry.TERMINAL:
                break
            time.sleep(0.2)
        st = registry.status(job_id)
        if st.get("state") == "FAILED":
            print("Job failed (rank {}, exit code {}).".format(st.get("failed_rank"), st.get("exit_code")))
    except (ValueError, OSError) as err:
        print("There was an error streaming the job logs.")
        raise err


def _generate_job_id():
    """Unique job id prefixed with 'tf_cloud_train' (deploy.py:214-218)."""
    unique_tag = str(uuid.uuid4()).replace("-", "_")
    return "tf_cloud_train_{}".format(unique_tag)


def cancel_job(job_id):
    st = registry.status(job_id)
    pid = st.get("supervisor_pid")
    if pid and st.get("state") == "RUNNING":
        import signal
        try:
            os.kill(int(pid), signal.SIGTERM)
        except ProcessLookupError:
            pass
