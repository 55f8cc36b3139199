"""Platform table: which (cpu, memory, accelerator, count) machine shapes are valid, and how a
machine shape maps onto the local box.

The reference validates against Google AI Platform's compatibility table
(src/python/tensorflow_cloud/core/gcp.py:35-70, 123-406: a literal list of 199 tuples) and asks
google-auth for the project.  Here the same table is *generated* from its rule structure (which
n1 machine shapes exist, and the largest shape each accelerator count may be attached to), B200
rows are added, and "project" / "region" are local notions (user + host).  Label rules are the
ones of gcp.py:409-481.
"""
from __future__ import annotations

import getpass
import os
import re
import socket

# n1 machine shapes: (cpu_cores, memory_gb) -> AI Platform machine type (gcp.py:93-116)
_MACHINE_TYPES = {
    (4, 15): "n1-standard-4", (8, 30): "n1-standard-8", (16, 60): "n1-standard-16",
    (32, 120): "n1-standard-32", (64, 240): "n1-standard-64", (96, 360): "n1-standard-96",
    (2, 13): "n1-highmem-2", (4, 26): "n1-highmem-4", (8, 52): "n1-highmem-8",
    (16, 104): "n1-highmem-16", (32, 208): "n1-highmem-32", (64, 416): "n1-highmem-64",
    (96, 624): "n1-highmem-96", (16, 14.4): "n1-highcpu-16", (32, 28.8): "n1-highcpu-32",
    (64, 57.6): "n1-highcpu-64", (96, 86.4): "n1-highcpu-96",
}

# accelerator -> {count: (largest core count allowed, extra shapes beyond that)}
_HIGHCPU64 = [(64, 57.6)]
_GPU_RULES = {
    "K80": {1: (8, []), 2: (16, []), 4: (32, []), 8: (32, _HIGHCPU64)},
    "P100": {1: (16, []), 2: (32, []), 4: (32, _HIGHCPU64)},
    "P4": {1: (16, []), 2: (32, []), 4: (96, [])},
    "T4": {1: (16, []), 2: (32, []), 4: (96, [])},
    "V100": {1: (8, []), 2: (16, []), 4: (32, []), 8: (96, [])},
    # the local hardware: any host shape may carry 1, 2, 4 or 8 B200s
    "B200": {1: (96, []), 2: (96, []), 4: (96, []), 8: (96, [])},
}

_ACCELERATOR_NAMES = {
    "CPU": "ACCELERATOR_TYPE_UNSPECIFIED", "K80": "NVIDIA_TESLA_K80", "P100": "NVIDIA_TESLA_P100",
    "V100": "NVIDIA_TESLA_V100", "P4": "NVIDIA_TESLA_P4", "T4": "NVIDIA_TESLA_T4",
    "TPU_V2": "TPU_V2", "TPU_V3": "TPU_V3", "B200": "NVIDIA_B200",
}


def _get_valid_machine_configurations():
    out = [(c, m, "CPU", 0) for (c, m) in _MACHINE_TYPES]
    for accel, rules in _GPU_RULES.items():
        for count, (max_cores, extra) in rules.items():
            shapes = [s for s in _MACHINE_TYPES if s[0] <= max_cores] + list(extra)
            out.extend((c, m, accel, count) for (c, m) in shapes)
    out.append((None, None, "TPU_V2", 8))
    out.append((None, None, "TPU_V3", 8))
    return out


def get_project_name() -> str:
    """Local stand-in for the GCP project: ``$TFC_PROJECT`` or ``<user>@<host>``."""
    env = os.environ.get("TFC_PROJECT")
    if env:
        return env
    try:
        return f"{getpass.getuser()}-{socket.gethostname()}"
    except Exception:  # pragma: no cover
        raise RuntimeError("Could not determine the project id.")


def validate_machine_configuration(cpu_cores, memory, accelerator_type, accelerator_count):
    if accelerator_type.value in ("TPU_V2", "TPU_V3"):
        cpu_cores = memory = None
    current = (cpu_cores, memory, accelerator_type.value, accelerator_count)
    if current not in _get_valid_machine_configurations():
        raise ValueError(
            "Invalid machine configuration: cpu_cores:{}, memory:{}, accelerator_type:{}, "
            "accelerator_count:{}. Valid shapes are the n1 (cpu, memory) pairs with 0 accelerators, "
            "K80/P100/P4/T4/V100 counts from the AI Platform compatibility table, 1/2/4/8 B200, "
            "or a TPU with accelerator count 8.".format(
                cpu_cores, memory, str(accelerator_type), accelerator_count))


def get_region() -> str:
    """The reference always answers "us-central1" (gcp.py:73-75); locally the region is a label."""
    return os.environ.get("TFC_REGION", "us-central1")


def get_accelerator_type(accl_type: str) -> str:
    return _ACCELERATOR_NAMES[accl_type]


def get_machine_type(cpu_cores, memory, accelerator_type) -> str:
    if accelerator_type.value in ("TPU_V2", "TPU_V3"):
        return "cloud_tpu"
    return _MACHINE_TYPES[(cpu_cores, memory)]


def get_cloud_tpu_supported_tf_versions():
    return ["2.1"]


def validate_job_labels(job_labels) -> None:
    if not job_labels:
        print("No labels provided for the training job. Please consider creating labels to help "
              "with retrieval of job information (they are stored in the job registry).")
    if len(job_labels) > 64:
        raise ValueError("Invalid job labels: too many labels.Expecting at most 64 labels."
                         "Received {}.".format(len(job_labels)))
    for k, v in job_labels.items():
        if not k or not k[0].islower():
            raise ValueError("Invalid job labels:Label key must start with lowercase letters."
                             "Received {}.".format(k))
        if not v or not v[0].islower():
            raise ValueError("Invalid job labels:Label value must start with lowercase letters."
                             "Received {}.".format(v))
        if len(k) > 63:
            raise ValueError("Invalid job labels:Label key is too long.Expecting at most 63 "
                             "characters.Received {}.".format(k))
        if len(v) > 63:
            raise ValueError("Invalid job labels:Label value is too long for key {}.Expecting at "
                             "most 63 characters.Received {}.".format(k, v))
        if not re.match(r"^[a-z0-9_-]+$", k):
            raise ValueError("Invalid job labels:Label key can only contain lowercase letters,"
                             "numeric characters, underscores and dashes.Received: {}.".format(k))
        if not re.match(r"^[a-z0-9_-]+$", v):
            raise ValueError("Invalid job labels:Label value can only contain lowercase letters,"
                             "numeric characters, underscores and dashes.Received: {}.".format(v))
