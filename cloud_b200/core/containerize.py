"""Staging - the local replacement for Docker containerisation
(reference: src/python/tensorflow_cloud/core/containerize.py:44-507).

The reference synthesises a Dockerfile, tars the entry point's directory tree together with the
wrapper script and ``requirements.txt`` (symlinks dereferenced), builds the image with the Docker
daemon or Cloud Build and pushes it to ``gcr.io``.  There is no Docker daemon, registry or network
on the B200 box, so the same *inputs* produce a **staged job image**: a directory
``<job_root>/images/tf_cloud_train_<uuid>/`` that holds

* ``app/``            - the entry point's directory tree + wrapper + requirements.txt, laid out by
                        the same file-path map the reference feeds to ``tar`` (containerize.py:235-277),
* ``Jobfile``         - the Dockerfile analogue: base "image" (framework + torch + CUDA versions
                        instead of ``tensorflow/tensorflow:<ver>-gpu``), WORKDIR, the requirements
                        check that replaces ``pip install -r``, and the ENTRYPOINT command,
* ``image.json``      - machine-readable manifest of the above.

Two builders mirror the reference's pair: ``LocalContainerBuilder`` copies straight into the image
directory (the "docker build" path); ``CloudContainerBuilder`` goes through a tarball in a named
staging bucket directory first and polls the unpack like Cloud Build is polled (the path notebooks
use, since they must name ``docker_image_bucket_name``).
"""
from __future__ import annotations

import importlib.metadata
import json
import logging
import os
import re
import shutil
import sys
import tarfile
import tempfile
import time
import uuid

from . import machine_config
from ..utils import framework_utils

logger = logging.getLogger(__name__)
logging.basicConfig(level=logging.INFO)


def job_root() -> str:
    root = os.environ.get("TFC_JOB_ROOT") or os.path.join(tempfile.gettempdir(), "cloud_b200_jobs")
    os.makedirs(root, exist_ok=True)
    return root


class ContainerBuilder(object):
    """Builds a staged job image from the entry point, wrapper and requirements."""

    def __init__(self, entry_point, preprocessed_entry_point, chief_config, worker_config,
                 docker_registry, project_id, requirements_txt=None, destination_dir="/app/",
                 docker_base_image=None, docker_image_bucket_name=None, called_from_notebook=False):
        self.entry_point = entry_point
        self.preprocessed_entry_point = preprocessed_entry_point
        self.chief_config = chief_config
        self.worker_config = worker_config
        self.docker_registry = docker_registry
        self.requirements_txt = requirements_txt
        self.destination_dir = destination_dir
        self.docker_base_image = docker_base_image
        self.docker_image_bucket_name = docker_image_bucket_name
        self.called_from_notebook = called_from_notebook
        self.project_id = project_id
        self.tar_file_path = None
        self.docker_file_path = None

    # -- public -------------------------------------------------------------------------------
    def get_docker_image(self, max_status_check_attempts=None, delay_between_status_checks=None):
        raise NotImplementedError

    def get_generated_files(self):
        return [self.docker_file_path, self.tar_file_path]

    # -- Jobfile (Dockerfile analogue) -----------------------------------------------------------
    def _base_image(self):
        if self.docker_base_image is not None:
            return self.docker_base_image
        import torch
        gpu = self.chief_config.accelerator_type != machine_config.AcceleratorType.NO_ACCELERATOR
        return "cloud_b200:{}-torch{}{}".format(framework_utils.get_version(),
                                                torch.__version__.split("+")[0],
                                                "-gpu" if gpu else "")

    def _base_image_exist(self):
        """The reference probes Docker Hub (containerize.py:287-301); locally the "image" exists
        iff the running interpreter can import the framework."""
        try:
            import cloud_b200  # noqa: F401
            return True
        except ImportError:
            return False

    def _create_docker_file(self):
        lines = ["FROM {}\n".format(self._base_image()),
                 "WORKDIR {}\n".format(self.destination_dir)]
        if self.requirements_txt is not None:
            _, req_name = os.path.split(self.requirements_txt)
            dst = os.path.join(self.destination_dir, req_name)
            lines.append("COPY {} {}\n".format(dst, dst))
            lines.append("RUN check-requirements {}\n".format(dst))   # offline: verify, not install
        if machine_config.is_tpu_config(self.worker_config):
            lines.append("RUN echo 'TPU workers are not available on this box'\n")
        lines.append("COPY {} {}\n".format(self.destination_dir, self.destination_dir))
        if self.preprocessed_entry_point is not None:
            _, file_name = os.path.split(self.preprocessed_entry_point)
        else:
            _, file_name = os.path.split(self.entry_point or sys.argv[0])
        lines.append('ENTRYPOINT ["python", "{}"]\n'.format(file_name))
        _, self.docker_file_path = tempfile.mkstemp()
        with open(self.docker_file_path, "w") as f:
            f.writelines(lines)
        self._entry_file = file_name
        return lines

    # -- file map / tarball ------------------------------------------------------------------------
    def _get_file_path_map(self):
        """source path -> path inside the image (same rules as containerize.py:235-277)."""
        location_map = {}
        if self.entry_point is None and sys.argv[0].endswith("py"):
            self.entry_point = sys.argv[0]
        if not self.called_from_notebook:
            entry_point_dir, _ = os.path.split(self.entry_point)
            if not entry_point_dir:
                entry_point_dir = "."
            location_map[entry_point_dir] = self.destination_dir
        if self.preprocessed_entry_point is not None:
            _, name = os.path.split(self.preprocessed_entry_point)
            location_map[self.preprocessed_entry_point] = os.path.join(self.destination_dir, name)
        if self.requirements_txt is not None:
            _, name = os.path.split(self.requirements_txt)
            location_map[self.requirements_txt] = os.path.join(self.destination_dir, name)
        location_map[self.docker_file_path] = "Jobfile"
        return location_map

    def _get_tar_file_path(self):
        self._create_docker_file()
        file_path_map = self._get_file_path_map()
        _, self.tar_file_path = tempfile.mkstemp()
        with tarfile.open(self.tar_file_path, "w:gz", dereference=True) as tar:
            for source, destination in file_path_map.items():
                tar.add(source, arcname=destination.lstrip("/"), filter=_tar_filter)

    def _generate_name(self):
        unique_tag = str(uuid.uuid4()).replace("-", "_")
        return "{}/{}:{}".format(self.docker_registry, "tf_cloud_train", unique_tag)

    # -- requirements (replaces `pip install -r`) ----------------------------------------------------
    def _check_requirements(self, path):
        missing = []
        with open(path) as f:
            for raw in f:
                line = raw.split("#", 1)[0].strip()
                if not line or line.startswith("-"):
                    continue
                name = re.split(r"[<>=!~;\[ ]", line, 1)[0].strip()
                if not name:
                    continue
                try:
                    importlib.metadata.version(name)
                except importlib.metadata.PackageNotFoundError:
                    missing.append(name)
        if missing:
            raise RuntimeError("Staging failed: requirements not installed on this box and there "
                               "is no network to install them: {}".format(", ".join(missing)))

    def _image_dir(self, image_uri):
        tag = image_uri.rsplit(":", 1)[-1]
        return os.path.join(job_root(), "images", "tf_cloud_train_" + tag)

    def _finalise_image(self, image_dir, image_uri):
        if self.requirements_txt is not None:
            self._check_requirements(self.requirements_txt)
        manifest = {
            "image_uri": image_uri, "base_image": self._base_image(),
            "workdir": os.path.join(image_dir, self.destination_dir.strip("/")),
            "entrypoint": ["python", self._entry_file],
            "project_id": self.project_id, "created": time.time(),
            "python": sys.executable,
        }
        with open(os.path.join(image_dir, "image.json"), "w") as f:
            json.dump(manifest, f, indent=1)
        return manifest


def _tar_filter(info):
    name = info.name
    if "/__pycache__" in name or name.endswith(".pyc") or "/.git/" in name or name.endswith("/.git"):
        return None
    return info


class LocalContainerBuilder(ContainerBuilder):
    """"docker build + push" on the local box: unpack the context into the image store."""

    def get_docker_image(self, max_status_check_attempts=None, delay_between_status_checks=None):
        self._get_tar_file_path()
        image_uri = self._generate_name()
        logger.info("Building staged job image.")
        image_dir = self._build_docker_image(image_uri)
        logger.info("Publishing staged job image: {}".format(image_uri))
        self._publish_docker_image(image_dir, image_uri)
        return image_uri

    def _build_docker_image(self, image_uri):
        image_dir = self._image_dir(image_uri)
        os.makedirs(image_dir, exist_ok=True)
        try:
            with tarfile.open(self.tar_file_path, "r:gz") as tar:
                tar.extractall(image_dir, filter="data")
        except (tarfile.TarError, OSError) as e:
            raise RuntimeError("Staging error: {}".format(e))
        return image_dir

    def _publish_docker_image(self, image_dir, image_uri):
        self._finalise_image(image_dir, image_uri)


class CloudContainerBuilder(ContainerBuilder):
    """Bucket path: upload the tarball to ``<job_root>/buckets/<bucket>/``, then "build" (unpack)
    it, polling for completion the way Cloud Build is polled (containerize.py:432-458)."""

    def get_docker_image(self, max_status_check_attempts=20, delay_between_status_checks=30):
        self._get_tar_file_path()
        storage_object_name = self._upload_tar_to_gcs()
        image_uri = self._generate_name()
        logger.info("Building and publishing staged job image: {}".format(image_uri))
        build = self._create_cloud_build_request_dict(image_uri, storage_object_name)
        image_dir = self._image_dir(image_uri)
        os.makedirs(image_dir, exist_ok=True)
        status = "QUEUED"
        attempts = 0
        while attempts < (max_status_check_attempts or 20):
            attempts += 1
            if status == "QUEUED":
                status = "WORKING"
                try:
                    with tarfile.open(build["source"]["storageSource"]["path"], "r:gz") as tar:
                        tar.extractall(image_dir, filter="data")
                    status = "SUCCESS"
                except (tarfile.TarError, OSError) as e:
                    logger.error("staging build failed: %s", e)
                    status = "FAILURE"
            if status not in ("WORKING", "QUEUED"):
                break
            time.sleep(min(delay_between_status_checks or 0, 0.1))
        if status != "SUCCESS":
            raise RuntimeError("There was an error executing the staging build. Job status: " + status)
        self._finalise_image(image_dir, image_uri)
        return image_uri

    def _upload_tar_to_gcs(self):
        bucket_dir = os.path.join(job_root(), "buckets", self.docker_image_bucket_name)
        os.makedirs(bucket_dir, exist_ok=True)          # "create bucket if missing"
        unique_tag = str(uuid.uuid4()).replace("-", "_")
        name = "tf_cloud_train_tar_{}".format(unique_tag)
        shutil.copyfile(self.tar_file_path, os.path.join(bucket_dir, name))
        return name

    def _create_cloud_build_request_dict(self, image_uri, storage_object_name):
        path = os.path.join(job_root(), "buckets", self.docker_image_bucket_name, storage_object_name)
        return {
            "projectId": self.project_id,
            "images": [[image_uri]],
            "steps": {"name": "cloud_b200/stage", "args": ["unpack", "-t", image_uri, "."]},
            "source": {"storageSource": {"bucket": self.docker_image_bucket_name,
                                         "object": storage_object_name, "path": path}},
        }


def resolve_image(image_uri):
    """image uri -> manifest dict of a previously staged image."""
    tag = image_uri.rsplit(":", 1)[-1]
    path = os.path.join(job_root(), "images", "tf_cloud_train_" + tag, "image.json")
    if not os.path.exists(path):
        raise RuntimeError("Unknown job image: {}".format(image_uri))
    with open(path) as f:
        return json.load(f)
